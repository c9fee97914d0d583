from .runtime import Runtime, topology  # noqa: F401
