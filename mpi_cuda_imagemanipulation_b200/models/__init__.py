from .pipeline import Pipeline, kern_cpp, kernel_cu, kernel_cu_5x5  # noqa: F401
