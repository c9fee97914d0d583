from .io import read_image, read_raw, synth_image, write_image, write_raw  # noqa: F401
