from .chain import Chain, as_chain, band_rows, contrast_lut, host_run, image_geometry  # noqa: F401
