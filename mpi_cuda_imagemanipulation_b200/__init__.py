"""b200img — Blackwell-native distributed image filtering.

Same capabilities as Dohruba/MPI-CUDA-ImageManipulation (gray -> contrast -> emboss -> BGR over
row bands), rebuilt for 8xB200: hand-written sm_100a kernels that fuse scatter, halo exchange and
gather with the filter math over NVLink peer memory, a C++ runtime in place of MPI ranks, and
this thin tensor front end.
"""
from ._lib import B200ImgError, LIB_PATH, lib  # noqa: F401
from .ops.chain import Chain, as_chain, band_rows, contrast_lut, host_run  # noqa: F401
from .models.pipeline import Pipeline, kern_cpp, kernel_cu, kernel_cu_5x5  # noqa: F401
from .utils.io import read_image, read_raw, synth_image, write_image, write_raw  # noqa: F401

__version__ = lib.b2i_version().decode()
