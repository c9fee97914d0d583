"""JPEG / PPM / PGM / raw file I/O and synthetic images (``csrc/core/io.cpp``, ``csrc/core/jpeg.cpp``).

The reference's file I/O is ``cv::imread`` / ``cv::imwrite`` of JPEG on rank 0 (``kernel.cu:110,236``):
JPEG files go through nvJPEG (needs a GPU); PPM / PGM / raw keep the same interleaved bytes lossless.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _lib
from .._lib import check, lib


def read_image(path: str, order: str = "bgr") -> np.ndarray:
    """Reads a JPEG (by content) or a binary PPM/PGM into an HxWxC uint8 array (3-channel data in ``order``)."""
    data = C.c_void_p()
    W, H, Cn = C.c_int(), C.c_int(), C.c_int()
    check(lib.b2i_read_image(path.encode(), _lib.ORDERS[order], C.byref(data), C.byref(W),
                           C.byref(H), C.byref(Cn)))
    try:
        n = W.value * H.value * Cn.value
        arr = np.ctypeslib.as_array(C.cast(data, C.POINTER(C.c_uint8)), shape=(n,)).copy()
    finally:
        lib.b2i_free(data)
    return arr.reshape(H.value, W.value, Cn.value)


def write_image(path: str, img, order: str = "bgr", quality: int = 95) -> None:
    """``.jpg`` / ``.jpeg``: JPEG at ``quality`` (lossy; nvJPEG, needs a GPU); anything else: binary PPM / PGM."""
    a = np.ascontiguousarray(np.asarray(img), dtype=np.uint8)
    if a.ndim == 2:
        a = a[:, :, None]
    H, W, Cn = a.shape
    check(lib.b2i_write_image(path.encode(), _lib.ORDERS[order], a.ctypes.data, W, H, Cn, W * Cn, quality))


def read_raw(path: str, W: int, H: int, Cn: int = 3) -> np.ndarray:
    data = C.c_void_p()
    check(lib.b2i_read_raw(path.encode(), W, H, Cn, C.byref(data)))
    try:
        arr = np.ctypeslib.as_array(C.cast(data, C.POINTER(C.c_uint8)), shape=(W * H * Cn,)).copy()
    finally:
        lib.b2i_free(data)
    return arr.reshape(H, W, Cn)


def write_raw(path: str, img) -> None:
    a = np.ascontiguousarray(np.asarray(img), dtype=np.uint8)
    if a.ndim == 2:
        a = a[:, :, None]
    H, W, Cn = a.shape
    check(lib.b2i_write_raw(path.encode(), a.ctypes.data, W, H, Cn, W * Cn))


def synth_image(W: int, H: int, Cn: int = 3, seed: int = 0, out=None) -> np.ndarray:
    """Deterministic synthetic test card (there is no network for real photographs)."""
    if out is None:
        out = np.empty((H, W, Cn), np.uint8)
    check(lib.b2i_synth_image(out.ctypes.data, W, H, Cn, W * Cn, seed))
    return out
