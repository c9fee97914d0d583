"""Filter chains: the unit of work of the framework.

A :class:`Chain` is the data form of what the reference hard-codes as three kernel launches plus a
host ``cvtColor`` (``kernel.cu:192-195,210``) or four OpenCV calls (``kern.cpp:73-77``). The same
object drives the host reference, the single-GPU kernels and the multi-GPU runtime.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Union

import numpy as np

from .. import _lib
from .._lib import ChainStruct, check, lib

try:  # torch is the tensor front end, but the host path must not depend on it
    import torch
except Exception:  # pragma: no cover
    torch = None

ArrayLike = Union[np.ndarray, "torch.Tensor"]


class Chain:
    """A parsed, validated filter chain (wraps ``b2i_chain_t``)."""

    def __init__(self, spec: Optional[str] = None, *, preset: Optional[str] = None,
                 border: Optional[str] = None, order: str = "bgr"):
        if (spec is None) == (preset is None):
            raise ValueError("give exactly one of spec= or preset=")
        self.c = ChainStruct()
        self.order = order.lower()
        o = _lib.ORDERS[self.order]
        if preset is not None:
            check(lib.b2i_chain_preset(preset.encode(), o, C.byref(self.c)))
        else:
            check(lib.b2i_chain_parse(spec.encode(), _lib.BORDERS[border], o, C.byref(self.c)))
        self.spec = spec if spec is not None else preset

    # -- introspection -------------------------------------------------------------------
    @property
    def n_ops(self) -> int:
        return self.c.n_ops

    @property
    def halo(self) -> int:
        """Rows a band needs from each neighbour (sum of stencil radii)."""
        return lib.b2i_chain_halo(C.byref(self.c))

    def out_channels(self, in_channels: int) -> int:
        return check(lib.b2i_chain_out_channels(C.byref(self.c), in_channels))

    def describe(self) -> str:
        buf = C.create_string_buffer(512)
        check(lib.b2i_chain_describe(C.byref(self.c), buf, 512))
        return buf.value.decode()

    def __repr__(self) -> str:
        return f"Chain({self.describe()!r}, order={self.order!r})"

    @property
    def ref(self):
        return C.byref(self.c)


def as_chain(x, **kw) -> Chain:
    if isinstance(x, Chain):
        return x
    if isinstance(x, str):
        if x in ("kernel_cu", "kern_cpp", "kernel_cu_5x5", "kernel.cu", "kern.cpp"):
            return Chain(preset=x, order=kw.get("order", "bgr"))
        return Chain(x, **kw)
    raise TypeError(f"cannot make a filter chain from {type(x).__name__}")


def image_geometry(img: ArrayLike):
    """(H, W, C) of an image given as HxW or HxWxC uint8."""
    if img.ndim == 2:
        return int(img.shape[0]), int(img.shape[1]), 1
    if img.ndim == 3 and img.shape[2] in (1, 3):
        return int(img.shape[0]), int(img.shape[1]), int(img.shape[2])
    raise ValueError(f"expected HxW or HxWx{{1,3}} uint8 image, got shape {tuple(img.shape)}")


def _host_view(img: ArrayLike) -> np.ndarray:
    if torch is not None and isinstance(img, torch.Tensor):
        if img.dtype != torch.uint8:
            raise TypeError("images are uint8")
        return img.detach().cpu().contiguous().numpy()
    a = np.asarray(img)
    if a.dtype != np.uint8:
        raise TypeError("images are uint8")
    return np.ascontiguousarray(a)


def _like(out: np.ndarray, proto: ArrayLike):
    if torch is not None and isinstance(proto, torch.Tensor):
        return torch.from_numpy(out)
    return out


def host_run(chain, img: ArrayLike, *, rows=None, ranks: int = 0, seams: bool = False,
             **kw) -> ArrayLike:
    """Runs ``chain`` on the CPU with the C++ host reference (``csrc/host_ref``).

    ``rows=(y0, y1)`` returns only those rows of the full-image result. ``ranks=N`` uses the
    host "ranks" backend: N workers, root-owned scatter/gather like ``kernel.cu:137,223``;
    ``seams=True`` filters each band in isolation the way the reference does.
    """
    ch = as_chain(chain, **kw)
    src = _host_view(img)
    H, W, Cn = image_geometry(src)
    oc = ch.out_channels(Cn)
    if rows is None:
        out = np.empty((H, W, oc), np.uint8)
        if ranks and ranks > 0:
            ms = C.c_double(0)
            check(lib.b2i_host_run_ranks(ch.ref, src.ctypes.data, W, H, Cn, W * Cn,
                                         out.ctypes.data, W * oc, ranks, int(seams), C.byref(ms)))
        else:
            check(lib.b2i_host_run(ch.ref, src.ctypes.data, W, H, Cn, W * Cn, out.ctypes.data,
                                   W * oc))
    else:
        y0, y1 = rows
        out = np.empty((y1 - y0, W, oc), np.uint8)
        check(lib.b2i_host_run_rows(ch.ref, src.ctypes.data, W, H, Cn, W * Cn, y0, y1,
                                    out.ctypes.data, W * oc))
    if oc == 1 and img.ndim == 2:
        out = out[:, :, 0]
    return _like(out, img)


def band_rows(H: int, n: int, i: int):
    """Rows [y0, y1) of band i of n; remainder rows are spread, never dropped (cf. kernel.cu:117)."""
    y0, y1 = C.c_int(0), C.c_int(0)
    check(lib.b2i_band_rows(H, n, i, C.byref(y0), C.byref(y1)))
    return y0.value, y1.value


def contrast_lut(factor: float, flavour: str = "ref") -> np.ndarray:
    lut = np.empty(256, np.uint8)
    if flavour == "ref":
        lib.b2i_lut_contrast_ref(float(factor), lut.ctypes.data)
    else:
        lib.b2i_lut_contrast_cv(float(factor), lut.ctypes.data)
    return lut
