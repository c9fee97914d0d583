"""Single-GPU CUDA execution of a filter chain on torch tensors (zero-copy, current stream).

The tensor is handed to the sm_100a kernels by address (``data_ptr``); nothing is copied or
re-laid-out. This is the replacement for the reference's ``cudaMalloc`` / synchronous
``cudaMemcpy`` / three launches / ``cudaDeviceSynchronize`` / D2H / host ``cvtColor`` sequence
(``kernel.cu:147-210``): one (or, for long chains, a few) launches on the caller's stream.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _lib
from .._lib import Launch, PlanInfo, View, check, lib
from .chain import Chain, as_chain, image_geometry


class Plan:
    """A chain compiled for ``in_channels`` input (wraps ``b2i_plan_t``)."""

    def __init__(self, chain, in_channels: int, **kw):
        self.chain: Chain = as_chain(chain, **kw)
        self.cin = in_channels
        self._p = C.c_void_p()
        check(lib.b2i_plan_create(self.chain.ref, in_channels, C.byref(self._p)))
        self.cout = self.chain.out_channels(in_channels)
        self._tmp = {}

    def __del__(self):
        p, self._p = getattr(self, "_p", None), None
        if p and lib is not None:  # module globals are gone at interpreter shutdown
            lib.b2i_plan_destroy(p)

    @property
    def handle(self):
        return self._p

    @property
    def halo(self) -> int:
        return lib.b2i_plan_halo(self._p)

    def info(self) -> dict:
        pi = PlanInfo()
        check(lib.b2i_plan_info(self._p, C.byref(pi)))
        n = pi.n_segments
        return {
            "cin": pi.cin, "cout": pi.cout, "halo": pi.halo, "n_segments": n,
            "segments": [{"pointwise": bool(pi.seg_pointwise[i]), "halo": pi.seg_halo[i],
                          "phases": pi.seg_phases[i], "cout": pi.seg_cout[i]} for i in range(n)],
            "tile_h": pi.last_tile_h, "smem_bytes": pi.last_smem_bytes, "grid": pi.last_grid,
            "blocks_per_sm": pi.last_blocks_per_sm, "tma_load": bool(pi.last_tma_load),
            "tma_store": bool(pi.last_tma_store), "launches": pi.last_launches,
        }

    @property
    def n_segments(self) -> int:
        return self.info()["n_segments"]

    # scratch images for multi-segment plans / the reference-style path, cached per geometry
    def scratch(self, device, W: int, H: int):
        key = (str(device), W, H)
        if key not in self._tmp:
            pitch = (W * 3 + 127) // 128 * 128     # pitched: TMA needs a 16-byte row pitch at any width
            self._tmp[key] = (torch.empty(H * pitch, dtype=torch.uint8, device=device),
                              torch.empty(H * pitch, dtype=torch.uint8, device=device),
                              torch.zeros(256 * _lib.B2I_MAX_OPS, dtype=torch.uint8, device=device))
        return self._tmp[key]

    def run(self, src: torch.Tensor, out: Optional[torch.Tensor] = None, *, naive: bool = False,
            no_tma: bool = False, tile_h: int = 0, stage: Optional[bool] = None) -> torch.Tensor:
        """dst = chain(src) for a whole HxWxC uint8 CUDA tensor on the current stream.

        ``stage``: tightly packed images whose row length is not a multiple of 16 bytes cannot be
        addressed by TMA. By default (None) such an image of 64 Ki pixels or more is copied into a
        pitched buffer of this plan, filtered there by the fast kernels and copied back (two extra
        passes over memory, still well ahead of the plain-loader kernel); False keeps it where it is.
        Images that already live in pitched buffers (``pitched_empty``) need neither."""
        if not src.is_cuda or src.dtype != torch.uint8:
            raise TypeError("expected a uint8 CUDA tensor")
        H, W, Cn = image_geometry(src)
        if Cn != self.cin:
            raise ValueError(f"plan was built for {self.cin} channels, image has {Cn}")
        if src.stride(-1) != 1 and src.ndim == 3 or (src.ndim == 3 and src.stride(1) != Cn):
            src = src.contiguous()
        src_pitch = src.stride(0)
        if out is None:
            shape = (H, W) if (self.cout == 1 and src.ndim == 2) else (H, W, self.cout)
            out = torch.empty(shape, dtype=torch.uint8, device=src.device)
        elif out.numel() != H * W * self.cout or out.dtype != torch.uint8:
            raise ValueError("out must be a uint8 tensor of H*W*cout bytes")
        elif not out.is_contiguous() and not (out.ndim == 3 and out.stride(2) == 1 and out.stride(1) == self.cout):
            raise ValueError("out must be contiguous, or rows of packed pixels at a row pitch (see pitched_empty)")
        dst_pitch = W * self.cout if out.is_contiguous() else out.stride(0)
        flags = (_lib.LAUNCH_NAIVE if naive else 0) | (_lib.LAUNCH_NO_TMA if no_tma else 0)
        if stage is None:
            stage = W * H >= 65536
        if stage and not naive and not no_tma and ((src_pitch | dst_pitch) & 15 or (src.data_ptr() | out.data_ptr()) & 15):
            key = ("stage", str(src.device), W, H)
            if key not in self._tmp:
                self._tmp[key] = (pitched_empty(H, W, Cn, src.device), pitched_empty(H, W, self.cout, src.device))
            sp, dp = self._tmp[key]
            stream = torch.cuda.current_stream().cuda_stream
            with torch.cuda.device(src.device):   # 2-D device copies (cudaMemcpy2DAsync), rows of W*C bytes
                check(lib.b2i_cuda_memcpy2d(sp.data_ptr(), sp.stride(0), src.data_ptr(), src_pitch, W * Cn, H, 2, stream))
                self.run(sp, dp, tile_h=tile_h, stage=False)
                check(lib.b2i_cuda_memcpy2d(out.data_ptr(), dst_pitch, dp.data_ptr(), dp.stride(0), W * self.cout, H, 2,
                                            stream))
            return out
        need_tmp = naive or self.n_segments > 1 or (self.halo > 0 and (W <= 16 or H <= 16))
        t0 = t1 = ls = None
        if need_tmp:
            t0, t1, ls = self.scratch(src.device, W, H)
        with torch.cuda.device(src.device):
            L = Launch()
            L.W, L.H, L.y0, L.y1 = W, H, 0, H
            L.src = View(src.data_ptr(), src_pitch, 0, H)
            L.dst = View(out.data_ptr(), dst_pitch, 0, H)
            if need_tmp:
                tp = (W * 3 + 127) // 128 * 128
                L.tmp[0] = View(t0.data_ptr(), tp, 0, H)
                L.tmp[1] = View(t1.data_ptr(), tp, 0, H)
                L.counters = ls.data_ptr()
            L.stream = torch.cuda.current_stream().cuda_stream
            L.tile_h = tile_h
            L.flags = flags
            check(lib.b2i_cuda_run(self._p, C.byref(L)))
        return out


def run_passes(plan: "Plan", img: torch.Tensor, passes: int, fused: bool = True) -> torch.Tensor:
    """`passes` applications of a one-stencil plan to a whole image on one GPU, ping-ponging between
    two buffers: with fused=True in ONE cooperative launch whose tasks order themselves with
    per-task flags (the single-GPU form of ``DistRuntime.run_resident``), else one launch per pass."""
    H, W, Cn = image_geometry(img)
    bufs = [img.contiguous().clone(), torch.empty_like(img)]
    words = torch.zeros(64, dtype=torch.int32, device=img.device)
    flags = torch.zeros(16384, dtype=torch.int32, device=img.device)
    ls = []
    with torch.cuda.device(img.device):
        for k in range(2):
            L = Launch()
            L.W, L.H, L.y0, L.y1 = W, H, 0, H
            L.src = View(bufs[k].data_ptr(), W * Cn, 0, H)
            L.dst = View(bufs[k ^ 1].data_ptr(), W * Cn, 0, H)
            L.counters = words.data_ptr()
            L.status = words.data_ptr() + 16
            L.stream = torch.cuda.current_stream().cuda_stream
            ls.append(L)
        if fused:
            check(lib.b2i_cuda_run_resident_fused(plan.handle, C.byref(ls[0]), C.byref(ls[1]), passes, 0,
                                                  flags.data_ptr(), flags.numel()))
        else:
            check(lib.b2i_cuda_run_resident_steps(plan.handle, C.byref(ls[0]), C.byref(ls[1]), passes, 0))
    torch.cuda.synchronize(img.device)
    if int(words[4].item()) != 0:
        raise RuntimeError(f"a task wait timed out (status {int(words[4].item())})")
    return bufs[passes & 1]


def pitched_empty(H: int, W: int, Cn: int = 3, device="cuda", align: int = 128) -> torch.Tensor:
    """An HxWxC uint8 image whose rows start `align` bytes apart-aligned: ``t.stride(0)`` is the pitch.

    TMA (and with it every fast kernel here) needs a 16-byte-aligned row pitch; a tightly packed
    image has one only when ``W * C`` is a multiple of 16. The reference takes any width
    (``kernel.cu:116``); images of odd widths should live in buffers like this one, the way
    ``cudaMallocPitch`` lays them out."""
    pitch = (W * Cn + align - 1) // align * align
    buf = torch.empty(H * pitch, dtype=torch.uint8, device=device)
    return buf.as_strided((H, W, Cn), (pitch, Cn, 1))


_PLAN_CACHE = {}


def cuda_run(chain, img: torch.Tensor, out=None, **kw) -> torch.Tensor:
    """One-shot convenience: plans are cached per (chain description, order, channels)."""
    run_kw = {k: kw.pop(k) for k in ("naive", "no_tma", "tile_h") if k in kw}
    ch = as_chain(chain, **kw)
    _, _, Cn = image_geometry(img)
    key = (ch.describe(), bytes(ch.c), Cn)
    plan = _PLAN_CACHE.get(key)
    if plan is None:
        plan = _PLAN_CACHE[key] = Plan(ch, Cn)
    return plan.run(img, out, **run_kw)
