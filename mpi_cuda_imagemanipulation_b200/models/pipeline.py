"""Pipelines: the framework's "models".

The reference ships exactly two fixed pipelines — ``kernel.cu`` (MPI+CUDA) and ``kern.cpp``
(MPI+OpenCV) — that differ in numerics. Here a :class:`Pipeline` is any filter
chain bound to a device, with the two reference programs available as presets
(:func:`kernel_cu`, :func:`kern_cpp`). ``Pipeline.__call__`` takes what the user has:

* a CUDA uint8 tensor  -> filtered in place on the GPU, zero-copy, on the current stream;
* a NumPy array / CPU tensor -> uploaded through pinned memory, filtered, downloaded;
* and :meth:`Pipeline.stream` pushes a sequence of host frames through a 3-stage
  copy/compute/copy software pipeline on side streams (BASELINE.json config 5).
"""
from __future__ import annotations

from typing import Iterable, Iterator, Optional

import numpy as np

from ..ops.chain import Chain, as_chain, host_run, image_geometry

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


def _cuda_ok() -> bool:
    return torch is not None and torch.cuda.is_available()


class Pipeline:
    def __init__(self, chain, *, device: Optional[str] = None, border: Optional[str] = None,
                 order: str = "bgr"):
        kw = {"order": order}
        if border is not None and isinstance(chain, str):
            kw["border"] = border
        self.chain: Chain = as_chain(chain, **kw)
        self.device = device if device is not None else ("cuda:0" if _cuda_ok() else "cpu")
        self._plans = {}
        self._slots = None

    # ---- plans are per input channel count ----------------------------------------------------
    def plan(self, in_channels: int):
        from ..ops.cuda import Plan
        if in_channels not in self._plans:
            self._plans[in_channels] = Plan(self.chain, in_channels)
        return self._plans[in_channels]

    @property
    def halo(self) -> int:
        return self.chain.halo

    def __repr__(self):
        return f"Pipeline({self.chain.describe()!r}, device={self.device!r})"

    # ---- one image ------------------------------------------------------------------------------
    def __call__(self, img, out=None):
        if torch is not None and isinstance(img, torch.Tensor) and img.is_cuda:
            _, _, c = image_geometry(img)
            return self.plan(c).run(img, out)
        if self.device == "cpu" or not _cuda_ok():
            return host_run(self.chain, img)
        a = img.numpy() if (torch is not None and isinstance(img, torch.Tensor)) else np.asarray(img)
        H, W, c = image_geometry(a)
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory().to(self.device, non_blocking=True)
        res = self.plan(c).run(t).cpu()
        if torch is not None and isinstance(img, torch.Tensor):
            return res
        r = res.numpy()
        return r

    run = __call__

    # ---- streaming: pinned host frames in, pinned host frames out -------------------------------
    def _make_slots(self, H, W, c, depth):
        dev = torch.device(self.device)
        oc = self.chain.out_channels(c)
        key = (H, W, c, depth)
        if self._slots is not None and self._slots["key"] == key:
            return self._slots
        s = {"key": key, "oc": oc,
             "d_in": [torch.empty((H, W, c), dtype=torch.uint8, device=dev) for _ in range(depth)],
             "d_out": [torch.empty((H, W, oc), dtype=torch.uint8, device=dev) for _ in range(depth)],
             "h2d": torch.cuda.Stream(dev), "run": torch.cuda.Stream(dev), "d2h": torch.cuda.Stream(dev),
             "ev_in": [torch.cuda.Event() for _ in range(depth)],
             "ev_run": [torch.cuda.Event() for _ in range(depth)],
             "ev_out": [torch.cuda.Event() for _ in range(depth)],
             "used": [False] * depth}
        self._slots = s
        return s

    def stream(self, frames_in: Iterable["torch.Tensor"], frames_out: Iterable["torch.Tensor"],
               depth: int = 3) -> int:
        """Pushes pinned host frames through H2D -> filter -> D2H on three side streams, ``depth``
        frames in flight (the reference copies synchronously from pageable memory, one image per
        process launch: kernel.cu:163,202). Returns the number of frames; synchronises at the end.
        ``frames_out[i]`` receives the result of ``frames_in[i]``."""
        n = 0
        s = None
        for fin, fout in zip(frames_in, frames_out):
            H, W, c = image_geometry(fin)
            if s is None:
                s = self._make_slots(H, W, c, depth)
                plan = self.plan(c)
            k = n % depth
            with torch.cuda.stream(s["h2d"]):
                if s["used"][k]:
                    s["h2d"].wait_event(s["ev_run"][k])      # slot's previous frame was consumed
                s["d_in"][k].copy_(fin.view(H, W, c), non_blocking=True)
                s["ev_in"][k].record(s["h2d"])
            with torch.cuda.stream(s["run"]):
                s["run"].wait_event(s["ev_in"][k])
                if s["used"][k]:
                    s["run"].wait_event(s["ev_out"][k])      # slot's previous result was downloaded
                plan.run(s["d_in"][k], s["d_out"][k])
                s["ev_run"][k].record(s["run"])
            with torch.cuda.stream(s["d2h"]):
                s["d2h"].wait_event(s["ev_run"][k])
                fout.view(H, W, s["oc"]).copy_(s["d_out"][k], non_blocking=True)
                s["ev_out"][k].record(s["d2h"])
            s["used"][k] = True
            n += 1
        if s is not None:
            s["d2h"].synchronize()
        return n


def kernel_cu(**kw) -> Pipeline:
    """The MPI+CUDA program's pipeline (kernel.cu:192-195,210): gray -> contrast 3.5 -> emboss3 -> BGR."""
    return Pipeline(Chain(preset="kernel_cu", order=kw.pop("order", "bgr")), **kw)


def kern_cpp(**kw) -> Pipeline:
    """The MPI+OpenCV program's pipeline (kern.cpp:73-77), OpenCV numerics."""
    return Pipeline(Chain(preset="kern_cpp", order=kw.pop("order", "bgr")), **kw)


def kernel_cu_5x5(**kw) -> Pipeline:
    """kernel.cu with smallEmboss=false (kernel.cu:76-82): compiled in the reference, never launched."""
    return Pipeline(Chain(preset="kernel_cu_5x5", order=kw.pop("order", "bgr")), **kw)
