"""Frame-level data parallelism for batches (BASELINE.json config 5): frame f -> GPU f mod N.

The reference handles one image per process launch with pageable, synchronous copies
(``kernel.cu:163,202``). Here every GPU runs its own three-stream H2D / filter / D2H software
pipeline (:meth:`Pipeline.stream`) fed from pinned host frames; one Python thread drives all of
them because every call is asynchronous.
"""
from __future__ import annotations

from typing import List, Sequence

import torch

from ..models.pipeline import Pipeline
from ..ops.chain import image_geometry


class BatchStreamer:
    def __init__(self, chain, gpus: int = 1, depth: int = 3, **kw):
        n = min(gpus, torch.cuda.device_count())
        self.pipes: List[Pipeline] = [Pipeline(chain, device=f"cuda:{g}", **kw) for g in range(n)]
        self.depth = depth

    def run(self, frames_in: Sequence[torch.Tensor], frames_out: Sequence[torch.Tensor]) -> int:
        """frames_in / frames_out: pinned uint8 HxWxC CPU tensors. Returns the number of frames."""
        n = len(self.pipes)
        # interleave the per-GPU pipelines frame by frame so that all GPUs are fed from the start
        gens = []
        for g, p in enumerate(self.pipes):
            gens.append(_stream_steps(p, frames_in[g::n], frames_out[g::n], self.depth))
        alive = list(gens)
        count = 0
        while alive:
            for it in list(alive):
                try:
                    next(it)
                    count += 1
                except StopIteration:
                    alive.remove(it)
        for p in self.pipes:
            if p._slots is not None:
                p._slots["d2h"].synchronize()
        return count


def _stream_steps(pipe: Pipeline, fins, fouts, depth):
    """Generator form of Pipeline.stream: enqueues one frame per next()."""
    s = None
    k_total = 0
    for fin, fout in zip(fins, fouts):
        H, W, c = image_geometry(fin)
        if s is None:
            s = pipe._make_slots(H, W, c, depth)
            plan = pipe.plan(c)
        k = k_total % depth
        with torch.cuda.device(pipe.device):
            with torch.cuda.stream(s["h2d"]):
                if s["used"][k]:
                    s["h2d"].wait_event(s["ev_run"][k])
                s["d_in"][k].copy_(fin.view(H, W, c), non_blocking=True)
                s["ev_in"][k].record(s["h2d"])
            with torch.cuda.stream(s["run"]):
                s["run"].wait_event(s["ev_in"][k])
                if s["used"][k]:
                    s["run"].wait_event(s["ev_out"][k])
                plan.run(s["d_in"][k], s["d_out"][k])
                s["ev_run"][k].record(s["run"])
            with torch.cuda.stream(s["d2h"]):
                s["d2h"].wait_event(s["ev_run"][k])
                fout.view(H, W, s["oc"]).copy_(s["d_out"][k], non_blocking=True)
                s["ev_out"][k].record(s["d2h"])
        s["used"][k] = True
        k_total += 1
        yield k_total
