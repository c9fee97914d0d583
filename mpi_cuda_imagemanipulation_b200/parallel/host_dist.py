"""CPU ranks over torch.distributed (gloo): BASELINE.json config 1, "mpirun -np 2 on CPU with host
reference filter". The reference's kern.cpp data flow (kern.cpp:46-55,81-83) — rank 0 owns the
image, bands are scattered, filtered on the host, gathered — with two differences: bands carry
halo rows (the result has no seams) and remainder rows are kept (kern.cpp:40 drops them).

    torchrun --nproc-per-node 2 -m mpi_cuda_imagemanipulation_b200.parallel.host_dist --chain gray
"""
from __future__ import annotations

import argparse
import json
import os
import time

import numpy as np
import torch
import torch.distributed as dist

from ..ops.chain import as_chain, band_rows, host_run


def run_host_ranks(img, chain, root: int = 0):
    """Collective: every rank calls it; `img` (HxWxC uint8 numpy) is read on the root only.
    Returns the filtered image on the root, None elsewhere."""
    rank, world = dist.get_rank(), dist.get_world_size()
    ch = as_chain(chain)
    meta = [None]
    if rank == root:
        H, W, C = img.shape
        meta[0] = (H, W, C)                       # MPI_Bcast of image_properties (kern.cpp:48)
    dist.broadcast_object_list(meta, src=root)
    H, W, C = meta[0]
    halo, oc = ch.halo, ch.out_channels(C)
    spans = []
    for r in range(world):
        y0, y1 = band_rows(H, world, r)
        spans.append((y0, y1, max(0, y0 - halo - 1), min(H, y1 + halo + 1)))
    y0, y1, ra, rb = spans[rank]
    mine = torch.empty((rb - ra, W, C), dtype=torch.uint8)
    if rank == root:                              # MPI_Scatter with overlap (kern.cpp:55)
        t = torch.from_numpy(np.ascontiguousarray(img))
        reqs = []
        for r, (_, _, a, b) in enumerate(spans):
            if r == root:
                mine.copy_(t[a:b])
            else:
                reqs.append(dist.isend(t[a:b].contiguous(), dst=r))
        for q in reqs:
            q.wait()
    else:
        dist.recv(mine, src=root)
    # rows [y0, y1) of the whole-image result from rows [ra, rb): host_run's rows= needs global
    # coordinates, so the band is embedded at its place in a virtual image via the C API's view
    from .. import _lib
    import ctypes as Ct
    out = np.empty((y1 - y0, W, oc), np.uint8)
    if y1 > y0:
        src = mine.numpy()
        # b2i_host_run_rows reads source rows relative to row 0: shift the pointer by ra rows
        base = src.ctypes.data - ra * W * C
        _lib.check(_lib.lib.b2i_host_run_rows(ch.ref, base, W, H, C, W * C, y0, y1, out.ctypes.data, W * oc))
    res = None
    if rank == root:                              # MPI_Gather (kern.cpp:81)
        res = np.empty((H, W, oc), np.uint8)
        for r, (a0, a1, _, _) in enumerate(spans):
            if r == root:
                res[a0:a1] = out
            elif a1 > a0:
                buf = torch.empty((a1 - a0, W, oc), dtype=torch.uint8)
                dist.recv(buf, src=r)
                res[a0:a1] = buf.numpy()
    elif y1 > y0:
        dist.send(torch.from_numpy(out), dst=root)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chain", default="gray")
    ap.add_argument("--width", type=int, default=512)
    ap.add_argument("--height", type=int, default=512)
    a = ap.parse_args()
    dist.init_process_group("gloo")
    import mpi_cuda_imagemanipulation_b200 as b2
    rank = dist.get_rank()
    img = b2.synth_image(a.width, a.height, 3, seed=1) if rank == 0 else None
    dist.barrier()
    t0 = time.perf_counter()
    out = run_host_ranks(img, a.chain)
    dt = time.perf_counter() - t0
    if rank == 0:
        ok = bool(np.array_equal(out, host_run(a.chain, img)))
        print(json.dumps({"config": "host ranks (gloo)", "ranks": dist.get_world_size(), "chain": a.chain,
                          "image": f"{a.width}x{a.height}", "ms": round(dt * 1e3, 3), "bit_exact": ok}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
