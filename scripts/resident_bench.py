"""Resident bands under torchrun: P passes of a chain with halo rows pushed between the GPUs, one
cooperative launch (fused) against one launch per pass; result checked around every band boundary.
    torchrun --nproc-per-node N scripts/resident_bench.py [--size WxH] [--chain gauss5] [--passes 8]"""
import argparse, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
import numpy as np
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.chain import band_rows
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime, _boundary_rows, _check_rows, _timed

ap = argparse.ArgumentParser()
ap.add_argument("--size", default="7680x4320")
ap.add_argument("--chain", default="gauss5")
ap.add_argument("--passes", type=int, default=8)
ap.add_argument("--check", type=int, default=1)
ap.add_argument("--cycle", type=int, default=0, help="also time halo exchanges every CYCLE passes (2 or 4)")
a = ap.parse_args()
W, H = (int(x) for x in a.size.split("x"))
rt = DistRuntime(root=0)
rank, world = rt.rank, rt.world
plan = Plan(a.chain, 3)
img = b2.synth_image(W, H, 3, seed=1) if rank == 0 else None
rt.setup_root(W, H, 3, 3)
rt.load_root(img)
rt.setup_resident(W, H, 3)
P = a.passes
out = {}
def run(mode):
    if mode == "cycle":
        rt.run_resident_cycle(plan, P, a.cycle)
        return False
    return rt.run_resident(plan, P, fused=mode == "fused")


for mode in ("fused", "steps") + (("cycle",) if a.cycle else ()):
    fused = mode == "fused"
    rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
    ran = run(mode)
    rt.barrier()
    ok = None
    if a.check:
        rt.gather_resident(P, rt.dst.ptrs[0], rt.geom["dp"], 3)
        got = rt.fetch_root()
        if rank == 0:
            bands = [band_rows(H, world, r) for r in range(world)]
            ok = _check_rows(",".join([a.chain] * P), img, got, _boundary_rows(H, bands))
    rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
    for _ in range(2):
        run(mode)
    rt.barrier()
    rr = [_timed(lambda: run(mode)) / P * 1e3 for _ in range(7)]
    import time
    torch.cuda.synchronize(); rt.barrier()
    t0 = time.perf_counter(); run(mode); host_us = (time.perf_counter() - t0) / P * 1e6
    torch.cuda.synchronize(); rt.barrier()
    st = rt.status()
    out[mode] = (statistics.median(rr), min(rr), ran, ok, st, host_us)
if rank == 0:
    for mode, (med, mn, ran, ok, st, host_us) in out.items():
        print(f"N={world} {a.size} {a.chain} x{P} {mode:5s}: {med:7.1f} us/pass (min {mn:.1f}) fused_ran={ran} status={st} check={ok} host_enqueue={host_us:.1f} us/pass",
              flush=True)
rt.close()
