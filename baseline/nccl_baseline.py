#!/usr/bin/env python
"""The multi-GPU BASELINE (not the product): NCCL scatter -> per-rank filter -> NCCL send/recv halos
-> NCCL gather, the reference's MPI data flow (kernel.cu:137,223) moved onto torch.distributed.
BASELINE.json: "A path that only calls MPI/NCCL scatter, gather or sendrecv for these ops is the
baseline, not the product." Run under torchrun:

    torchrun --nproc-per-node N baseline/nccl_baseline.py --chain gauss5 --passes 4 [--kernel naive|fast]

--kernel naive uses the reference-style one-thread-per-pixel kernels (what a straight port of
kernel.cu would run); --kernel fast uses this repo's fused kernels but still moves every band and
halo through NCCL, which isolates what fusing the transfers into the kernels buys.
Prints one JSON line on rank 0 (CUDA events per rank, max over ranks).
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")

import numpy as np
import torch
import torch.distributed as dist

import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chain", default="kernel_cu")
    ap.add_argument("--passes", type=int, default=1)
    ap.add_argument("--kernel", default="naive", choices=["naive", "fast"])
    ap.add_argument("--width", type=int, default=7680)
    ap.add_argument("--height", type=int, default=4320)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    W, H = a.width, a.height
    assert H % world == 0, "the baseline scatters equal bands (like kernel.cu:117 it cannot take a remainder)"
    hb = H // world
    plan = Plan(a.chain, 3)
    # one row more than the chain's halo: the REF border rule leaves r+1 rows at a sub-image's edge unfiltered
    halo, oc = plan.halo + 1, plan.cout
    assert a.passes == 1 or oc == 3
    img = b2.synth_image(W, H, 3, seed=1) if rank == 0 else None
    full = torch.from_numpy(img).to(dev) if rank == 0 else None
    band = torch.empty((hb + 2 * halo, W, 3), dtype=torch.uint8, device=dev)   # [halo | band | halo]
    outb = torch.empty((hb + 2 * halo, W, oc), dtype=torch.uint8, device=dev)
    gathered = [torch.empty((hb, W, oc), dtype=torch.uint8, device=dev) for _ in range(world)] if rank == 0 else None

    def step():
        core = band[halo:halo + hb]
        dist.scatter(core, list(full.split(hb, 0)) if rank == 0 else None, src=0)          # MPI_Scatter
        cur, nxt = band, outb
        for p in range(a.passes):
            if halo and world > 1:                                                            # MPI_Sendrecv halos
                ops = []
                if rank > 0:
                    ops += [dist.P2POp(dist.isend, cur[halo:2 * halo].contiguous(), rank - 1),
                            dist.P2POp(dist.irecv, cur[0:halo], rank - 1)]
                if rank + 1 < world:
                    ops += [dist.P2POp(dist.isend, cur[hb:hb + halo].contiguous(), rank + 1),
                            dist.P2POp(dist.irecv, cur[hb + halo:hb + 2 * halo], rank + 1)]
                for r in dist.batch_isend_irecv(ops):
                    r.wait()
            # every rank filters [halo|band|halo] as if it were a whole image; only the band rows are kept,
            # so the artificial borders at the band edges never reach the result (interior ranks)
            lo = 0 if rank > 0 else halo
            hi = hb + 2 * halo if rank + 1 < world else hb + halo
            plan.run(cur[lo:hi], nxt[lo:hi], naive=a.kernel == "naive")
            cur, nxt = nxt, cur
        dist.gather(cur[halo:halo + hb].contiguous(), gathered, dst=0)                        # MPI_Gather
        return cur

    step()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        got = torch.cat(gathered, 0).cpu().numpy()
        want = img
        for _ in range(a.passes):
            want = b2.host_run(a.chain, want)
        print(json.dumps({"impl": "nccl_baseline", "kernel": a.kernel, "chain": a.chain, "passes": a.passes,
                          "n_gpus": world, "image": f"{W}x{H}", "ms_per_step": round(ms.item(), 4),
                          "mpix_per_s": round(W * H / 1e6 / ms.item() * 1e3, 1),
                          "bit_exact": bool(np.array_equal(got, want))}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
