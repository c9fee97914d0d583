"""Run under torchrun (one rank per GPU): multi-process correctness of the fused scatter/gather
(ROOT mode) and of the halo push with device flags (RESIDENT mode) against the host reference."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime


def main():
    rt = DistRuntime(root=0)
    rank, world = rt.rank, rt.world
    fails = 0
    cases = [("kernel_cu", 2048, 1030), ("kern_cpp", 1024, 403), ("gauss5,sobel", 1024, 403),
             ("gray", 2048, 777), ("invert", 1001, 203), ("kernel_cu", 1001, 77),
             ("gray,gauss5,gauss5,gauss5,gauss5,gauss5,sobel,expand", 1024, 400)]
    for chain, W, H in cases:
        img = b2.synth_image(W, H, 3, seed=W + H)
        plan = Plan(chain, 3)
        rt.setup_root(W, H, 3, plan.cout)
        rt.load_root(img if rank == 0 else None)
        for no_tma in (False, True):
            rt.run_root(plan, 2, no_tma=no_tma)
            got = rt.fetch_root()
            if rank == 0:
                ok = np.array_equal(got, b2.host_run(chain, img))
                fails += not ok
                print(f"root  {chain:28s} {W}x{H} no_tma={no_tma}: {'OK' if ok else 'MISMATCH'} tma={plan.info()['tma_load']}", flush=True)
    for chain, W, H, passes in [("gauss5", 1024, 330, 1), ("gauss5", 1024, 330, 4), ("gauss5,sobel", 2048, 500, 3),
                                ("emboss3@reflect101", 1001, 200, 5), ("invert", 512, 100, 3)]:
        img = b2.synth_image(W, H, 3, seed=passes)
        plan = Plan(chain, 3)
        rt.setup_root(W, H, 3, 3)
        rt.load_root(img if rank == 0 else None)
        rt.setup_resident(W, H, 3)
        rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
        rt.run_resident(plan, passes)
        rt.gather_resident(passes, rt.dst.ptrs[0], rt.geom["dp"], 3)
        got = rt.fetch_root()
        st = rt.status()
        print(f"   [rank {rank}] {chain} status={st} epoch={rt.epoch}", flush=True)
        if rank == 0:
            want = img
            for _ in range(passes):
                want = b2.host_run(chain, want)
            ok = np.array_equal(got, want) and st == 0
            fails += not ok
            print(f"resid {chain:28s} {W}x{H} passes={passes}: {'OK' if ok else 'MISMATCH'} status={st}", flush=True)
    # halo rows exchanged every `cycle` passes only (run_resident_cycle): the whole image, bit for bit
    for chain, W, H, passes, cycle in [("gauss5", 1024, 330, 4, 2), ("gauss5", 1024, 330, 8, 4), ("gauss3", 2048, 500, 8, 4),
                                       ("gauss5", 7680, 1080, 4, 4), ("gauss5@replicate", 2048, 64 * world + 3, 6, 2)]:
        img = b2.synth_image(W, H, 3, seed=passes + cycle)
        plan = Plan(chain, 3)
        rt.setup_root(W, H, 3, 3)
        rt.load_root(img if rank == 0 else None)
        rt.setup_resident(W, H, 3)
        rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
        rt.run_resident_cycle(plan, passes, cycle)
        rt.gather_resident(passes, rt.dst.ptrs[0], rt.geom["dp"], 3)
        got = rt.fetch_root()
        st = rt.status()
        if rank == 0:
            want = img
            for _ in range(passes):
                want = b2.host_run(chain, want)
            ok = np.array_equal(got, want) and st == 0
            fails += not ok
            print(f"cycle {chain:28s} {W}x{H} passes={passes} exchange every {cycle}: {'OK' if ok else 'MISMATCH'} status={st}", flush=True)
    rt.close()
    if rank == 0:
        print("DIST_CHECK", "PASS" if fails == 0 else f"FAIL({fails})", flush=True)
    sys.exit(1 if fails else 0)


if __name__ == "__main__":
    main()
