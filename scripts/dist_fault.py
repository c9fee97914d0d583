"""Fault injection under torchrun: the LAST rank skips a resident run (a lost peer). The others must
not hang: their waits give up after 50 ms, every rank then learns of it (check_status raises
B2I_ERR_TIMEOUT everywhere), the flags are reset, and the next run is correct again.
The reference's behaviour in the same situation: the surviving ranks sit in MPI_Barrier for ever
(kernel.cu:111-114,127)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200 import _lib
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime

rt = DistRuntime(root=0)
rank, world = rt.rank, rt.world
W, H, P = 1920, 1080, 4
img = b2.synth_image(W, H, 3, seed=11)
plan = Plan("gauss5", 3)
rt.setup_root(W, H, 3, 3)
rt.load_root(img if rank == 0 else None)
rt.setup_resident(W, H, 3)
fails = 0
for fused in (True, False):
    rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
    rt.timeout_ns = 50_000_000
    t0 = time.time()
    if rank != world - 1:
        rt.run_resident(plan, P, fused=fused)     # the last rank never shows up
    else:
        rt.epoch += P + 1
    torch.cuda.synchronize()
    dt = time.time() - t0
    raised = False
    try:
        rt.check_status()
    except _lib.B200ImgError as e:
        raised = e.code == _lib.B2I_ERR_TIMEOUT
    ok = raised and dt < 5.0
    print(f"[rank {rank}] fused={fused}: lost peer -> B2I_ERR_TIMEOUT raised={raised} after {dt * 1e3:.0f} ms", flush=True)
    fails += not ok
    # the runtime is usable again: same passes, everyone present
    rt.timeout_ns = 0
    rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
    rt.run_resident(plan, P, fused=fused)
    rt.gather_resident(P, rt.dst.ptrs[0], rt.geom["dp"], 3)
    got = rt.fetch_root()
    rt.check_status()
    if rank == 0:
        good = np.array_equal(got, b2.host_run(",".join(["gauss5"] * P), img))
        print(f"after the reset, fused={fused}: {'OK' if good else 'MISMATCH'}", flush=True)
        fails += not good
t = torch.tensor([fails], device=rt.device)
torch.distributed.all_reduce(t)
rt.close()
if rank == 0:
    print("DIST_FAULT", "PASS" if int(t.item()) == 0 else "FAIL", flush=True)
sys.exit(1 if int(t.item()) else 0)
