"""torchrun: ROOT-mode step time of the reference pipeline on 8K for a sweep of root shares, with
each rank's own kernel time next to the step time (where does the step go?)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch, torch.distributed as dist
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime, _timed, check, lib
rt = DistRuntime(root=0)
W, H, K = 7680, 4320, 40
plan = Plan("kernel_cu", 3)
rt.setup_root(W, H, 3, 3)
rt.load_root(b2.synth_image(W, H, 3, seed=1) if rt.rank == 0 else None)
for share in [None] + [float(x) for x in sys.argv[1:]]:
    rt.set_root_share(share)
    rt.run_root(plan, 5); rt.barrier()
    step = min(_timed(lambda: rt.run_root(plan, K)) / K for _ in range(3))
    # own kernels only, all ranks concurrently, no flags
    L = rt._root_launch(False, plan)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    own = 0.0
    if rt.y1 > rt.y0:
        torch.cuda.synchronize(); dist.barrier()
        e0.record()
        for _ in range(K): check(lib.b2i_cuda_run(plan.handle, C.byref(L)))
        e1.record(); torch.cuda.synchronize()
        own = e0.elapsed_time(e1) / K
    else:
        dist.barrier()
    t = torch.zeros(rt.world, device="cuda"); t[rt.rank] = own
    dist.all_reduce(t)
    if rt.rank == 0:
        print(f"share={share}: step {step*1e3:.1f} us; own kernels (concurrent) " + " ".join(f"{x*1e3:.1f}" for x in t.tolist()) + f" us; rows {[b[1]-b[0] for b in rt.bands]}", flush=True)
rt.close()
