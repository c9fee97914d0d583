"""torchrun probe (2+ GPUs): where does ROOT mode spend its time? Times the non-root rank's band kernel
with the source and/or destination on the root GPU (over NVLink) or local."""
import os, sys, ctypes as C, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime
from mpi_cuda_imagemanipulation_b200._lib import lib, check, View
rt = DistRuntime(root=0)
W, H = 7680, 4320
chain = sys.argv[1] if len(sys.argv) > 1 else "kernel_cu"
plan = Plan(chain, 3)
rt.setup_root(W, H, 3, plan.cout)
rt.load_root(b2.synth_image(W, H, 3, seed=1) if rt.rank == 0 else None)
K = 20
for src_remote, dst_remote in ((0, 0), (1, 0), (0, 1), (1, 1)):
    L = rt._root_launch()
    if not src_remote: L.src = View(rt.src.local, rt.geom["sp"], 0, H)
    if not dst_remote: L.dst = View(rt.dst.local, rt.geom["dp"], 0, H)
    for env in ("", "B200IMG_NO_SWEEP"):
        if env: os.environ[env] = "1"
        def run():
            for _ in range(K): check(lib.b2i_cuda_run(plan.handle, C.byref(L)))
        run(); torch.cuda.synchronize(); dist.barrier()
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); dist.barrier()
            if rt.rank != 0 or True:
                e0.record(); run(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) / K)
        if env: del os.environ[env]
        rows = rt.y1 - rt.y0
        mb = W * rows * 3 / 1e6
        print(f"rank{rt.rank} {chain} src_remote={src_remote} dst_remote={dst_remote} {'tile ' if env else 'sweep'}: "
              f"{statistics.median(ts)*1e3:.1f} us  ({mb/statistics.median(ts)/1e3:.0f} GB/s per direction)", flush=True)
rt.close()
