"""ROOT mode, equal bands, under torchrun: step time with the remote ranks on the TMA kernels vs on the
plain ld/st loaders (B2I_LAUNCH_NO_TMA), against the link roofline.
    torchrun --nproc-per-node N scripts/root_link_probe.py [chain]"""
import os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime, _timed

chain = sys.argv[1] if len(sys.argv) > 1 else "kernel_cu"
W, H, K = 7680, 4320, 20
rt = DistRuntime(root=0)
rank, world = rt.rank, rt.world
plan = Plan(chain, 3)
img = b2.synth_image(W, H, 3, seed=1) if rank == 0 else None
rt.setup_root(W, H, 3, plan.cout)
rt.load_root(img)
rt.set_root_share(None)
link_us = (world - 1) / world * W * H * 3 / 770e9 * 1e6
for label, remote_no_tma in (("tma", False), ("plain ld/st on the remote ranks", True)):
    nt = remote_no_tma and rank != 0
    rt.run_root(plan, 3, no_tma=nt)
    rt.barrier()
    ms = statistics.median([_timed(lambda: rt.run_root(plan, K, no_tma=nt)) / K for _ in range(5)])
    if rank == 0:
        print(f"N={world} {chain} equal bands, {label}: {ms * 1e3:.1f} us/step; link roofline {link_us:.1f} us "
              f"({link_us / (ms * 1e3):.2f})", flush=True)
rt.close()
