"""torchrun probe: aggregate host<->device bandwidth with all ranks copying at once, from (a) private
pinned buffers and (b) one POSIX-shm buffer registered by every rank (the root-owned host image)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime
from mpi_cuda_imagemanipulation_b200._lib import lib, check
rt = DistRuntime()
MB = int(sys.argv[1]) if len(sys.argv) > 1 else 100
n = MB << 20
dev = rt.device
d_a = torch.empty(n, dtype=torch.uint8, device=dev); d_b = torch.empty(n, dtype=torch.uint8, device=dev)
priv_in = torch.empty(n, dtype=torch.uint8).pin_memory(); priv_out = torch.empty(n, dtype=torch.uint8).pin_memory()
sh_in = rt.shared_host("pin", n * rt.world); sh_out = rt.shared_host("pout", n * rt.world)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
def run(src_ptr, dst_ptr, reps=10):
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); s1.wait_event(e0); s2.wait_event(e0)
    for _ in range(reps):
        if src_ptr: check(lib.b2i_cuda_memcpy2d(d_a.data_ptr(), n, src_ptr, n, n, 1, 0, s1.cuda_stream))
        if dst_ptr: check(lib.b2i_cuda_memcpy2d(dst_ptr, n, d_b.data_ptr(), n, n, 1, 1, s2.cuda_stream))
    j1, j2 = torch.cuda.Event(), torch.cuda.Event(); j1.record(s1); j2.record(s2)
    torch.cuda.current_stream().wait_event(j1); torch.cuda.current_stream().wait_event(j2)
    e1.record(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return reps * n / (ms.item() / 1e3) / 1e9   # GB/s per rank per direction
for name, i, o in (("private pinned", priv_in.data_ptr(), priv_out.data_ptr()),
                   ("shared shm    ", sh_in.ctypes.data + n * rt.rank, sh_out.ctypes.data + n * rt.rank)):
    h2d, d2h, both = run(i, 0), run(0, o), run(i, o)
    if rt.rank == 0:
        print(f"{name} world={rt.world}: per-rank GB/s  H2D {h2d:.1f}  D2H {d2h:.1f}  both {both:.1f}+{both:.1f}; "
              f"aggregate both {both*rt.world:.0f}+{both*rt.world:.0f}", flush=True)
rt.close()
