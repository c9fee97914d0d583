import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime
rt = DistRuntime(root=0)
chain, W, H, passes = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
no_tma = len(sys.argv) > 5 and sys.argv[5] == "no_tma"
img = b2.synth_image(W, H, 3, seed=passes)
plan = Plan(chain, 3)
rt.setup_root(W, H, 3, 3); rt.load_root(img if rt.rank == 0 else None)
rt.setup_resident(W, H, 3)
want = img
for _ in range(passes): want = b2.host_run(chain, want)
for it in range(4):
    rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
    t0 = time.time()
    rt.run_resident(plan, passes, no_tma=no_tma)
    torch.cuda.synchronize(); dt = time.time() - t0
    st = rt.status()
    rt.gather_resident(passes, rt.dst.ptrs[0], rt.geom["dp"], 3)
    got = rt.fetch_root()
    print(f"rank{rt.rank} iter {it}: {dt*1e3:.2f} ms status={st} " + (f"ok={np.array_equal(got, want)} {plan.info()}" if rt.rank == 0 else ""), flush=True)
rt.close()
