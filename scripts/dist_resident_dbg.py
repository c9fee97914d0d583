import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel.dist import DistRuntime
rt = DistRuntime(root=0)
W, H = int(sys.argv[1]), int(sys.argv[2])
P = int(sys.argv[3])
img = b2.synth_image(W, H, 3, seed=2)
plan = Plan("gauss5", 3)
rt.setup_root(W, H, 3, 3); rt.load_root(img if rt.rank == 0 else None)
if os.environ.get("PRE_ROOT"):
    kp = Plan("kernel_cu", 3)
    for _ in range(10):
        rt.run_root(kp, 30)
    rt.barrier()
    print(f"rank{rt.rank} pre-root done, epoch={rt.epoch} status={rt.status()}", flush=True)
rt.setup_resident(W, H, 3)
for it in range(3):
    if it == 0 or not os.environ.get("NO_RESCATTER"):
        rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
    else:
        rt.barrier()
    t0 = time.time()
    rt.run_resident(plan, P)
    torch.cuda.synchronize()
    dt = time.time() - t0
    st = rt.status()
    rt.gather_resident(P, rt.dst.ptrs[0], rt.geom["dp"], 3)
    got = rt.fetch_root()
    if rt.rank == 0:
        want = img
        for _ in range(P): want = b2.host_run("gauss5", want)
        print(f"iter {it}: {dt*1e3:.2f} ms status={st} ok={np.array_equal(got, want)} info={plan.info()['tile_h']},{plan.info()['grid']}", flush=True)
    else:
        print(f"  rank{rt.rank} iter {it}: {dt*1e3:.2f} ms status={st}", flush=True)
rt.close()
