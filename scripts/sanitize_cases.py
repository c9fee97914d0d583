"""A few small launches of every kernel family, for compute-sanitizer (one tool per gpurun call):
    compute-sanitizer --tool memcheck python scripts/sanitize_cases.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan, run_passes

CASES = [("kernel_cu", 3, 512, 96), ("kern_cpp", 3, 496, 70), ("gray,contrast:3.5,expand", 3, 1000, 33), ("invert", 3, 333, 21),
         ("gauss5", 3, 512, 64), ("sobel", 3, 512, 64), ("sharpen", 3, 512, 140), ("gauss5", 1, 640, 140),
         ("emboss5", 3, 500, 61), ("gray,gauss5,sobel,expand", 3, 528, 150)]
bad = 0
for chain, c, W, H in CASES:
    img = b2.synth_image(W, H, c, seed=4)
    got = Plan(chain, c).run(torch.from_numpy(img).cuda())
    torch.cuda.synchronize()
    ok = np.array_equal(got.cpu().numpy().reshape(H, -1), b2.host_run(chain, img).reshape(H, -1))
    bad += not ok
    print(f"{chain:28s} c{c} {W}x{H}: {'ok' if ok else 'MISMATCH'}", flush=True)
img = b2.synth_image(960, 200, 3, seed=4)
got = run_passes(Plan("gauss5", 3), torch.from_numpy(img).cuda(), 3, fused=True)
ok = np.array_equal(got.cpu().numpy(), b2.host_run("gauss5,gauss5,gauss5", img))
bad += not ok
print(f"fused passes: {'ok' if ok else 'MISMATCH'}", flush=True)
sys.exit(1 if bad else 0)
