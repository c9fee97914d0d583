"""Runs small GPU cases, each in its own process (a faulting kernel poisons the CUDA context)."""
import os
import subprocess
import sys

CASE = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import cuda_run
chain, W, H, C, kw = %r
img = b2.synth_image(W, H, C, seed=1)
out = cuda_run(chain, torch.from_numpy(img).cuda(), **kw)
torch.cuda.synchronize()
want = b2.host_run(chain, img)
got = out.cpu().numpy()
bad = np.argwhere(got != want)
print("MISMATCH %%d first %%s" %% (len(bad), bad[:3].tolist()) if len(bad) else "OK")
'''
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
cases = [
    ("gray", 1024, 256, 3, {}),
    ("kernel_cu", 1024, 256, 3, {"naive": True}),
    ("kernel_cu", 1024, 256, 3, {"no_tma": True}),
    ("kernel_cu", 1024, 256, 3, {}),
    ("emboss3", 1024, 256, 1, {"no_tma": True}),
    ("emboss3", 1024, 256, 1, {}),
    ("gauss5", 1024, 256, 3, {}),
]
if len(sys.argv) > 1:
    cases = [cases[int(i)] for i in sys.argv[1:]]
for c in cases:
    r = subprocess.run([sys.executable, "-c", CASE % (ROOT, c)], capture_output=True, text=True, timeout=120)
    tail = (r.stdout.strip().splitlines() or [""])[-1]
    err = [l for l in r.stderr.splitlines() if "rror" in l][-1:] if r.returncode else []
    print(c, "->", tail, err, flush=True)
