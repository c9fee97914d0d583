"""Small driver for ncu: runs one chain a few times on an 8K (or given) image, no CUDA graphs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan

chain = sys.argv[1] if len(sys.argv) > 1 else "kernel_cu"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 7680
H = int(sys.argv[3]) if len(sys.argv) > 3 else 4320
C = int(sys.argv[4]) if len(sys.argv) > 4 else 3
n = int(sys.argv[5]) if len(sys.argv) > 5 else 6
img = torch.from_numpy(b2.synth_image(W, H, C, seed=1)).cuda()
plan = Plan(chain, C)
out = None
for _ in range(n):
    out = plan.run(img, out)
torch.cuda.synchronize()
print("ok", chain, W, H, plan.info())
