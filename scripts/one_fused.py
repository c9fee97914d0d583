"""One fused multi-pass launch (the multi-GPU instantiation of sweep5 on one GPU), for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan, run_passes
W, H, P = 7680, int(sys.argv[1]) if len(sys.argv) > 1 else 4320, int(sys.argv[2]) if len(sys.argv) > 2 else 4
img = torch.from_numpy(b2.synth_image(W, H, 3, seed=1)).cuda()
out = run_passes(Plan("gauss5", 3), img, P, fused=True)
print("fused passes done", out.shape)
