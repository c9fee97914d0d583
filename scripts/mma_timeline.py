"""Timeline of blocks 0 and 1 of the tensor-core stencil kernel (globaltimer stamps, ns from kernel start).
    python scripts/mma_timeline.py [chain] (B200IMG_MMA_DBG=... for the experiments)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
buf = torch.zeros(64, dtype=torch.int64, device="cuda")
os.environ["B200IMG_MMA_DBGPTR"] = str(buf.data_ptr())
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
chain = sys.argv[1] if len(sys.argv) > 1 else "gauss5"
img = torch.from_numpy(b2.synth_image(7680, 4320, 3, seed=1)).cuda()
plan = Plan(chain, 3)
out = torch.empty_like(img)
names = ["start", "after sync1", "mma: setup done", "mma: first slab full", "mma: all issued", "epi: first chunk full",
         "epi: last store issued", "epi: stores done", "producer: last load issued", "end"]
for rep in range(3):
    buf.zero_()
    plan.run(img, out)
    torch.cuda.synchronize()
    t = buf.cpu().numpy().reshape(4, 16)
    if rep == 2:
        for b in range(2):
            t0 = t[b][0]
            print(f"block {b}: " + ", ".join(f"{names[i]}={int(t[b][i] - t0)}" for i in range(10)))
