"""Config 3 only (8K gauss5, bands resident, halo push): in-process runtime at N = 1..gpus."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.parallel import Runtime
chain = sys.argv[1] if len(sys.argv) > 1 else "gauss5"
W, H = 7680, 4320
img = b2.synth_image(W, H, 3, seed=1)
for n in (1, 2, 4, 8):
    if n > torch.cuda.device_count():
        break
    with Runtime(gpus=n) as rt:
        rt.run(img, chain, mode="resident", passes=4)
        rt.run(img, chain, mode="resident", passes=32)
        e = rt.last_ms
        rt.run(img, chain, mode="resident", passes=32, sync="flags")
        f = rt.last_ms
    print(json.dumps({"chain": chain, "n": n, "us_per_pass_events": round(e * 1e3, 1), "us_per_pass_flags": round(f * 1e3, 1)}), flush=True)
