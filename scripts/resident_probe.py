"""Where does a resident pass go when bands are small? 8K-wide images of a few heights, N = 1..gpus,
events vs flags (in-process runtime). us per pass."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.parallel import Runtime
chain = sys.argv[1] if len(sys.argv) > 1 else "gauss5"
for H in (540, 1080, 4320):
    img = b2.synth_image(7680, H, 3, seed=1)
    for n in (1, 2, 4, 8):
        if n > torch.cuda.device_count():
            break
        with Runtime(gpus=n) as rt:
            rt.run(img, chain, mode="resident", passes=4)
            rt.run(img, chain, mode="resident", passes=64)
            e = rt.last_ms
            rt.run(img, chain, mode="resident", passes=64, sync="flags")
            f = rt.last_ms
        print(json.dumps({"H": H, "n": n, "rows_per_gpu": H // n, "events_us": round(e * 1e3, 1), "flags_us": round(f * 1e3, 1)}), flush=True)
