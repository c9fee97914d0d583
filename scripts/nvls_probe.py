"""ROOT mode on 8K, reference pipeline: result to the root only vs replicated on every GPU through
the NVLS multicast image, for a few root shares (in-process runtime). us per step, max over GPUs."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.parallel import Runtime
chain = sys.argv[1] if len(sys.argv) > 1 else "kernel_cu"
img = b2.synth_image(7680, 4320, 3, seed=1)
want = b2.host_run(chain, img, rows=(2000, 2064))
for n in (2, 4, 8):
    if n > torch.cuda.device_count():
        break
    with Runtime(gpus=n) as rt:
        for share in (None, 0.9):
            row = {"n": n, "root_share": share}
            for rep in (False, True):
                rt.run(img, chain, mode="root", iters=3, root_share=share, replicate=rep)
                out = rt.run(img, chain, mode="root", iters=20, root_share=share, replicate=rep)
                ok = bool(np.array_equal(out[2000:2064], want))
                if rep:
                    ok = ok and all(np.array_equal(r[2000:2064], want) for r in rt.replicas)
                    row["nvls"] = bool(rt.nvls)
                row["replicated_us" if rep else "root_only_us"] = round(rt.last_ms * 1e3, 1)
                row["ok_rep" if rep else "ok"] = ok
            print(json.dumps(row), flush=True)
