import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
W, H = 7680, 4320
img = torch.from_numpy(b2.synth_image(W, H, 3, seed=1)).cuda()
for chain in sys.argv[1:] or ["gauss5", "gauss5,sobel", "gray,gauss5", "sobel"]:
    plan = Plan(chain, 3)
    out = plan.run(img)
    for th in (0, 8, 16, 24, 32):
        try:
            for _ in range(3): plan.run(img, out, tile_h=th)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10): plan.run(img, out, tile_h=th)
            e1.record(); torch.cuda.synchronize()
            i = plan.info()
            print(f"{chain:16s} tile_h={th:2d} -> {e0.elapsed_time(e1)/10*1e3:8.1f} us  (th {i['tile_h']}, smem {i['smem_bytes']}, bps {i['blocks_per_sm']}, grid {i['grid']})", flush=True)
        except Exception as e:
            print(chain, th, "failed:", str(e)[:100])
