"""Device time per launch on widths that are not multiples of 8 pixels, in pitched buffers, next to the
aligned width: python scripts/odd_width_bench.py"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan, pitched_empty

K, NB = 40, 4
for chain, c in (("kernel_cu", 3), ("sharpen", 3), ("gauss5", 1)):
    for W, H in ((7680, 4320), (7681, 4320), (1000, 1000), (1001, 1001)):
        img = torch.from_numpy(b2.synth_image(W, H, c, seed=1)).cuda()
        plan = Plan(chain, c)
        srcs, outs = [], []
        for _ in range(NB):
            s = pitched_empty(H, W, c); s.copy_(img.reshape(H, W, c)); srcs.append(s)
            outs.append(pitched_empty(H, W, plan.cout))
        for i in range(NB):
            plan.run(srcs[i], outs[i])
        torch.cuda.synchronize()
        g, st = torch.cuda.CUDAGraph(), torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st), torch.cuda.graph(g, stream=st):
            for i in range(K):
                plan.run(srcs[i % NB], outs[i % NB])
        torch.cuda.synchronize(); g.replay(); torch.cuda.synchronize()
        e0, e1, reps = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), []
        for _ in range(7):
            e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
            reps.append(e0.elapsed_time(e1) / K * 1e3)
        us = statistics.median(reps)
        print(f"{chain:10s} c{c} {W}x{H}: {us:7.1f} us  {us * 1e3 / (W * H / 1e3):7.3f} ps/pixel  (th {plan.info()['tile_h']})", flush=True)

# tightly packed odd widths: staged through pitched buffers (default) vs filtered in place (plain loaders)
for chain, c in (("kernel_cu", 3), ("gauss5", 3)):
    W, H = 7681, 4320
    img = torch.from_numpy(b2.synth_image(W, H, c, seed=1)).cuda()
    plan = Plan(chain, c)
    out = torch.empty((H, W, plan.cout), dtype=torch.uint8, device="cuda")
    for stage in (True, False):
        for _ in range(3):
            plan.run(img, out, stage=stage)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            plan.run(img, out, stage=stage)
        e1.record(); torch.cuda.synchronize()
        print(f"{chain:10s} packed {W}x{H} stage={stage}: {e0.elapsed_time(e1) / 10 * 1e3:7.1f} us", flush=True)
