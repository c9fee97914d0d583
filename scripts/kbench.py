"""Device time per launch of chains on an 8K (or given) image: K launches in one CUDA graph over
rotating buffers (no host in the timed region, traffic between reuses > L2), median of replays.
    python scripts/kbench.py kernel_cu gauss5 sobel gauss5,sobel [--size 7680x4320] [--k 40]"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan

args = [a for a in sys.argv[1:] if not a.startswith("--")]
opt = {a.split("=")[0]: a.split("=")[1] for a in sys.argv[1:] if a.startswith("--") and "=" in a}
W, H = (int(x) for x in opt.get("--size", "7680x4320").split("x"))
K, NB = int(opt.get("--k", 40)), 4
img = b2.synth_image(W, H, 3, seed=1)
for chain in args or ["kernel_cu", "gauss5", "sobel", "gauss5,sobel"]:
    plan = Plan(chain, 3)
    srcs = [torch.from_numpy(img).cuda() for _ in range(NB)]
    outs = [plan.run(srcs[i]) for i in range(NB)]
    torch.cuda.synchronize()
    g, s = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s), torch.cuda.graph(g, stream=s):
        for i in range(K):
            plan.run(srcs[i % NB], outs[i % NB])
    torch.cuda.synchronize()
    g.replay(); torch.cuda.synchronize()
    e0, e1, reps = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), []
    for _ in range(7):
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        reps.append(e0.elapsed_time(e1) / K * 1e3)
    i = plan.info()
    print(f"{chain:20s} {statistics.median(reps):8.1f} us  (min {min(reps):.1f}; th {i['tile_h']}, grid {i['grid']}, launches {i['launches']})", flush=True)
    del g
