"""Per-pass device time of P resident gauss5 passes on one GPU: fused (one cooperative launch) vs one
launch per pass. python scripts/passes_bench.py [WxH ...]"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200._lib import Launch, View, check, lib
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan

P = 8
for size in sys.argv[1:] or ["7680x540", "7680x1080", "7680x2160", "7680x4320"]:
    W, H = (int(x) for x in size.split("x"))
    img = torch.from_numpy(b2.synth_image(W, H, 3, seed=1)).cuda()
    bufs = [img.clone(), torch.empty_like(img)]
    words = torch.zeros(64, dtype=torch.int32, device="cuda")
    flags = torch.zeros(16384, dtype=torch.int32, device="cuda")
    plan = Plan("gauss5", 3)
    ls = []
    for k in range(2):
        L = Launch()
        L.W, L.H, L.y0, L.y1 = W, H, 0, H
        L.src = View(bufs[k].data_ptr(), W * 3, 0, H)
        L.dst = View(bufs[k ^ 1].data_ptr(), W * 3, 0, H)
        L.counters, L.status = words.data_ptr(), words.data_ptr() + 16
        L.stream = torch.cuda.current_stream().cuda_stream
        ls.append(L)
    res = {}
    for name in ("fused", "steps"):
        ep = [0]
        def go():
            if name == "fused":
                check(lib.b2i_cuda_run_resident_fused(plan.handle, C.byref(ls[0]), C.byref(ls[1]), P, ep[0],
                                                      flags.data_ptr(), flags.numel()))
            else:
                check(lib.b2i_cuda_run_resident_steps(plan.handle, C.byref(ls[0]), C.byref(ls[1]), P, ep[0]))
            ep[0] += P + 1
        for _ in range(3):
            go()
        torch.cuda.synchronize()
        e0, e1, reps = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), []
        for _ in range(9):
            e0.record(); go(); go(); go(); go(); e1.record(); torch.cuda.synchronize()
            reps.append(e0.elapsed_time(e1) / (4 * P) * 1e3)
        res[name] = statistics.median(reps)
    i = plan.info()
    print(f"{size}: fused {res['fused']:.1f} us/pass, per-pass launches {res['steps']:.1f} us/pass "
          f"(seg_rows {-i['tile_h']}, grid {i['grid']}, status {int(words[4])})", flush=True)
