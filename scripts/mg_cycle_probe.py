"""One GPU, one resident band: the exchange cycle of parallel/dist.py (run_resident_cycle) with the hooks
pointed at local memory, against one launch per pass with hooks and without. What each kind of pass costs
without NVLink.   python scripts/mg_cycle_probe.py [WxROWS ...]"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
from mpi_cuda_imagemanipulation_b200._lib import Launch, View, check, lib
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan

P, G = 64, 8
for size in sys.argv[1:] or ["7680x540"]:
    W, R = (int(x) for x in size.split("x"))
    H, y0, y1, bp = 3 * R, R, 2 * R, W * 3
    plan = Plan("gauss5", 3)
    halo = plan.halo
    bands = [torch.randint(0, 255, (R + 2 * G, bp), dtype=torch.uint8, device="cuda") for _ in range(4)]
    dummy = torch.zeros((4 * G, bp), dtype=torch.uint8, device="cuda")
    words = torch.zeros(4096, dtype=torch.int32, device="cuda")
    words[8:10] = 1 << 20

    def launch(inb, outb, ext=0, push_rows=0, push=False, wait=False, signal=False):
        L = Launch()
        L.W, L.H, L.y0, L.y1 = W, H, y0 - ext, y1 + ext
        L.src = View(bands[inb].data_ptr(), bp, y0 - G, y1 + G)
        L.dst = View(bands[outb].data_ptr() + bp * (G - ext), bp, y0 - ext, y1 + ext)
        L.counters, L.status = words.data_ptr(), words.data_ptr() + 64
        L.stream = torch.cuda.current_stream().cuda_stream
        L.timeout_ns = 2_000_000_000
        L.push_rows = push_rows
        if push:
            L.push_up = View(dummy.data_ptr(), bp, y0, y0 + G)
            L.push_dn = View(dummy.data_ptr() + 2 * G * bp, bp, y1 - push_rows, y1)
        if wait:
            L.wait_up_flag, L.wait_dn_flag = words.data_ptr() + 32, words.data_ptr() + 36
        if signal:
            L.signal_up_flag, L.signal_dn_flag = words.data_ptr() + 128, words.data_ptr() + 132
        return L

    def cycle_descs(n):
        d = (Launch * (2 * n))()
        for c in range(2):
            pair, other = ((0, 1), 2) if c == 0 else ((2, 3), 0)
            for j in range(n):
                if j == n - 1:
                    d[c * n + j] = launch(pair[j & 1], other, 0, n * halo, push=True, signal=True)
                elif j == 0:
                    d[c * n + j] = launch(pair[0], pair[1], (n - 1) * halo, n * halo, wait=True)
                else:
                    d[c * n + j] = launch(pair[j & 1], pair[(j & 1) ^ 1], (n - 1 - j) * halo)
        return d

    def one(descs, period):
        ep = [0]
        def go():
            check(lib.b2i_cuda_run_resident_cycle(plan.handle, descs, period, P, ep[0]))
            ep[0] += P + 1
        for _ in range(3):
            go()
        torch.cuda.synchronize()
        e0, e1, reps = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), []
        for _ in range(7):
            e0.record(); go(); go(); e1.record(); torch.cuda.synchronize()
            reps.append(e0.elapsed_time(e1) / (2 * P) * 1e3)
        return statistics.median(reps)

    def pair_descs(**kw):
        d = (Launch * 2)()
        d[0], d[1] = launch(0, 1, **kw), launch(1, 0, **kw)
        return d

    res = {
        "plain": one(pair_descs(), 2),
        "plain, 6 rows wider": one(pair_descs(ext=6), 2),
        "wait only": one(pair_descs(ext=6, push_rows=8, wait=True), 2),
        "wait only, push_rows 2, not wider": one(pair_descs(push_rows=2, wait=True), 2),
        "wait only, push_rows 8, not wider": one(pair_descs(push_rows=8, wait=True), 2),
        "signal only, push_rows 2": one(pair_descs(push_rows=2, signal=True), 2),
        "wait + signal, push_rows 2": one(pair_descs(push_rows=2, wait=True, signal=True), 2),
        "push 2 + signal + wait (one launch per pass)": one(pair_descs(push_rows=2, push=True, wait=True, signal=True), 2),
        "push 4 + signal": one(pair_descs(push_rows=4, push=True, signal=True), 2),
        "push 8 + signal": one(pair_descs(push_rows=8, push=True, signal=True), 2),
        "cycle of 2": one(cycle_descs(2), 4),
        "cycle of 4": one(cycle_descs(4), 8),
    }
    print(f"gauss5 {W}x{R}-row band, us per pass (status {int(words[16])}):")
    for k, v in res.items():
        print(f"  {k:46s} {v:6.1f}", flush=True)
