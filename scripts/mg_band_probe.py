"""One GPU, one resident band with the multi-GPU hooks switched on but pointed at local memory: the band
pushes its halo rows into local dummy rows, signals local words and waits on local words that are already
satisfied. Separates what the hooks cost inside the kernel from what NVLink and the neighbours cost.
    python scripts/mg_band_probe.py [WxROWS ...]      (per-pass launches, 64 passes per call)"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200._lib import Launch, View, check, lib
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan

P, G = 64, 8
chain = os.environ.get("CHAIN", "gauss5")
for size in sys.argv[1:] or ["7680x540", "7680x1080", "7680x2160"]:
    W, R = (int(x) for x in size.split("x"))
    H = 3 * R                     # the band is the middle third of a taller image
    y0, y1 = R, 2 * R
    bp = W * 3
    plan = Plan(chain, 3)
    halo = plan.halo
    bands = [torch.randint(0, 255, (R + 2 * G, bp), dtype=torch.uint8, device="cuda") for _ in range(2)]
    dummy = torch.zeros((4 * G, bp), dtype=torch.uint8, device="cuda")
    words = torch.zeros(4096, dtype=torch.int32, device="cuda")
    words[8:10] = 1 << 20         # the wait flags (bytes 32, 36): every epoch is already there
    res = {}
    for hooks in (0, 1):
        ls = []
        for k in range(2):
            L = Launch()
            L.W, L.H, L.y0, L.y1 = W, H, y0, y1
            L.src = View(bands[k].data_ptr(), bp, y0 - G, y1 + G)
            L.dst = View(bands[k ^ 1].data_ptr() + bp * G, bp, y0, y1)
            L.counters, L.status = words.data_ptr(), words.data_ptr() + 64
            L.stream = torch.cuda.current_stream().cuda_stream
            L.timeout_ns = 2_000_000_000
            if hooks:
                L.push_rows = halo
                L.push_up = View(dummy.data_ptr(), bp, y0, y0 + G)
                L.push_dn = View(dummy.data_ptr() + 2 * G * bp, bp, y1 - halo, y1)
                if not os.environ.get("PROBE_NO_WAIT"):
                    L.wait_up_flag, L.wait_dn_flag = words.data_ptr() + 32, words.data_ptr() + 36
                if not os.environ.get("PROBE_NO_SIGNAL"):
                    L.signal_up_flag, L.signal_dn_flag = words.data_ptr() + 128, words.data_ptr() + 132
                if os.environ.get("PROBE_NO_PUSH"):
                    L.push_rows = 0; L.push_up = View(0, 0, 0, 0); L.push_dn = View(0, 0, 0, 0)
                if os.environ.get("PROBE_ONE_SIDE"):
                    L.push_dn = View(0, 0, 0, 0); L.wait_dn_flag = 0; L.signal_dn_flag = 0
            ls.append(L)
        ep = [0]
        def go():
            check(lib.b2i_cuda_run_resident_steps(plan.handle, C.byref(ls[0]), C.byref(ls[1]), P, ep[0]))
            ep[0] += P + 1
        for _ in range(3):
            go()
        torch.cuda.synchronize()
        e0, e1, reps = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), []
        for _ in range(7):
            e0.record(); go(); go(); e1.record(); torch.cuda.synchronize()
            reps.append(e0.elapsed_time(e1) / (2 * P) * 1e3)
        res[hooks] = statistics.median(reps)
    i = plan.info()
    print(f"{chain} {W}x{R}-row band: no hooks {res[0]:.1f} us/pass, hooks to local memory {res[1]:.1f} us/pass "
          f"(seg_rows {-i['tile_h']}, grid {i['grid']}, status {int(words[16])})", flush=True)
