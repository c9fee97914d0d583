#!/usr/bin/env python
"""Measures the five BASELINE.json configurations (one process, 1..N GPUs through the in-process
C++ runtime) and the reference-style baselines next to them. Writes one JSON document.

    python bench_configs.py --gpus 8 --out profiles/configs_r1.json

Device times are CUDA-event times per GPU, max over GPUs (b2i_ctx_run_*); config 5 is end to end
by definition (frames start and end in pinned host memory).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np
import torch

import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
from mpi_cuda_imagemanipulation_b200.parallel import Runtime
from mpi_cuda_imagemanipulation_b200.parallel.batch import BatchStreamer

PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    pass
HBM = PEAKS.get("hbm_gbs", 6650.0)
NVL = 770.0  # GB/s per direction per GPU, measured peer copy (B200_PROFILING.md)


def naive_ms(chain, img, iters=5):
    """Reference-style kernels (one launch per op, one thread per pixel) on one GPU."""
    t = torch.from_numpy(img).cuda()
    p = Plan(chain, img.shape[2])
    out = p.run(t, naive=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        p.run(t, out, naive=True)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=torch.cuda.device_count())
    ap.add_argument("--out", default="")
    ap.add_argument("--frames", type=int, default=256)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--ns", default="", help="comma-separated GPU counts to measure (default 1,2,4,8 up to --gpus)")
    ap.add_argument("--only", default="", help="comma-separated config numbers to run (default all)")
    a = ap.parse_args()
    ngpu = max(1, min(a.gpus, torch.cuda.device_count()))
    Ns = [n for n in (1, 2, 4, 8) if n <= ngpu]
    if a.ns:
        Ns = [int(x) for x in a.ns.split(",") if int(x) <= ngpu]
    only = {int(x) for x in a.only.split(",")} if a.only else {1, 2, 3, 4, 5}
    res = {"gpus_available": ngpu, "hbm_gbs": HBM, "nvlink_gbs": NVL, "configs": {}}

    # 1. 512x512 RGB, grayscale, 2 host ranks, no GPU ------------------------------------------------
    if 1 in only:
        img = b2.synth_image(512, 512, 3, seed=1)
        best = 1e9
        for _ in range(20):
            t0 = time.perf_counter()
            out = b2.host_run("gray", img, ranks=2)
            best = min(best, time.perf_counter() - t0)
        res["configs"]["1_host_512_gray_np2"] = {"ms": round(best * 1e3, 4), "mpix_per_s": round(0.262144 / best, 1),
                                                 "bit_exact": bool(np.array_equal(out, b2.host_run("gray", img)))}

    W, H = 7680, 4320
    img = b2.synth_image(W, H, 3, seed=1)
    mp = W * H / 1e6
    # 2. 8K gray / invert / reference pipeline, root-owned scatter -> filter -> gather ---------------
    if 2 in only:
        c2 = {}
        for chain, bpp in (("gray", 4), ("invert", 6), ("kernel_cu", 6)):
            row = {"naive_1gpu_ms": round(naive_ms(chain, img), 4)}
            for n in Ns:
                with Runtime(gpus=n) as rt:
                    rt.run(img, chain, mode="root", iters=3)
                    rt.run(img, chain, mode="root", iters=20)
                    ms_equal = rt.last_ms
                    rt.run(img, chain, mode="root", iters=20, root_share="auto")   # link-aware bands
                    ms, share = rt.last_ms, getattr(rt, "root_share", None)
                    rep_ms = None
                    if n > 1 and chain == "kernel_cu":   # same step, result on every GPU through one NVLS multicast store
                        try:
                            rt.run(img, chain, mode="root", iters=20, replicate=True)
                            rep_ms = (rt.last_ms, bool(rt.nvls))
                        except Exception as e:
                            rep_ms = (None, str(e)[:80])
                # roofline of the band plan actually used: the root's share at HBM speed, the rest through the
                # root's NVLink port (input one way, output the other way, concurrently)
                f_root = 1.0 if n == 1 else (share if share else 1.0 / n)
                link_ms = (1 - f_root) * W * H * max(3, bpp - 3) / (NVL * 1e6)
                roof_ms = max(f_root * W * H * bpp / (HBM * 1e6), link_ms)
                row[f"n{n}"] = {"ms": round(ms, 4), "mpix_per_s": round(mp / ms * 1e3, 1),
                                "equal_bands_ms": round(ms_equal, 4), "root_share": None if share is None else round(share, 3),
                                "roofline_ms": round(roof_ms, 4), "of_roofline": round(roof_ms / ms, 3)}
                if rep_ms:
                    row[f"n{n}"]["replicated_output_equal_bands"] = {"ms": rep_ms[0] and round(rep_ms[0], 4), "nvls_multicast": rep_ms[1]}
            c2[chain] = row
        res["configs"]["2_8k_root_scatter_filter_gather"] = c2

    # 3. 8K RGB 5x5 Gaussian, bands resident, halo push fused into the kernel ---------------------------
    if 3 in only:
        c3 = {"naive_1gpu_ms": round(naive_ms("gauss5", img), 4)}
        for n in Ns:
            with Runtime(gpus=n) as rt:
                rt.run(img, "gauss5", mode="resident", passes=4)
                rt.run(img, "gauss5", mode="resident", passes=16)
                ms_e = rt.last_ms
                rt.run(img, "gauss5", mode="resident", passes=16, sync="flags")
                ms_f = rt.last_ms
            roof_ms = W * H * 6 / n / (HBM * 1e6)
            c3[f"n{n}"] = {"ms_per_pass_events": round(ms_e, 4), "ms_per_pass_flags": round(ms_f, 4),
                           "mpix_per_s": round(mp / min(ms_e, ms_f) * 1e3, 1), "roofline_ms": round(roof_ms, 4),
                           "of_roofline": round(roof_ms / min(ms_e, ms_f), 3), "halo_bytes_per_pass": 2 * 2 * W * 3}
        res["configs"]["3_8k_gauss5_resident_halo"] = c3

    del img

    # 4. 16K x 16K RGB blur -> Sobel, resident ----------------------------------------------------------
    if 4 in only and not a.quick:
        W4 = H4 = 16384
        img4 = b2.synth_image(W4, H4, 3, seed=2)
        mp4 = W4 * H4 / 1e6
        c4 = {}
        for n in Ns:
            with Runtime(gpus=n) as rt:
                rt.run(img4, "gauss5,sobel", mode="resident", passes=2)
                rt.run(img4, "gauss5,sobel", mode="resident", passes=6)
                ms = rt.last_ms
            roof_ms = W4 * H4 * 6 / n / (HBM * 1e6)
            c4[f"n{n}"] = {"ms_per_pass": round(ms, 4), "mpix_per_s": round(mp4 / ms * 1e3, 1),
                           "roofline_ms": round(roof_ms, 4), "of_roofline": round(roof_ms / ms, 3)}
        res["configs"]["4_16k_gauss5_sobel_resident"] = c4
        del img4

    # 5. batch of 4K frames from pinned host memory through grayscale + blur ---------------------------
    if 5 in only:
        Wf, Hf, F = 3840, 2160, a.frames if not a.quick else 32
        base = torch.from_numpy(b2.synth_image(Wf, Hf, 3, seed=3))
        NPOOL = 16  # distinct pinned frames cycled through (F x 24.9 MB would be 6.4 GB of pinned memory)
        fin = [base.clone().pin_memory() for _ in range(NPOOL)]
        fout = [torch.empty((Hf, Wf, 1), dtype=torch.uint8).pin_memory() for _ in range(NPOOL)]
        c5 = {}
        for n in Ns:
            bs = BatchStreamer("gray,gauss5", gpus=n)
            ins = [fin[i % NPOOL] for i in range(F)]
            outs = [fout[i % NPOOL] for i in range(F)]
            bs.run(ins[:2 * n], outs[:2 * n])
            for g in range(n):
                torch.cuda.synchronize(g)
            t0 = time.perf_counter()
            bs.run(ins, outs)
            for g in range(n):
                torch.cuda.synchronize(g)
            dt = time.perf_counter() - t0
            c5[f"n{n}"] = {"frames": F, "ms_per_frame": round(dt / F * 1e3, 4), "fps": round(F / dt, 1),
                           "mpix_per_s": round(F * Wf * Hf / 1e6 / dt, 1),
                           "h2d_gbs": round(F * Wf * Hf * 3 / dt / 1e9, 1), "d2h_gbs": round(F * Wf * Hf / dt / 1e9, 1)}
        want = b2.host_run("gray,gauss5", base.numpy(), rows=(100, 132))
        c5["bit_exact_rows_100_132"] = bool(np.array_equal(fout[0].numpy()[100:132], want))
        # the reference's pattern: pageable synchronous copies, one frame at a time (kernel.cu:163,202)
        p = Plan("gray,gauss5", 3)
        pag = base.clone()
        t0 = time.perf_counter()
        for _ in range(16):
            d = pag.cuda()
            o = p.run(d)
            _ = o.cpu()
        dt = (time.perf_counter() - t0) / 16
        c5["pageable_sync_1gpu"] = {"ms_per_frame": round(dt * 1e3, 3), "mpix_per_s": round(Wf * Hf / 1e6 / dt, 1)}
        res["configs"]["5_batch_4k_gray_blur_stream"] = c5

    txt = json.dumps(res, indent=1)
    print(txt)
    if a.out:
        open(a.out, "w").write(txt)


if __name__ == "__main__":
    main()
