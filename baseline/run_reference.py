"""The reference arm of bench.py: the UNMODIFIED reference program, built against shim MPI/OpenCV
headers (baseline/ref_shim), run through its own entry point (main(), `mpirun -np N` style) on the
same synthetic 8K image. This module loads none of this repo's native code: the image comes from
NumPy (baseline/synth_np.py) and the reference runs as subprocesses of its own binary.

The reference handles one image per process launch, so one step = one launch of N ranks. What the
numbers mean (every one is a host clock: the reference has no device timing):

  value     MP/s from the reference's own "______took N ns to run modification" line
            (kernel.cu:190,226-231): three kernel launches + cudaDeviceSynchronize + pageable D2H +
            host GRAY2BGR + MPI_Gather; it excludes decode, MPI_Scatter, cudaMalloc and H2D. Taken
            WARM: the shim creates the CUDA context inside MPI_Init (B2I_REF_PREWARM=1) and the
            launcher sets CUDA_MODULE_LOADING=EAGER, so neither driver start-up nor the lazy load
            of the three kernels sits in the window. `cold.*` has the as-shipped numbers.
  e2e       MP/s over MPI_Scatter entry -> MPI_Gather exit on rank 0, measured inside the shim:
            scatter, cudaMalloc, pageable H2D, kernels, D2H, host expand, gather. Warm as above;
            `ctx_init_ms` is the context creation that the warm window leaves out.
  kernel_ms device time of grayscaleKernel + contrastKernel + embossKernel from ncu on this binary
            (profiles/ref_kernels_ncu.json, recorded once on a B200; ncu cannot run inside a bench).
  as_written  N > 1 only: all ranks on GPU 0, which is what kernel.cu:147 does.
  config1   BASELINE.json config 1: the reference's CPU program kern.cpp, -np 2, 512x512.
"""
import json
import os
import re
import statistics
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)


def ensure_built(name="ref_kernel_cu"):
    exe = os.path.join(HERE, "_ref", name)
    shims = [os.path.join(HERE, "ref_shim", "mpi.h"), os.path.join(HERE, "ref_shim", "opencv2", "opencv.hpp")]
    stale = os.path.exists(exe) and any(os.path.getmtime(h) > os.path.getmtime(exe) for h in shims)
    if not os.path.exists(exe) or stale:
        r = subprocess.run(["bash", os.path.join(HERE, "build_ref.sh")], capture_output=True, text=True)
        if r.returncode != 0 or not os.path.exists(exe):
            raise RuntimeError("reference build failed: " + (r.stdout + r.stderr)[-200:].replace("\n", " "))
    return exe


def run_once(exe, np_, inp, outdir, gpu_per_rank, shm_bytes, warm=True, timeout=600):
    """One launch of the reference on np_ ranks. Returns its printed window and the shim's clocks (ns)."""
    from mini_mpirun import launch
    env = dict(os.environ, B2I_REF_INPUT=inp, B2I_REF_OUTDIR=outdir)
    env.pop("B2I_REF_PREWARM", None)
    env.pop("CUDA_MODULE_LOADING", None)
    if warm:
        env.update(B2I_REF_PREWARM="1", CUDA_MODULE_LOADING="EAGER")
    if gpu_per_rank:
        env.pop("CUDA_VISIBLE_DEVICES", None)
    outs = launch(np_, exe, gpu_per_rank=gpu_per_rank, shm_bytes=shm_bytes, env=env, timeout=timeout)
    rc0, out0, err0 = outs[0]
    if any(rc for rc, _, _ in outs):
        raise RuntimeError(f"reference rank failed: rc={[rc for rc, _, _ in outs]} {err0[-200:]}")
    res = {}
    m = re.search(r"took (\d+)ns to run modification", out0)
    if m:
        res["window_ns"] = int(m.group(1))
    m = re.search(r"took (\d+)ms to run modification", out0)
    if m:
        res["window_ms_printed"] = int(m.group(1))
    m = re.search(r"scatter_to_gather_ns=(\d+)", err0)
    if m:
        res["s2g_ns"] = int(m.group(1))
    m = re.search(r"after_scatter_to_gather_ns=(\d+)", err0)
    if m:
        res["after_scatter_ns"] = int(m.group(1))
    ctx = [int(x) for _, _, e in outs for x in re.findall(r"ctx_init_ns=(\d+) where=" +
                                                           ("MPI_Init" if warm else "cudaSetDevice"), e)]
    if ctx:
        res["ctx_init_ns"] = max(ctx)
    return res


def _med(rs, key):
    v = [r[key] for r in rs if key in r]
    return statistics.median(v) if v else None


def _series(exe, n, inp, tmp, shm, gpu_per_rank, warm, warmup, steps):
    rs = []
    for i in range(warmup + steps):
        r = run_once(exe, n, inp, tmp, gpu_per_rank=gpu_per_rank, shm_bytes=shm, warm=warm)
        if i >= warmup:
            rs.append(r)
    return rs


def bench_config1(tmp):
    """BASELINE.json config 1 on the reference's own CPU program (kern.cpp:17-97), -np 2, 512x512."""
    from synth_np import synth_image, write_ppm_from_bgr
    exe = ensure_built("ref_kern_cpp")
    W = H = 512
    inp = os.path.join(tmp, "in512.ppm")
    write_ppm_from_bgr(inp, synth_image(W, H, 3, seed=1))
    rs = _series(exe, 2, inp, tmp, 8 << 20, gpu_per_rank=False, warm=False, warmup=2, steps=9)
    ns = _med(rs, "after_scatter_ns")
    return {"program": "kern.cpp (unmodified, shim MPI/OpenCV), -np 2, 512x512 RGB",
            "window_ms": round(ns / 1e6, 4), "mpix_per_s": round(W * H / 1e6 / (ns / 1e9), 1),
            "printed_ms": _med(rs, "window_ms_printed"),
            "window": "leaving MPI_Scatter -> leaving MPI_Gather on rank 0 (kern.cpp:60-86; the program itself prints "
                      "whole milliseconds only), median of 9 launches"}


def bench_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:   # under torchrun only rank 0 drives the reference's own N-rank launch
        return None
    from bench_common import clock_sampler_cls, config_block
    from synth_np import synth_image, write_ppm_from_bgr
    exe = ensure_built()
    n = max(1, args.gpus)
    W, H = args.width, args.height
    steps, warm = max(1, args.steps), max(0, args.warmup)
    mp_done = W * (H // n) * n / 1e6   # the reference drops H % N rows (kernel.cu:117)
    extras = {}
    with tempfile.TemporaryDirectory(prefix="b2i_ref_") as tmp:
        inp = os.path.join(tmp, "in.ppm")
        write_ppm_from_bgr(inp, synth_image(W, H, 3, seed=1))
        shm = max(256 << 20, 2 * W * H * 3 + (1 << 20))
        with clock_sampler_cls()(0) as clk:
            rs = _series(exe, n, inp, tmp, shm, gpu_per_rank=n > 1, warm=True, warmup=warm, steps=steps)
        try:
            cold = _series(exe, n, inp, tmp, shm, gpu_per_rank=n > 1, warm=False, warmup=0, steps=3)
            extras["cold"] = {"window_ms": round(_med(cold, "window_ns") / 1e6, 3),
                              "scatter_to_gather_ms": round(_med(cold, "s2g_ns") / 1e6, 3),
                              "ctx_init_ms": round((_med(cold, "ctx_init_ns") or 0) / 1e6, 3),
                              "what": "as shipped: context created by cudaSetDevice after MPI_Scatter (kernel.cu:137,147), "
                                      "kernels loaded lazily at first launch; median of 3 launches"}
        except Exception as e:
            extras["cold"] = {"error": str(e)[:160]}
        if n > 1:
            try:
                aw = _series(exe, n, inp, tmp, shm, gpu_per_rank=False, warm=True, warmup=1, steps=3)
                ns = _med(aw, "window_ns")
                extras["as_written"] = {"window_ms": round(ns / 1e6, 3), "mpix_per_s": round(mp_done / (ns / 1e9), 1),
                                        "scatter_to_gather_ms": round(_med(aw, "s2g_ns") / 1e6, 3),
                                        "what": f"all {n} ranks on GPU 0, as kernel.cu:147 is written; warm; median of 3"}
            except Exception as e:
                extras["as_written"] = {"error": str(e)[:160]}
        try:
            extras["config1_kern_cpp_np2_512"] = bench_config1(tmp)
        except Exception as e:
            extras["config1_kern_cpp_np2_512"] = {"error": str(e)[:160]}
    ns = _med(rs, "window_ns")
    if ns is None:
        raise RuntimeError("reference printed no timing line")
    out = {
        "impl": "reference", "metric": "megapixels/sec (device-timed, max over ranks)",
        "value": round(mp_done / (ns / 1e9), 1), "unit": "MP/s", "n_gpus": n, "steps": steps, "warmup": warm,
        "ms_per_step": round(ns / 1e6, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "uint8", "data": "synthetic",
        "config": config_block(W, H, n),
        "clocks": clk.summary(),
        "detail": {"build": "unmodified /root/reference/kernel.cu, nvcc -O3 sm_100a, shim MPI/OpenCV headers "
                            "(baseline/ref_shim); one process launch per step"
                            + (", one GPU per rank via CUDA_VISIBLE_DEVICES" if n > 1 else ""),
                   "timed_as": "the reference's own printed window (kernel.cu:190-231): kernels + sync + pageable D2H + "
                               "host GRAY2BGR + MPI_Gather; host clock (it has no device timing); warm: context "
                               "created in MPI_Init by the shim, CUDA_MODULE_LOADING=EAGER; median over steps"},
        "gpu_launches": 3 * n * steps,
        "extras": extras,
    }
    e2e = _med(rs, "s2g_ns")
    if e2e:
        out["e2e"] = {"value": round(mp_done / (e2e / 1e9), 1), "unit": "MP/s", "ms_per_step": round(e2e / 1e6, 3),
                      "h2d_bytes_per_step": W * (H // n) * n * 3, "d2h_bytes_per_step": W * (H // n) * n,
                      "ctx_init_ms": round((_med(rs, "ctx_init_ns") or 0) / 1e6, 3),
                      "how": "MPI_Scatter entry -> MPI_Gather exit on rank 0, measured in the shim: scatter, cudaMalloc, "
                             "pageable H2D, kernels, D2H, host expand, gather; warm (ctx_init_ms, spent in MPI_Init, "
                             "is outside the window)"}
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "ref_kernels_ncu.json")))
        k = rec.get(f"{W}x{H}") or {}
        if k:
            out["kernel_ms"] = k["total_ms"]
            out["detail"]["kernel_ms"] = ("sum of gpu__time_duration of the reference's three kernels on one B200, from ncu "
                                          f"on this binary, recorded {rec.get('when', '?')} (profiles/ref_kernels_ncu.json); "
                                          "not re-measured in this run")
    except Exception:
        pass
    return out


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--width", type=int, default=7680)
    ap.add_argument("--height", type=int, default=4320)
    print(json.dumps(bench_reference(ap.parse_args())))
