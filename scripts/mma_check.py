"""Correctness probe of the tensor-core stencil kernel against the host reference: mismatch counts
and where they are, for a list of chains / sizes (run under gpurun; every case in a subprocess with a
timeout so that a hang cannot take the box down).
    python scripts/mma_check.py [--dbg=N]"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [("gauss5", 3, 512, 256), ("gauss5", 3, 1024, 300), ("gauss5", 1, 640, 200), ("gauss3", 3, 512, 130),
         ("sobel", 3, 512, 256), ("gauss5@replicate", 3, 512, 140), ("gauss5@zero", 3, 512, 140),
         ("sharpen", 3, 768, 260), ("filter5:1/2/3/4/5/-1/-2/-3/-4/-5/0/1/0/-1/0/2/2/2/2/2/1/1/1/1/-30/0@reflect101", 3, 512, 129),
         ("gauss5", 3, 7680, 4320), ("sobel", 3, 7680, 4320), ("gauss5", 3, 1000, 77), ("gauss5", 3, 1001, 333)]


def one(chain, cn, W, H):
    import numpy as np, torch
    import mpi_cuda_imagemanipulation_b200 as b2
    from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
    img = b2.synth_image(W, H, cn, seed=3)
    want = b2.host_run(chain, img)
    plan = Plan(chain, cn)
    got = plan.run(torch.from_numpy(img).cuda())
    torch.cuda.synchronize()
    got = got.cpu().numpy().reshape(want.shape)
    bad = got != want
    info = plan.info()
    msg = f"{chain[:28]:28s} c{cn} {W}x{H}: tile_h={info['tile_h']} grid={info['grid']} mismatches={int(bad.sum())}/{bad.size}"
    if bad.any():
        ys, xs = np.nonzero(bad.reshape(H, -1))
        msg += f" rows[{ys.min()},{ys.max()}] cols[{xs.min()},{xs.max()}]"
        y, x = ys[0], xs[0]
        msg += f" first(y={y},x={x}) got={got.reshape(H, -1)[y, x]} want={want.reshape(H, -1)[y, x]}"
        rows = np.unique(ys)
        cols = np.unique(xs)
        msg += f" nrows={len(rows)} ncols={len(cols)} cols%128={sorted(set((cols % 128).tolist()))[:12]}"
    print(msg, flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        one(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]))
        sys.exit(0)
    for c in CASES:
        try:
            r = subprocess.run([sys.executable, __file__, "--one", c[0], str(c[1]), str(c[2]), str(c[3])],
                               capture_output=True, text=True, timeout=120)
            print((r.stdout.strip() or "no output") + (f"  [rc={r.returncode}] {r.stderr.strip()[-300:]}" if r.returncode else ""), flush=True)
        except subprocess.TimeoutExpired:
            print(f"{c}: TIMEOUT", flush=True)
