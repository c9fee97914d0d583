"""CPU ranks over torch.distributed/gloo (BASELINE.json config 1), world_size 2 and 3, no GPU."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n,chain,W,H", [(2, "gray", 512, 512), (3, "kernel_cu", 200, 101), (2, "gauss5,sobel", 97, 40)])
def test_gloo_ranks_match_single_process(n, chain, W, H):
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29600 + n + len(chain)),
                        "-m", "mpi_cuda_imagemanipulation_b200.parallel.host_dist", "--chain", chain,
                        "--width", str(W), "--height", str(H)], capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    line = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert line, r.stdout[-2000:] + r.stderr[-2000:]
    rep = json.loads(line[-1])
    assert rep["bit_exact"] and rep["ranks"] == n
