"""Real multi-GPU, one process per GPU (torchrun): needs >= 2 GPUs, skipped otherwise.

Exercises what a one-GPU box cannot: IPC-mapped peer memory under TMA loads and stores, device
flags acquired across GPUs, NVLink halo pushes.
"""
import os
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n", [2, 4, 8])
def test_torchrun_ranks_match_host_reference(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29500 + n),
                        os.path.join(ROOT, "scripts", "dist_check.py")], capture_output=True, text=True, timeout=600)
    assert "DIST_CHECK PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("n", [2, 4])
def test_lost_peer_is_a_timeout_error_not_a_hang(n):
    """One rank never joins a resident run: every rank gets B2I_ERR_TIMEOUT within the (50 ms) timeout
    and the runtime works again afterwards (scripts/dist_fault.py)."""
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29520 + n),
                        os.path.join(ROOT, "scripts", "dist_fault.py")], capture_output=True, text=True, timeout=600)
    assert "DIST_FAULT PASS" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_in_process_runtime_on_real_peers():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import numpy as np
    import mpi_cuda_imagemanipulation_b200 as b2
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime
    n = min(torch.cuda.device_count(), 8)
    img = b2.synth_image(2048, 1030, seed=3)
    with Runtime(gpus=n) as rt:
        assert np.array_equal(rt.run(img, "kernel_cu", mode="root"), b2.host_run("kernel_cu", img))
        want = b2.host_run("gauss5", b2.host_run("gauss5", b2.host_run("gauss5", img)))
        assert np.array_equal(rt.run(img, "gauss5", mode="resident", passes=3), want)
        assert np.array_equal(rt.run(img, "gauss5", mode="resident", passes=3, sync="flags"), want)
        assert np.array_equal(rt.run(img, "kernel_cu", mode="direct"), b2.host_run("kernel_cu", img))


def test_nvls_multicast_replicated_output():
    """Every GPU ends up with the whole result through ONE store per byte (NVLS multicast)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import numpy as np
    import mpi_cuda_imagemanipulation_b200 as b2
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime, topology
    n = min(torch.cuda.device_count(), 8)
    if not all(d.get("multicast") == 1 for d in topology()["devices"][:n]):
        pytest.skip("no multicast support on this box")
    img = b2.synth_image(2048, 1030, seed=6)
    with Runtime(gpus=n) as rt:
        for chain in ("kernel_cu", "gauss5,sobel", "gray"):
            want = b2.host_run(chain, img)
            got = rt.run(img, chain, mode="root", replicate=True)
            assert rt.nvls, "expected a multicast object on real peers"
            assert np.array_equal(got, want), chain
            for g, r in enumerate(rt.replicas):
                assert np.array_equal(r, want), (chain, g)
