"""The unmodified reference's CPU program (kern.cpp), built against the shim MPI/OpenCV headers,
must compute what the kern_cpp preset computes (which test_host_numerics.py pins to real OpenCV).
This validates the shims the baseline arm relies on. Runs on the CPU container."""
import os
import subprocess
import sys

import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "baseline", "_ref", "ref_kern_cpp")


def _build():
    if not os.path.exists(EXE):
        subprocess.run(["bash", os.path.join(ROOT, "baseline", "build_ref.sh")], capture_output=True)
    if not os.path.exists(EXE):
        pytest.skip("reference sources not available on this machine")


@pytest.mark.parametrize("np_", [1, 2, 4])
def test_shim_built_kern_cpp_matches_preset(tmp_path, np_):
    _build()
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    from mini_mpirun import launch
    W, H = 96, 64
    img = b2.synth_image(W, H, 3, seed=4)
    inp = str(tmp_path / "in.ppm")
    b2.write_image(inp, img, order="bgr")
    env = dict(os.environ, B2I_REF_INPUT=inp, B2I_REF_OUTDIR=str(tmp_path))
    outs = launch(np_, EXE, env=env, shm_bytes=1 << 20, timeout=60)
    assert all(rc == 0 for rc, _, _ in outs), outs
    assert "Process #0 received the gathered image" in outs[0][1]
    got = b2.read_image(str(tmp_path / "imageFinalEmboss2.jpg.ppm"), order="bgr")
    # the reference filters every band in isolation (no halo exchange): that is seams=True
    want = b2.host_run("kern_cpp", img, ranks=np_, seams=True)
    assert np.array_equal(got, want)
    if np_ > 1:   # and that differs from the seam-free result right at the band boundaries
        clean = b2.host_run("kern_cpp", img)
        assert not np.array_equal(got, clean)
        assert np.array_equal(got[2:H // np_ - 2], clean[2:H // np_ - 2])
