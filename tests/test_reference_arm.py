"""The reference arm of bench.py must stay clear of this repo's native library (the driver voids
the comparison if the arm maps libb200img.so), and its NumPy image must be the image our arm uses."""
import json
import os
import subprocess
import sys

import numpy as np

import mpi_cuda_imagemanipulation_b200 as b2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_numpy_synth_equals_native():
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import synth_np
    for W, H, seed in [(64, 48, 1), (1001, 77, 7), (640, 360, 1), (1, 1, 0)]:
        assert np.array_equal(synth_np.synth_image(W, H, 3, seed), b2.synth_image(W, H, 3, seed=seed))


def test_numpy_ppm_equals_native(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import synth_np
    img = b2.synth_image(100, 30, 3, seed=1)
    synth_np.write_ppm_from_bgr(str(tmp_path / "a.ppm"), img)
    b2.write_image(str(tmp_path / "b.ppm"), img, order="bgr")
    assert (tmp_path / "a.ppm").read_bytes() == (tmp_path / "b.ppm").read_bytes()


def test_reference_arm_maps_no_native_library():
    """Imports everything the reference arm imports, then looks at /proc/self/maps."""
    code = (
        "import sys, os\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "import bench\n"
        "import baseline.run_reference as rr\n"
        "from bench_common import clock_sampler_cls, config_block\n"
        "clock_sampler_cls()\n"
        "import synth_np\n"
        "synth_np.synth_image(32, 16, 3, 1)\n"
        "maps = open('/proc/self/maps').read()\n"
        "print('MAPPED' if 'libb200img' in maps else 'CLEAN')\n"
        "print('PKG' if any(m.startswith('mpi_cuda_imagemanipulation_b200') for m in sys.modules) else 'NOPKG')\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    assert r.stdout.split() == ["CLEAN", "NOPKG"], r.stdout


def test_config1_reference_cpu_program_runs(tmp_path):
    """BASELINE.json config 1 on the reference's own CPU program: needs no GPU."""
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import run_reference as rr
    try:
        rr.ensure_built("ref_kern_cpp")
    except RuntimeError:
        import pytest
        pytest.skip("reference sources not available on this machine")
    res = rr.bench_config1(str(tmp_path))
    assert res["window_ms"] > 0 and res["mpix_per_s"] > 0
    json.dumps(res)
