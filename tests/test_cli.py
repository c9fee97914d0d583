"""The command-line entry point (SURVEY §4 'CLI'): zero-argument run, flags, error path."""
import json
import os
import subprocess

import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "bin", "b200img")


def _exe():
    if not os.path.exists(EXE):
        subprocess.run(["make", "-C", ROOT, "-j8"], capture_output=True)
    if not os.path.exists(EXE):
        pytest.skip("bin/b200img not built")
    return EXE


def test_no_arguments_reproduces_the_reference_run(tmp_path):
    # kernel.cu takes no arguments: it filters a hard-coded image and writes imageFinalEmboss2.*
    img = b2.synth_image(320, 200, 3, seed=8)
    b2.write_image(str(tmp_path / "3.ppm"), img)            # the reference reads ...\3.jpg
    r = subprocess.run([_exe(), "--backend", "host"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Process #0 received the gathered image" in r.stdout
    assert "ns to run modification" in r.stdout and "ms to run modification" in r.stdout
    out = b2.read_image(str(tmp_path / "imageFinalEmboss2.ppm"))
    assert np.array_equal(out, b2.host_run("kernel_cu", img))


def test_np_flag_filters_presets_and_json(tmp_path):
    img = b2.synth_image(512, 512, 3, seed=1)
    b2.write_image(str(tmp_path / "in.ppm"), img)
    r = subprocess.run([_exe(), "--input", "in.ppm", "--output", "g.pgm", "--filters", "gray", "-np", "2",
                        "--backend", "host", "--json"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    rep = json.loads(r.stdout)
    assert rep["workers"] == 2 and rep["W"] == 512 and rep["chain"] == "gray_ref"
    assert np.array_equal(b2.read_image(str(tmp_path / "g.pgm")), b2.host_run("gray", img))
    r = subprocess.run([_exe(), "--input", "in.ppm", "--output", "k.ppm", "--preset", "kern_cpp", "--ranks", "3",
                        "--backend", "host"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert np.array_equal(b2.read_image(str(tmp_path / "k.ppm")), b2.host_run("kern_cpp", img))
    # --seams reproduces the reference's missing halo exchange
    r = subprocess.run([_exe(), "--input", "in.ppm", "--output", "s.ppm", "--ranks", "2", "--seams",
                        "--backend", "host"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0
    assert np.array_equal(b2.read_image(str(tmp_path / "s.ppm")), b2.host_run("kernel_cu", img, ranks=2, seams=True))


def test_raw_io_border_and_rgb_order(tmp_path):
    img = b2.synth_image(100, 60, 3, seed=2)
    b2.write_raw(str(tmp_path / "in.raw"), img)
    r = subprocess.run([_exe(), "--input", "in.raw", "--width", "100", "--height", "60", "--output", "o.raw",
                        "--filters", "gray,gauss5,sobel", "--border", "replicate", "--order", "rgb",
                        "--backend", "host"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    want = b2.host_run(b2.Chain("gray,gauss5,sobel", border="replicate", order="rgb"), img)
    assert np.array_equal(b2.read_raw(str(tmp_path / "o.raw"), 100, 60, 1), want)


def test_missing_input_fails_like_the_reference_but_cleanly(tmp_path):
    r = subprocess.run([_exe(), "--input", "nope.ppm", "--backend", "host"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 1 and "Failed imread(): image not found" in r.stderr   # kernel.cu:112
    r = subprocess.run([_exe(), "--filters", "nosuch", "--backend", "host"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 1 and "unknown filter" in r.stderr
    assert subprocess.run([_exe(), "--help"], capture_output=True, text=True).returncode == 0


@pytest.mark.gpu
def test_cuda_backend_single_and_multi_band(tmp_path):
    img = b2.synth_image(1024, 403, 3, seed=5)
    b2.write_image(str(tmp_path / "in.ppm"), img)
    want = b2.host_run("kernel_cu", img)
    for extra in ([], ["-np", "3"], ["-np", "4", "--resident", "--iters", "1"]):
        r = subprocess.run([_exe(), "--input", "in.ppm", "--output", "o.ppm", "--backend", "cuda", *extra],
                           cwd=tmp_path, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr + r.stdout
        assert np.array_equal(b2.read_image(str(tmp_path / "o.ppm")), want), extra
    r = subprocess.run([_exe(), "--input", "in.ppm", "--output", "o2.ppm", "--backend", "cuda", "-np", "2",
                        "--replicate", "--json"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    line = json.loads(r.stdout)
    assert line["hbm_bytes"] == 1024 * 403 * 6 and line["hbm_roofline_ms"] > 0 and line["nvlink_bytes_equal_bands"] > 0
    assert 0 <= line["of_hbm_roofline"] < 1.5      # a 0.4 MP image is launch bound: the fraction rounds to 0
    rep = line["runtime"]
    assert rep["mode"].startswith("root/replicated") and rep["replicas_identical"] == "yes", rep
    assert np.array_equal(b2.read_image(str(tmp_path / "o2.ppm")), want)
    r = subprocess.run([_exe(), "--topology"], capture_output=True, text=True)
    assert r.returncode == 0 and len(json.loads(r.stdout)["devices"]) >= 1


def test_batch_directory(tmp_path):
    (tmp_path / "in").mkdir()
    (tmp_path / "out").mkdir()
    imgs = {}
    for i in range(3):
        imgs[f"f{i}.ppm"] = b2.synth_image(64 + 16 * i, 40, 3, seed=i)
        b2.write_image(str(tmp_path / "in" / f"f{i}.ppm"), imgs[f"f{i}.ppm"])
    r = subprocess.run([_exe(), "--batch", "in", "--outdir", "out", "--filters", "gray,gauss5", "--backend", "host"],
                       cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert json.loads(r.stdout)["frames"] == 3
    for name, img in imgs.items():
        got = b2.read_image(str(tmp_path / "out" / name.replace(".ppm", ".pgm")))
        assert np.array_equal(got, b2.host_run("gray,gauss5", img))
