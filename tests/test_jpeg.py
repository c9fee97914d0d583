"""JPEG in and out (nvJPEG on the GPU): the reference's own file format (kernel.cu:110,236; kern.cpp:33,92).

JPEG is lossy, so files are compared by PSNR against what cv2 (the OpenCV the reference calls)
decodes, not bit for bit."""
import os
import subprocess

import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2

cv2 = pytest.importorskip("cv2")
torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _psnr(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def _smooth(W, H, C=3):
    """A smooth test card (the hashed noise of synth_image is the worst case for any JPEG codec)."""
    y, x = np.mgrid[0:H, 0:W]
    img = np.stack([(x * 255 // max(W - 1, 1)), (y * 255 // max(H - 1, 1)), ((x + y) * 255 // max(W + H - 2, 1))], -1)
    img = img.astype(np.uint8)
    cv2.circle(img, (W // 2, H // 2), min(W, H) // 4, (200, 60, 30), -1)
    return img if C == 3 else img[:, :, 1].copy()


def test_reads_what_cv2_wrote(tmp_path):
    if not b2.lib.b2i_jpeg_available():
        pytest.skip("nvJPEG not on this machine")
    img = _smooth(641, 357)
    p = str(tmp_path / "in.jpg")
    assert cv2.imwrite(p, img, [cv2.IMWRITE_JPEG_QUALITY, 95])
    want = cv2.imread(p, cv2.IMREAD_UNCHANGED)          # what the reference would see (kernel.cu:110)
    got = b2.read_image(p, order="bgr")
    assert got.shape == want.shape
    assert _psnr(got, want) > 40, _psnr(got, want)       # two decoders, same file: IDCT rounding only
    rgb = b2.read_image(p, order="rgb")
    assert _psnr(rgb[:, :, ::-1], want) > 40


def test_gray_jpeg_reads_as_one_channel(tmp_path):
    if not b2.lib.b2i_jpeg_available():
        pytest.skip("nvJPEG not on this machine")
    g = _smooth(320, 200, 1)
    p = str(tmp_path / "g.jpg")
    assert cv2.imwrite(p, g)
    got = b2.read_image(p)
    assert got.shape == (200, 320, 1) and _psnr(got[:, :, 0], cv2.imread(p, cv2.IMREAD_UNCHANGED)) > 40


def test_cv2_reads_what_we_write(tmp_path):
    if not b2.lib.b2i_jpeg_available():
        pytest.skip("nvJPEG not on this machine")
    img = _smooth(800, 450)
    p = str(tmp_path / "out.jpg")
    b2.write_image(p, img, order="bgr", quality=95)
    back = cv2.imread(p, cv2.IMREAD_UNCHANGED)
    assert back.shape == img.shape and _psnr(back, img) > 35, _psnr(back, img)


def test_cli_jpeg_in_jpeg_out(tmp_path):
    """bin/b200img --input x.jpg --output y.jpg: the reference's own flow, file formats included."""
    if not b2.lib.b2i_jpeg_available():
        pytest.skip("nvJPEG not on this machine")
    img = _smooth(1280, 720)
    pin, pout = str(tmp_path / "3.jpg"), str(tmp_path / "imageFinalEmboss2.jpg")
    assert cv2.imwrite(pin, img, [cv2.IMWRITE_JPEG_QUALITY, 97])
    r = subprocess.run([os.path.join(ROOT, "bin", "b200img"), "--input", pin, "--output", pout, "--preset", "kernel_cu"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    want = b2.host_run("kernel_cu", b2.read_image(pin))   # the pipeline on the pixels our decoder delivers
    got = cv2.imread(pout, cv2.IMREAD_UNCHANGED)
    assert got.shape == want.shape and _psnr(got, want) > 30, _psnr(got, want)
