"""C++ host reference (csrc/host_ref) against independent NumPy / OpenCV oracles."""
import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2
from tests import oracles as O

RNG = np.random.default_rng(1234)
SIZES = [(64, 48), (37, 29), (16, 16), (5, 7), (130, 9), (1, 1), (3, 2), (2, 40)]


def rand_img(W, H, C=3):
    a = RNG.integers(0, 256, (H, W, C), dtype=np.uint8)
    return a if C == 3 else a[:, :, 0]


def test_gray_ref_all_values_every_channel():
    # all 256 values x 3 channels: ramp in one channel, zeros elsewhere
    for ch in range(3):
        img = np.zeros((1, 256, 3), np.uint8)
        img[0, :, ch] = np.arange(256)
        got = b2.host_run("gray", img)[:, :, 0]
        assert np.array_equal(got, O.gray_ref(img))
    white = np.full((2, 2, 3), 255, np.uint8)
    assert int(b2.host_run("gray", white)[0, 0, 0]) == 254  # 28+150+76, never 255


def test_gray_ref_random_and_rgb_order():
    img = rand_img(97, 31)
    assert np.array_equal(b2.host_run("gray", img)[:, :, 0], O.gray_ref(img))
    assert np.array_equal(b2.host_run("gray", img, order="rgb")[:, :, 0], O.gray_ref(img, "rgb"))


def test_gray_cv_matches_opencv():
    img = rand_img(123, 45)
    assert np.array_equal(b2.host_run("gray_cv", img)[:, :, 0], O.gray_cv(img))
    assert np.array_equal(b2.host_run("gray_cv", img, order="rgb")[:, :, 0], O.gray_cv(img, "rgb"))


@pytest.mark.parametrize("factor", [3.5, 1.0, 0.5, 2.25, 1.7, 10.0])
def test_contrast_ref_lut(factor):
    ramp = np.arange(256, dtype=np.uint8)
    assert np.array_equal(b2.contrast_lut(factor, "ref"), O.contrast_ref(ramp, factor))


def test_contrast_ref_spot_values():
    lut = b2.contrast_lut(3.5)
    for v, want in [(0, 0), (91, 0), (92, 2), (127, 124), (128, 128), (129, 131), (164, 254),
                    (165, 255), (255, 255)]:
        assert lut[v] == want, (v, lut[v], want)


def test_contrast_cv_lut_matches_opencv():
    import cv2
    ramp = np.arange(256, dtype=np.uint8).reshape(1, -1)
    want = cv2.addWeighted(ramp, 3.0, ramp, 0, -256.0)[0]
    assert np.array_equal(b2.contrast_lut(3, "cv"), want)
    assert b2.contrast_lut(3, "cv")[85] == 0 and b2.contrast_lut(3, "cv")[86] == 2
    assert b2.contrast_lut(3, "cv")[171] == 255


@pytest.mark.parametrize("W,H", SIZES)
@pytest.mark.parametrize("border", ["ref", "reflect101", "replicate", "zero"])
@pytest.mark.parametrize("name,K,shift,bias", [("emboss3", O.EMBOSS3, 0, 0),
                                               ("emboss5", O.EMBOSS5, 0, 0),
                                               ("gauss5", O.GAUSS5, 8, 128),
                                               ("gauss3", O.GAUSS3, 4, 8)])
def test_stencils_gray(W, H, border, name, K, shift, bias):
    if border in ("reflect101",) and (W <= K.shape[0] // 2 or H <= K.shape[0] // 2):
        pytest.skip("numpy reflect padding needs size > radius")
    g = rand_img(W, H, 1)
    got = b2.host_run(f"{name}@{border}", g)
    assert np.array_equal(got, O.stencil(g, K, border, shift, bias))


@pytest.mark.parametrize("border", ["ref", "reflect101", "replicate", "zero"])
def test_stencils_three_channel(border):
    img = rand_img(45, 33)
    assert np.array_equal(b2.host_run(f"gauss5@{border}", img), O.stencil(img, O.GAUSS5, border, 8, 128))
    assert np.array_equal(b2.host_run(f"sobel@{border}", img), O.sobel(img, border))
    assert np.array_equal(b2.host_run(f"emboss3@{border}", img), O.stencil(img, O.EMBOSS3, border))


def test_cv2_crosschecks():
    import cv2
    img = rand_img(80, 64)
    g = img[:, :, 1].copy()
    assert np.array_equal(b2.host_run("gauss5", img), cv2.GaussianBlur(img, (5, 5), 0))
    assert np.array_equal(b2.host_run("gauss3", g), cv2.GaussianBlur(g, (3, 3), 0))
    assert np.array_equal(b2.host_run("emboss3@reflect101", g),
                          cv2.filter2D(g, -1, O.EMBOSS3.astype(np.int8)))
    gx = cv2.Sobel(g, cv2.CV_16S, 1, 0, ksize=3).astype(np.int32)
    gy = cv2.Sobel(g, cv2.CV_16S, 0, 1, ksize=3).astype(np.int32)
    assert np.array_equal(b2.host_run("sobel", g),
                          np.clip(np.abs(gx) + np.abs(gy), 0, 255).astype(np.uint8))


def test_generic_filter_token():
    g = rand_img(40, 30, 1)
    K = np.array([[1, 2, 1], [2, 4, 2], [1, 2, 1]], np.int64)
    got = b2.host_run("filter3:1/2/1/2/4/2/1/2/1/4", g)
    assert np.array_equal(got, O.stencil(g, K, "reflect101", 4, 8))


def test_lut_ops():
    g = rand_img(33, 21, 1)
    assert np.array_equal(b2.host_run("invert", g), 255 - g)
    assert np.array_equal(b2.host_run("threshold:100", g), np.where(g > 100, 255, 0).astype(np.uint8))
    img = rand_img(33, 21)
    assert np.array_equal(b2.host_run("invert", img), 255 - img)
    assert np.array_equal(b2.host_run("brightness:-40", g),
                          np.clip(g.astype(np.int32) - 40, 0, 255).astype(np.uint8))


@pytest.mark.parametrize("W,H", [(64, 48), (33, 17), (7, 9), (256, 8)])
def test_preset_kernel_cu(W, H):
    img = rand_img(W, H)
    assert np.array_equal(b2.host_run("kernel_cu", img), O.pipeline_kernel_cu(img))


@pytest.mark.parametrize("W,H", [(64, 48), (33, 17), (100, 75)])
def test_preset_kern_cpp_matches_real_opencv(W, H):
    img = rand_img(W, H)
    assert np.array_equal(b2.host_run("kern_cpp", img), O.pipeline_kern_cpp(img))


def test_presets_differ_like_the_reference_programs():
    # the two reference programs are different algorithms
    img = b2.synth_image(64, 48)
    a, c = b2.host_run("kernel_cu", img), b2.host_run("kern_cpp", img)
    assert a.shape == c.shape and not np.array_equal(a, c)


def test_chain_validation():
    with pytest.raises(b2.B200ImgError):
        b2.Chain("nosuchfilter")
    with pytest.raises(b2.B200ImgError):
        b2.host_run("gray", rand_img(8, 8, 1))      # gray needs 3 channels
    with pytest.raises(b2.B200ImgError):
        b2.host_run("gray,gray", rand_img(8, 8))
    with pytest.raises(b2.B200ImgError):
        b2.Chain("emboss3@nowhere")
    c = b2.Chain("gauss5,sobel,emboss5")
    assert c.halo == 5 and c.n_ops == 3
    assert b2.Chain(preset="kernel_cu").describe() == "gray_ref,contrast,emboss3@ref,expand"
