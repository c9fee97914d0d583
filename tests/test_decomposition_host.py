"""Row-band decomposition on the host: planner, halo-exact bands, host "ranks" backend."""
import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2

CHAINS = ["kernel_cu", "kern_cpp", "gauss5,sobel", "gray,gauss5@ref,emboss5@reflect101,sobel@replicate",
          "invert", "gauss5@zero,gauss5@zero"]


def test_band_rows_cover_everything_without_dropping_remainders():
    for H in (0, 1, 7, 8, 100, 4320, 16384):
        for n in (1, 2, 3, 4, 8):
            rows = [b2.band_rows(H, n, i) for i in range(n)]
            assert rows[0][0] == 0 and rows[-1][1] == H
            for (a0, a1), (b0, b1) in zip(rows, rows[1:]):
                assert a1 == b0 and a0 <= a1
            sizes = [b - a for a, b in rows]
            assert max(sizes) - min(sizes) <= 1          # kernel.cu:117 would drop H % n rows


@pytest.mark.parametrize("chain", CHAINS)
@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("W,H", [(40, 37), (33, 8), (20, 5)])   # H % n != 0, H < n*halo
def test_rows_api_equals_full_image(chain, n, W, H):
    img = b2.synth_image(W, H, seed=W * H + n)
    full = b2.host_run(chain, img)
    for i in range(n):
        y0, y1 = b2.band_rows(H, n, i)
        band = b2.host_run(chain, img, rows=(y0, y1))
        assert np.array_equal(band, full[y0:y1]), (chain, n, i)


@pytest.mark.parametrize("chain", CHAINS)
@pytest.mark.parametrize("ranks", [1, 2, 3, 4, 8])
def test_host_ranks_equal_single_worker(chain, ranks):
    img = b2.synth_image(61, 43, seed=7)
    assert np.array_equal(b2.host_run(chain, img, ranks=ranks), b2.host_run(chain, img))


def test_baseline_config1_512_gray_two_ranks():
    # BASELINE.json config 1: 512x512 RGB, pointwise grayscale, 2 host ranks, no GPU
    img = b2.synth_image(512, 512, seed=1)
    out = b2.host_run("gray", img, ranks=2)
    b, g, r = (img[..., i].astype(np.int32) for i in range(3))
    assert np.array_equal(out[:, :, 0], (b * 11 // 100 + g * 59 // 100 + r * 3 // 10).astype(np.uint8))


def test_seams_mode_reproduces_the_reference_decomposition_artefact():
    # the reference filters each band in isolation (no halo exchange): rows next to a band
    # boundary differ from the seam-free result, interior rows do not
    img = b2.synth_image(64, 64, seed=3)
    clean = b2.host_run("kernel_cu", img)
    seamy = b2.host_run("kernel_cu", img, ranks=2, seams=True)
    assert not np.array_equal(clean[30:35], seamy[30:35])
    assert np.array_equal(clean[3:30], seamy[3:30]) and np.array_equal(clean[35:62], seamy[35:62])
