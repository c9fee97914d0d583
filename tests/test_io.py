"""PPM / PGM / raw round trips (SURVEY §4 'Unit (I/O)')."""
import os

import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2


@pytest.mark.parametrize("W,H", [(1, 1), (7, 5), (64, 48), (333, 17)])
def test_ppm_roundtrip_and_channel_order(tmp_path, W, H):
    img = b2.synth_image(W, H, seed=W + H)
    p = str(tmp_path / "a.ppm")
    b2.write_image(p, img, order="bgr")
    assert np.array_equal(b2.read_image(p, order="bgr"), img)
    # on disk PPM is RGB: reading it back as RGB yields the channel-swapped array
    assert np.array_equal(b2.read_image(p, order="rgb"), img[:, :, ::-1])
    raw = open(p, "rb").read()
    header = f"P6\n{W} {H}\n255\n".encode()
    assert raw.startswith(header)
    assert np.array_equal(np.frombuffer(raw[len(header):], np.uint8).reshape(H, W, 3), img[:, :, ::-1])


def test_pgm_roundtrip(tmp_path):
    g = b2.synth_image(31, 9, 1, seed=5)
    p = str(tmp_path / "g.pgm")
    b2.write_image(p, g)
    assert open(p, "rb").read(2) == b"P5"
    assert np.array_equal(b2.read_image(p), g)


def test_pnm_header_comments_and_errors(tmp_path):
    p = str(tmp_path / "c.ppm")
    with open(p, "wb") as f:
        f.write(b"P6\n# a comment\n2 1\n# another\n255\n" + bytes([1, 2, 3, 4, 5, 6]))
    assert b2.read_image(p, order="rgb").tolist() == [[[1, 2, 3], [4, 5, 6]]]
    with open(p, "wb") as f:
        f.write(b"P6\n2 2\n255\n" + bytes(5))
    with pytest.raises(b2.B200ImgError):
        b2.read_image(p)
    with pytest.raises(b2.B200ImgError):
        b2.read_image(str(tmp_path / "missing.ppm"))
    with open(p, "wb") as f:
        f.write(b"P3\n1 1\n255\n1 2 3\n")
    with pytest.raises(b2.B200ImgError):
        b2.read_image(p)


def test_raw_roundtrip(tmp_path):
    img = b2.synth_image(19, 11, seed=2)
    p = str(tmp_path / "a.raw")
    b2.write_raw(p, img)
    assert os.path.getsize(p) == 19 * 11 * 3            # the reference's scatter payload size
    assert np.array_equal(b2.read_raw(p, 19, 11, 3), img)
    with pytest.raises(b2.B200ImgError):
        b2.read_raw(p, 20, 11, 3)


def test_synth_is_deterministic_and_full_range():
    a, b = b2.synth_image(256, 128, seed=9), b2.synth_image(256, 128, seed=9)
    assert np.array_equal(a, b) and not np.array_equal(a, b2.synth_image(256, 128, seed=10))
    assert len(np.unique(a)) == 256
