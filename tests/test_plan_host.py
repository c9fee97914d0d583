"""The plan compiler runs on the host: how chains are cut into launches can be checked without a GPU."""
import pytest

import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan


def _info(chain, c=3):
    return Plan(chain, c).info()


def test_pointwise_chains_are_one_streaming_launch():
    for chain in ("gray", "invert", "gray,contrast:3.5,expand", "invert,gray,invert"):
        i = _info(chain)
        assert i["n_segments"] == 1 and i["segments"][0]["pointwise"], (chain, i)


def test_reference_pipeline_is_one_tile_segment_with_halo_1():
    i = _info("kernel_cu")
    assert i["n_segments"] == 1 and not i["segments"][0]["pointwise"]
    assert i["halo"] == 1 and i["cout"] == 3


def test_stencils_the_sweep_kernels_take_stand_alone():
    """gauss5 and sobel on 3-channel data each get their own launch (sweep5.cu, sweep_sobel.cu); the
    fused cut for the tile kernel is kept inside the plan and chosen per launch."""
    i = _info("gauss5,sobel")
    assert i["n_segments"] == 2 and i["halo"] == 3
    assert [s["halo"] for s in i["segments"]] == [2, 1]
    assert _info("gauss5")["n_segments"] == 1
    assert _info("gauss5,invert")["n_segments"] == 1          # a LUT rides on the stencil's result
    assert _info("gauss5,gauss5,gauss5")["n_segments"] == 3
    # on one channel there is no sweep kernel: the tile kernel takes the whole chain
    assert _info("gray,gauss5,sobel,expand")["n_segments"] == 1
    assert _info("gauss5,sobel", c=1)["n_segments"] == 1
    # border=ref stencils are not separable sweeps
    assert _info("gauss5@ref,sobel@ref")["n_segments"] == 1


def test_long_chains_split_by_phase_and_halo_budget():
    """A tile segment holds at most 6 phases (a separable stencil is two) and a summed radius of 8."""
    i = _info("gray,gauss5,gauss5,gauss5,gauss5,gauss5,sobel,expand")
    assert i["n_segments"] == 3 and i["halo"] == 11
    assert sum(s["halo"] for s in i["segments"]) == 11 and i["segments"][-1]["cout"] == 3


def test_channel_mismatch_is_an_error():
    with pytest.raises(b2.B200ImgError):
        Plan("expand", 3)
