"""The tensor-core stencil kernel (csrc/kernels/stencil_mma.cu) against the host reference, bit for bit.

It takes what the sweep kernels do not: one-channel planes, stencils that are not separable, and
images whose width is not a multiple of 8 pixels (in pitched buffers: TMA needs a 16-byte row
pitch). `tile_h == 128` in the plan's report is this kernel's signature (bands of 128 rows = the M
of its MMAs), so each test also proves which kernel ran. The reference's stencil for comparison is
kernel.cu:64-94 (one thread per pixel) and kern.cpp:75 (cv::filter2D).
"""
import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

FILTER5 = "filter5:1/2/3/4/5/-1/-2/-3/-4/-5/0/1/0/-1/0/2/2/2/2/2/1/1/1/1/-30/0"


def _run(chain, img, C, pitched=False):
    from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan, pitched_empty
    H, W = img.shape[:2]
    plan = Plan(chain, C)
    if pitched:
        src = pitched_empty(H, W, C)
        src.copy_(torch.from_numpy(img.reshape(H, W, C)).cuda())
        out = pitched_empty(H, W, plan.cout)
        plan.run(src, out)
    else:
        out = plan.run(torch.from_numpy(img).cuda())
    torch.cuda.synchronize()
    return out.cpu().numpy(), plan.info()


def _check(chain, W, H, C, pitched=False, seed=5):
    img = b2.synth_image(W, H, C, seed=seed)
    want = b2.host_run(chain, img)
    got, info = _run(chain, img, C, pitched)
    assert info["tile_h"] == 128, f"{chain}: the tensor-core kernel did not run ({info})"
    got = got.reshape(want.shape)
    if not np.array_equal(got, want):
        bad = np.argwhere(got.reshape(H, -1) != want.reshape(H, -1))
        raise AssertionError(f"{chain} {W}x{H}x{C}: {len(bad)} mismatches, rows {bad[:, 0].min()}-{bad[:, 0].max()} "
                             f"cols {bad[:, 1].min()}-{bad[:, 1].max()}")


@pytest.mark.parametrize("chain", ["gauss5", "gauss3", "sobel", "sharpen", "emboss3@reflect101", "emboss5@replicate",
                                   "gauss5@zero", FILTER5 + "@reflect101"])
@pytest.mark.parametrize("W,H", [(640, 200), (4096, 131), (1024, 700)])
def test_one_channel_stencils(chain, W, H):
    _check(chain, W, H, 1)


@pytest.mark.parametrize("chain", ["sharpen", "emboss3@reflect101", "emboss5@zero", FILTER5 + "@replicate",
                                   "sharpen@replicate"])
@pytest.mark.parametrize("W,H", [(512, 256), (2048, 300), (7680, 140)])
def test_non_separable_stencils_on_three_channels(chain, W, H):
    _check(chain, W, H, 3)


@pytest.mark.parametrize("chain,C", [("gauss5", 3), ("sobel", 3), ("gauss5", 1), ("sharpen", 3), ("gauss3@replicate", 1)])
@pytest.mark.parametrize("W,H", [(1001, 333), (515, 130), (7681, 64)])
def test_odd_widths_in_pitched_buffers(chain, C, W, H):
    """Any width: the tensor maps are over bytes, only the pitch has to be a multiple of 16."""
    _check(chain, W, H, C, pitched=True)


def test_whole_8k_image_many_runs_per_sm():
    """148 CTAs x ~41 slabs each: runs that start and end in the middle of a band."""
    _check("sharpen", 7680, 4320, 3)


def test_gray_then_blur_is_two_launches():
    """BASELINE.json config 5's chain: the streaming kernel, then the tensor-core kernel on the gray plane."""
    from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
    img = b2.synth_image(3840, 2160, 3, seed=2)
    plan = Plan("gray,gauss5", 3)
    out = plan.run(torch.from_numpy(img).cuda())
    torch.cuda.synchronize()
    info = plan.info()
    assert info["launches"] == 2 and info["tile_h"] == 128, info
    assert np.array_equal(out.cpu().numpy().reshape(2160, 3840), b2.host_run("gray,gauss5", img).reshape(2160, 3840))


@pytest.mark.parametrize("W,H", [(1001, 1001), (7681, 90), (1917, 300), (523, 64)])
def test_reference_pipeline_on_odd_widths_in_pitched_buffers(W, H):
    """kernel_cu (border = ref, kernel.cu:83) keeps the sweep kernel on any width when the rows are pitched:
    `tile_h < 0` is the sweep kernels' signature (minus the rows per segment)."""
    from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan, pitched_empty
    img = b2.synth_image(W, H, 3, seed=8)
    plan = Plan("kernel_cu", 3)
    src = pitched_empty(H, W, 3)
    src.copy_(torch.from_numpy(img).cuda())
    out = pitched_empty(H, W, 3)
    plan.run(src, out)
    torch.cuda.synchronize()
    assert plan.info()["tile_h"] < 0, plan.info()
    assert np.array_equal(out.cpu().numpy(), b2.host_run("kernel_cu", img))
