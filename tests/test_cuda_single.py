"""Single-GPU kernels against the C++ host reference, bit-exact.

Every test here needs a B200: `pytest -m gpu`. The host reference itself is tested against
NumPy / OpenCV in test_host_numerics.py, so equality here pins the kernels to the reference's
numerics (kernel.cu:31-94, kern.cpp:73-77).
"""
import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _run(chain, img_np, **kw):
    from mpi_cuda_imagemanipulation_b200.ops.cuda import cuda_run
    t = torch.from_numpy(img_np).cuda()
    out = cuda_run(chain, t, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _check(chain, W, H, C=3, seed=0, **kw):
    img = b2.synth_image(W, H, C, seed=seed)
    order = kw.pop("order", "bgr")
    want = b2.host_run(chain, img, order=order)
    got = _run(chain, img, order=order, **kw)
    assert got.shape == want.shape, (got.shape, want.shape)
    if not np.array_equal(got, want):
        bad = np.argwhere(got != want)
        raise AssertionError(f"{chain} {W}x{H}x{C} {kw}: {len(bad)} mismatches, first at "
                             f"{bad[0].tolist()} got {got[tuple(bad[0])]} want {want[tuple(bad[0])]}; "
                             f"rows {bad[:, 0].min()}-{bad[:, 0].max()} cols {bad[:, 1].min()}-{bad[:, 1].max()}")


def test_library_sees_the_gpu():
    assert b2.lib.b2i_cuda_device_count() >= 1


# ---- pointwise streaming kernel ---------------------------------------------------------------
@pytest.mark.parametrize("chain", ["gray", "gray_cv", "invert", "contrast:3.5", "gray,expand",
                                   "gray_cv,contrast_cv:3,expand", "invert,gray,contrast:2",
                                   "gray,contrast:3.5,expand,invert"])
@pytest.mark.parametrize("W,H", [(512, 512), (1000, 37), (1001, 19), (7680, 64), (16, 16), (5, 3)])
def test_pointwise_chains(chain, W, H):
    _check(chain, W, H)


@pytest.mark.parametrize("chain", ["invert", "contrast:1.5,expand", "expand", "threshold:99"])
@pytest.mark.parametrize("W,H", [(512, 64), (333, 21)])
def test_pointwise_on_gray_input(chain, W, H):
    _check(chain, W, H, C=1)


def test_gray_rgb_order_and_all_values():
    _check("gray", 640, 48, order="rgb")
    _check("gray_cv", 640, 48, order="rgb")
    for ch in range(3):  # every byte value through every channel's coefficient
        img = np.zeros((16, 256, 3), np.uint8)
        img[:, :, ch] = np.arange(256)
        assert np.array_equal(_run("gray", img), b2.host_run("gray", img))
        assert np.array_equal(_run("kernel_cu", img), b2.host_run("kernel_cu", img))


# ---- tile-chain kernel --------------------------------------------------------------------------
SIZES = [(512, 512), (1024, 96), (1920, 1080), (1000, 333), (1001, 77), (272, 40), (256, 32), (48, 500)]


@pytest.mark.parametrize("preset", ["kernel_cu", "kern_cpp", "kernel_cu_5x5"])
@pytest.mark.parametrize("W,H", SIZES)
def test_presets(preset, W, H):
    _check(preset, W, H)


@pytest.mark.parametrize("border", ["ref", "reflect101", "replicate", "zero"])
@pytest.mark.parametrize("stencil", ["emboss3", "emboss5", "gauss5", "gauss3", "sobel", "sharpen"])
@pytest.mark.parametrize("C", [1, 3])
def test_every_stencil_every_border(stencil, border, C):
    _check(f"{stencil}@{border}", 784, 150, C=C, seed=3)
    _check(f"{stencil}@{border}", 301, 67, C=C, seed=4)     # unaligned: plain ld/st loaders


@pytest.mark.parametrize("chain", [
    "gauss5,sobel", "gray,gauss5,sobel,expand", "gauss5@ref,emboss3@reflect101,sobel@zero",
    "gray_cv,gauss3@replicate,contrast:2,emboss5@reflect101,invert,expand",
    "emboss5,gray_cv,sobel,expand,gauss5", "invert,gauss5,gray,sobel",
    "gauss5,gauss5,gauss5,gauss5,gauss5",                     # 2 segments (halo 10 > 8)
    "gray,gauss5,gauss5,gauss5,gauss5,gauss5,sobel,expand",   # 2 segments, channel change
])
@pytest.mark.parametrize("W,H", [(1024, 200), (530, 91)])
def test_chains(chain, W, H):
    _check(chain, W, H, seed=11)


@pytest.mark.parametrize("tile_h", [8, 16, 24, 32])
def test_tile_heights(tile_h):
    _check("kernel_cu", 1024, 333, tile_h=tile_h)
    _check("gauss5,sobel", 1024, 333, tile_h=tile_h)


@pytest.mark.parametrize("border", ["reflect101", "replicate", "zero"])
@pytest.mark.parametrize("chain", ["gauss5", "gauss3", "sobel", "gauss5,sobel", "sobel,invert,gauss3"])
def test_sweep_kernels_strip_and_segment_edges(chain, border):
    """sweep5.cu / sweep_sobel.cu: widths around the 240-pixel strip, heights around the TMA box and
    segment heights, every border rule that reaches them."""
    from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
    ch = ",".join(f"{op}@{border}" if op != "invert" else op for op in chain.split(","))
    for W, H in [(32, 17), (240, 9), (496, 21), (256, 43), (480, 64), (976, 18), (960, 101), (1920, 19), (248, 30)]:
        _check(ch, W, H, seed=W + H)
    p = Plan(ch, 3)
    p.run(torch.from_numpy(b2.synth_image(1024, 256)).cuda())
    assert p.info()["tile_h"] < 0, p.info()     # negative: rows per sweep segment, i.e. a sweep kernel ran


def test_plain_loaders_match_tma():
    for chain in ("kernel_cu", "kern_cpp", "gauss5,sobel"):
        _check(chain, 1024, 256, no_tma=True)


def test_tma_is_actually_used_on_aligned_images():
    from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
    p = Plan("kernel_cu", 3)
    p.run(torch.from_numpy(b2.synth_image(1024, 256)).cuda())
    info = p.info()
    assert info["tma_load"] and info["tma_store"], info
    p.run(torch.from_numpy(b2.synth_image(1001, 64)).cuda())
    assert not p.info()["tma_load"]


def test_large_image_8k():
    _check("kernel_cu", 7680, 4320)


def test_reference_style_kernels_match_too():
    for chain in ("kernel_cu", "kern_cpp", "gauss5,sobel", "invert"):
        _check(chain, 640, 200, naive=True)


def test_degenerate_sizes_fall_back():
    _check("kernel_cu", 9, 7)
    _check("gauss5", 12, 40)


def test_output_into_preallocated_tensor_and_stream():
    from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan
    img = b2.synth_image(1024, 128)
    p = Plan("kernel_cu", 3)
    t = torch.from_numpy(img).cuda()
    out = torch.zeros(128, 1024, 3, dtype=torch.uint8, device="cuda")
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        r = p.run(t, out)
    s.synchronize()
    assert r.data_ptr() == out.data_ptr()
    assert np.array_equal(out.cpu().numpy(), b2.host_run("kernel_cu", img))


def test_batch_streamer_frames_from_pinned_memory():
    from mpi_cuda_imagemanipulation_b200.parallel.batch import BatchStreamer
    frames = [torch.from_numpy(b2.synth_image(640, 360, 3, seed=s)).pin_memory() for s in range(7)]
    outs = [torch.empty((360, 640, 1), dtype=torch.uint8).pin_memory() for _ in range(7)]
    n = BatchStreamer("gray,gauss5", gpus=torch.cuda.device_count()).run(frames, outs)
    assert n == 7
    for f, o in zip(frames, outs):
        assert np.array_equal(o.numpy(), b2.host_run("gray,gauss5", f.numpy()))


def test_pipeline_public_api_on_host_and_device_images():
    img = b2.synth_image(800, 300, 3, seed=3)
    want = b2.host_run("kernel_cu", img)
    pipe = b2.kernel_cu()
    assert np.array_equal(pipe(img), want)                                  # numpy in, numpy out
    assert np.array_equal(pipe(torch.from_numpy(img).cuda()).cpu().numpy(), want)   # CUDA tensor, zero-copy
    assert np.array_equal(b2.kern_cpp()(img), b2.host_run("kern_cpp", img))


# ---- resident multi-pass launch (one cooperative kernel walks the passes) ------------------------------
@pytest.mark.parametrize("W,H,passes", [(1920, 1080, 2), (1920, 1080, 5), (7680, 540, 8), (512, 64, 3), (7680, 2160, 4)])
def test_fused_passes_equal_chained_blurs(W, H, passes):
    """P passes of gauss5 in one launch (tasks ordered by per-task flags) == gauss5 applied P times."""
    from mpi_cuda_imagemanipulation_b200.ops.cuda import Plan, run_passes
    img = b2.synth_image(W, H, 3, seed=9)
    want = b2.host_run(",".join(["gauss5"] * passes), img)
    plan = Plan("gauss5", 3)
    got = run_passes(plan, torch.from_numpy(img).cuda(), passes, fused=True).cpu().numpy()
    assert plan.info()["launches"] == 1
    assert np.array_equal(got, want)
    got2 = run_passes(plan, torch.from_numpy(img).cuda(), passes, fused=False).cpu().numpy()
    assert np.array_equal(got2, want)
