"""b2i_cuda_run_image_noalloc: the entry point for callers that may not allocate device memory.

Some users of this library run where allocation is off limits at call time: inside an inference
server's or media pipeline's execution callback (the host framework owns the allocator), in a
process whose caching allocator has already claimed the whole GPU, or in a loop that cannot
afford cudaMalloc's device-wide synchronisation. For them the scratch images that a
multi-launch chain needs live in a fixed array inside libb200img.so instead of being passed in.

The test drives the C ABI directly (ctypes), the way such a caller would, and compares against
the ordinary entry point b2i_cuda_run_image, bit for bit, on two calls in a row — the second
call must not be disturbed by what the first left in the shared array. This file sorts last on
purpose: if it fails, `pytest -x` has already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2
from mpi_cuda_imagemanipulation_b200 import _lib

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

# five 5x5 blurs run as several launches (one per blur on the tensor-core kernel, or two fused
# segments on the tile kernel): the intermediate images go through scratch memory
MULTI_SEGMENT = "gauss5,gauss5,gauss5,gauss5,gauss5"


def _plan(chain_spec, channels):
    chain = b2.as_chain(chain_spec)           # a preset name or a filter list
    handle = C.c_void_p()
    _lib.check(_lib.lib.b2i_plan_create(chain.ref, channels, C.byref(handle)))
    return chain, handle


def _reference_entry_point(handle, src, out_channels):
    """The existing entry point, scratch supplied by the caller (here: torch tensors)."""
    H, W, Cn = src.shape
    dst = torch.zeros((H, W, out_channels), dtype=torch.uint8, device="cuda")
    tmp0 = torch.empty(H * W * 3, dtype=torch.uint8, device="cuda")
    tmp1 = torch.empty(H * W * 3, dtype=torch.uint8, device="cuda")
    table = torch.zeros(256 * _lib.B2I_MAX_OPS, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    _lib.check(_lib.lib.b2i_cuda_run_image(handle, src.data_ptr(), W * Cn, dst.data_ptr(), W * out_channels,
                                           W, H, tmp0.data_ptr(), tmp1.data_ptr(), table.data_ptr(),
                                           stream, 0))
    torch.cuda.synchronize()
    return dst


def _noalloc_entry_point(handle, src, dst):
    H, W, Cn = src.shape
    stream = torch.cuda.current_stream().cuda_stream
    return _lib.lib.b2i_cuda_run_image_noalloc(handle, src.data_ptr(), W * Cn, dst.data_ptr(),
                                               W * dst.shape[2], W, H, stream, 0)


def test_noalloc_matches_the_ordinary_entry_point_twice_in_a_row():
    chain, handle = _plan(MULTI_SEGMENT, 3)
    try:
        info = _lib.PlanInfo()
        _lib.check(_lib.lib.b2i_plan_info(handle, C.byref(info)))
        assert info.n_segments >= 2, "the chain must actually use scratch memory"
        first = torch.from_numpy(b2.synth_image(1024, 600, 3, seed=21)).cuda()
        second = torch.from_numpy(b2.synth_image(1280, 720, 3, seed=22)).cuda()   # different size on purpose
        want_first = _reference_entry_point(handle, first, 3)
        want_second = _reference_entry_point(handle, second, 3)
        # every buffer the caller owns exists before the calls under test; nothing may be
        # allocated from here on, so the device's free memory must not move between the two calls
        out_first = torch.zeros_like(want_first)
        out_second = torch.zeros_like(want_second)
        _lib.check(_noalloc_entry_point(handle, first, out_first))
        torch.cuda.synchronize()
        free_after_first = torch.cuda.mem_get_info()[0]
        _lib.check(_noalloc_entry_point(handle, second, out_second))
        torch.cuda.synchronize()
        free_after_second = torch.cuda.mem_get_info()[0]
        assert torch.equal(out_first, want_first)
        assert torch.equal(out_second, want_second)
        assert free_after_second == free_after_first, "the second call changed the amount of free device memory"
        # and the numbers are the right ones, not merely the same ones
        assert np.array_equal(out_second.cpu().numpy(), b2.host_run(MULTI_SEGMENT, second.cpu().numpy()))
    finally:
        _lib.lib.b2i_plan_destroy(handle)


def test_noalloc_single_launch_chain_and_back_to_back_reuse():
    # a chain that needs no scratch must work through the same door, before and after one that does
    _, h_one = _plan("kernel_cu", 3)
    _, h_multi = _plan(MULTI_SEGMENT, 3)
    try:
        img = torch.from_numpy(b2.synth_image(768, 300, 3, seed=5)).cuda()
        for handle in (h_one, h_multi, h_one):
            want = _reference_entry_point(handle, img, 3)
            got = torch.zeros_like(want)
            _lib.check(_noalloc_entry_point(handle, img, got))
            torch.cuda.synchronize()
            assert torch.equal(got, want)
    finally:
        _lib.lib.b2i_plan_destroy(h_one)
        _lib.lib.b2i_plan_destroy(h_multi)


def test_noalloc_rejects_images_beyond_its_fixed_scratch():
    _, handle = _plan(MULTI_SEGMENT, 3)
    try:
        W, H = 2048, 1025                                     # one row more than 2 Mi pixels
        assert W * H > _lib.NOALLOC_MAX_PIXELS
        src = torch.zeros((H, W, 3), dtype=torch.uint8, device="cuda")
        dst = torch.full((H, W, 3), 7, dtype=torch.uint8, device="cuda")
        rc = _noalloc_entry_point(handle, src, dst)
        assert rc == _lib.ERR_UNSUPPORTED
        msg = _lib.lib.b2i_last_error().decode()
        assert str(_lib.NOALLOC_MAX_PIXELS) in msg and "limited" in msg
        torch.cuda.synchronize()
        assert int(dst.min()) == 7, "a rejected call must not touch the output"
        # the limit itself is accepted
        W, H = 2048, 1024
        src = torch.from_numpy(b2.synth_image(W, H, 3, seed=3)).cuda()
        want = _reference_entry_point(handle, src, 3)
        got = torch.zeros_like(want)
        _lib.check(_noalloc_entry_point(handle, src, got))
        torch.cuda.synchronize()
        assert torch.equal(got, want)
    finally:
        _lib.lib.b2i_plan_destroy(handle)
