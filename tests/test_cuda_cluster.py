"""Thread-block clusters with a multicast TMA load for the rows two stacked tiles share
(csrc/kernels/cluster_probe.cu): both variants of the probe must produce the same bytes as the
host rule out[y] = in[y-R] ^ in[y+R]. The buffers are the probe's own cudaMalloc allocations on the
current GPU. Reference for scale: kernel.cu:195 uses one thread per pixel and no tiles."""
import ctypes as C

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("W3,H,T,R", [(1024, 200, 16, 2), (2304, 131, 32, 2), (768, 64, 8, 1), (4096, 300, 64, 4)])
def test_cluster_multicast_tiles_match_plain_tiles(W3, H, T, R):
    from mpi_cuda_imagemanipulation_b200._lib import lib, check
    torch.cuda.init()
    a, b, bad = C.c_float(), C.c_float(), C.c_longlong(-1)
    check(lib.b2i_cuda_cluster_probe(W3, H, T, R, 2, C.byref(a), C.byref(b), C.byref(bad)))
    assert bad.value == 0
    assert a.value > 0 and b.value > 0
