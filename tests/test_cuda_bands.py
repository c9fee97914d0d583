"""Row-band decomposition on the GPU: N-band result == 1-band result, bit-exactly.

On a one-GPU box the N bands are "virtual GPUs" sharing the device: peer pointers are local
pointers and launches serialise, but planner, views, ghost rows, halo pushes and the gather
are the very code that runs on N real GPUs (tests/test_multigpu.py covers real peers).
"""
import numpy as np
import pytest

import mpi_cuda_imagemanipulation_b200 as b2

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

CHAINS = ["kernel_cu", "kern_cpp", "gauss5,sobel", "gray", "invert",
          "gray,gauss5,gauss5,gauss5,gauss5,gauss5,sobel,expand"]


@pytest.mark.parametrize("chain", CHAINS)
@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
def test_root_mode_bands_equal_host(chain, n):
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime
    img = b2.synth_image(1024, 403, seed=n)        # 403 % n != 0 for n in 2,3,4,8
    with Runtime(gpus=n) as rt:
        got = rt.run(img, chain, mode="root")
    assert np.array_equal(got, b2.host_run(chain, img))


@pytest.mark.parametrize("n", [2, 3, 8])
def test_root_mode_unaligned_width_and_short_bands(n):
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime
    img = b2.synth_image(1001, 77, seed=5)
    with Runtime(gpus=n) as rt:
        assert np.array_equal(rt.run(img, "kernel_cu", mode="root"), b2.host_run("kernel_cu", img))
        assert np.array_equal(rt.run(img, "gauss5,sobel", mode="root"), b2.host_run("gauss5,sobel", img))


@pytest.mark.parametrize("chain,passes", [("gauss5", 1), ("gauss5", 4), ("gauss5,sobel", 3),
                                          ("emboss3@reflect101", 5), ("kernel_cu", 1),
                                          ("gauss3@ref,sharpen@zero", 3), ("invert", 3)])
@pytest.mark.parametrize("n", [1, 2, 4, 8])
def test_resident_passes_with_halo_push(chain, passes, n):
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime
    img = b2.synth_image(1024, 330, seed=n + passes)
    want = img
    for _ in range(passes):
        want = b2.host_run(chain, want)
    with Runtime(gpus=n) as rt:
        got = rt.run(img, chain, mode="resident", passes=passes)
    assert np.array_equal(got, want)


def test_resident_unaligned_width_uses_plain_stores_for_the_push():
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime
    img = b2.synth_image(1001, 200, seed=2)
    want = b2.host_run("gauss5", b2.host_run("gauss5", img))
    with Runtime(gpus=4) as rt:
        assert np.array_equal(rt.run(img, "gauss5", mode="resident", passes=2), want)
        assert np.array_equal(rt.run(img, "gauss5", mode="resident", passes=2, no_tma=True), want)


@pytest.mark.parametrize("n", [1, 2, 8])
def test_direct_mode(n):
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime
    img = b2.synth_image(1920, 540, seed=9)
    with Runtime(gpus=n) as rt:
        for chain in ("kernel_cu", "gauss5,sobel", "gray"):
            assert np.array_equal(rt.run(img, chain, mode="direct"), b2.host_run(chain, img))


@pytest.mark.parametrize("chain", ["kernel_cu", "gauss5,sobel", "invert", "gauss5"])
def test_replicated_output_plain_store_path(chain):
    """replicate=True: the last kernel leaves by plain stores to the (here: stand-in) multicast
    address. On one physical GPU all virtual GPUs share one copy; tests/test_multigpu.py checks
    the NVLS object on real peers."""
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime
    img = b2.synth_image(1024, 403, seed=4)
    want = b2.host_run(chain, img)
    with Runtime(gpus=4) as rt:
        got = rt.run(img, chain, mode="root", replicate=True)
        assert np.array_equal(got, want)
        assert len(rt.replicas) == 4 and all(np.array_equal(r, want) for r in rt.replicas)
        assert np.array_equal(rt.run(img, chain, mode="root"), want)      # and back to root-only output


def test_topology_report():
    from mpi_cuda_imagemanipulation_b200.parallel import topology
    t = topology()
    assert len(t["devices"]) >= 1 and t["devices"][0]["sms"] > 0
    assert len(t["p2p"]) == len(t["devices"])


@pytest.mark.parametrize("share", [0.5, 0.85, "auto"])
def test_root_mode_link_aware_band_plan(share):
    from mpi_cuda_imagemanipulation_b200.parallel import Runtime
    img = b2.synth_image(1024, 403, seed=4)
    with Runtime(gpus=4) as rt:
        assert np.array_equal(rt.run(img, "kernel_cu", mode="root", root_share=share), b2.host_run("kernel_cu", img))
        assert np.array_equal(rt.run(img, "gauss5,sobel", mode="root", root_share=share), b2.host_run("gauss5,sobel", img))
