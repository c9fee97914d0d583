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
    # the tensor-core stencil kernel works on bytes, so one-channel stencils stand alone too, and a
    # pointwise prefix is cut off in front of them (streaming kernel, then stencil_mma)
    assert _info("gauss5,sobel", c=1)["n_segments"] == 2
    assert [s["pointwise"] for s in _info("gray,gauss5")["segments"]] == [True, False]
    assert _info("gray,gauss5,sobel,expand")["n_segments"] == 4
    # ... but the reference pipeline's shape stays one launch (sweep3.cu)
    assert _info("kernel_cu")["n_segments"] == 1 and _info("kern_cpp")["n_segments"] == 1
    # REF borders are not the tensor-core kernel's: emboss5@ref after gray stays fused
    assert _info("gray,emboss5")["n_segments"] == 1
    # border=ref stencils are not separable sweeps
    assert _info("gauss5@ref,sobel@ref")["n_segments"] == 1


def test_long_chains_split_by_phase_and_halo_budget():
    """A tile segment holds at most 6 phases (a separable stencil is two) and a summed radius of 8."""
    # stencils with the reference's border rule stay on the tile kernel: its budgets cut the chain
    i = _info("gray,emboss5,emboss5,emboss5,emboss5,emboss5,sobel@ref,expand")
    assert i["n_segments"] == 2 and i["halo"] == 11
    assert sum(s["halo"] for s in i["segments"]) == 11 and i["segments"][-1]["cout"] == 3
    # the same chain of stencils the tensor-core kernel takes is one launch per op
    i = _info("gray,gauss5,gauss5,gauss5,gauss5,gauss5,sobel,expand")
    assert i["n_segments"] == 8 and i["halo"] == 11 and sum(s["halo"] for s in i["segments"]) == 11


def test_channel_mismatch_is_an_error():
    with pytest.raises(b2.B200ImgError):
        Plan("expand", 3)


def _plan_debug(chains):
    import os
    import subprocess
    import sys
    env = dict(os.environ, B200IMG_DEBUG="2")
    code = ("import sys\nfrom mpi_cuda_imagemanipulation_b200.ops.cuda import Plan\n"
            f"for c in {list(chains)!r}:\n    print('chain', c, file=sys.stderr, flush=True); Plan(c, 3)\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stderr
    out, cur = {}, None
    for line in r.stderr.splitlines():
        if line.startswith("chain "):
            cur = line[6:]
            out[cur] = ""
        elif cur is not None:
            out[cur] += line + "\n"
    return out


def test_contrast_tables_are_recognised_as_affine():
    """The sweep kernel computes clamp((a v + b) >> s) tables instead of looking them up; the planner
    finds (a, b, s) by exhaustive check (kernel.cu:56 with 3.5, kern.cpp:74)."""
    d = _plan_debug(["kernel_cu", "kern_cpp", "gray,invert,emboss3,expand", "gray,contrast:2.2,emboss3,expand"])
    assert "is clamp((7 v + -640) >> 1)" in d["kernel_cu"]
    assert "is clamp((3 v + -256) >> 0)" in d["kern_cpp"]
    assert "is clamp((-1 v + 255) >> 0)" in d["gray,invert,emboss3,expand"]
    assert "is not clamp" in d["gray,contrast:2.2,emboss3,expand"]


def test_library_path_override():
    import os
    import subprocess
    import sys
    from mpi_cuda_imagemanipulation_b200 import _lib
    env = dict(os.environ, B200IMG_LIB=_lib.LIB_PATH)
    r = subprocess.run([sys.executable, "-c", "import mpi_cuda_imagemanipulation_b200._lib as l; print(l.LIB_PATH)"],
                       capture_output=True, text=True, env=env,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0 and r.stdout.strip() == _lib.LIB_PATH, r.stderr


def test_resident_entry_points_reject_bad_arguments_without_a_gpu():
    """The multi-pass drivers validate before they launch (csrc/runtime/ipc.cu): no plan, no launch
    descriptions, a cycle period of 0. The reference's only argument check is the image load (kernel.cu:111)."""
    import ctypes as C
    from mpi_cuda_imagemanipulation_b200._lib import Launch, lib
    L = (Launch * 2)()
    assert lib.b2i_cuda_run_resident_cycle(None, L, 2, 4, 0) != 0
    plan = Plan("gauss5", 3)
    assert lib.b2i_cuda_run_resident_cycle(plan.handle, None, 2, 4, 0) != 0
    assert lib.b2i_cuda_run_resident_cycle(plan.handle, L, 0, 4, 0) != 0
    assert lib.b2i_cuda_run_resident_cycle(plan.handle, L, 2, 0, 0) == 0      # no passes: nothing to do
    assert lib.b2i_cuda_run_resident_steps(None, C.byref(L[0]), C.byref(L[1]), 1, 0) != 0
