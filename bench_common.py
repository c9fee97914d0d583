"""What both arms of bench.py share: the config block (BASELINE.json config 2 with the filter chain
the reference actually runs) and the clock sampler. Pure Python: importing this must not load the
native library, because the reference arm may not map any of this repo's `.so` files."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.abspath(__file__))
W8K, H8K = 7680, 4320


def config_block(W: int, H: int, n: int) -> dict:
    """Identical in the two arms (the driver compares them)."""
    return {
        "model": "reference pipeline gray->contrast3.5->emboss3x3->BGR (kernel.cu:192-210)",
        "image": f"{W}x{H} RGB", "global_batch": 1, "seq_len": W * H,
        "parallelism": f"bands{n}" if n > 1 else "bands1",
        "l2": "working set 199 MB per step > 126 MB L2: our arm rotates over 4 input/output image pairs "
              "(796 MB) at N=1 and streams 99.5 MB in + 99.5 MB out per step at N>1; the reference arm is one "
              "fresh process per step (cold L2)",
    }


def clock_sampler_cls():
    """utils/clocks.py loaded by path, so that the package (and with it libb200img.so) stays out."""
    p = os.path.join(ROOT, "mpi_cuda_imagemanipulation_b200", "utils", "clocks.py")
    spec = importlib.util.spec_from_file_location("_b2i_clocks", p)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.ClockSampler
