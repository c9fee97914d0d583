"""Clusters sharing the overlapping rows of stacked stencil tiles: one TMA box per CTA against a
two-CTA cluster whose shared rows arrive by one multicast TMA load (csrc/kernels/cluster_probe.cu).
    python scripts/cluster_probe.py            # 8K RGB rows (23040 bytes x 4320), several tile heights"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_cuda_imagemanipulation_b200._lib import lib, check

W3, H = 23040, 4320
print(f"image {W3} bytes x {H} rows; out[y] = in[y-R] ^ in[y+R]; us per launch, 50 launches")
for T, R in ((16, 2), (32, 2), (64, 2), (16, 4), (32, 4), (8, 2)):
    a, b, bad = C.c_float(), C.c_float(), C.c_longlong()
    check(lib.b2i_cuda_cluster_probe(W3, H, T, R, 50, C.byref(a), C.byref(b), C.byref(bad)))
    over = 2 * R / (T + 2 * R)
    print(f"T={T:3d} R={R}: plain {a.value:6.1f} us   cluster+multicast {b.value:6.1f} us   "
          f"(shared rows = {over:.0%} of a plain box; wrong bytes: {bad.value})")
