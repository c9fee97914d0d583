"""NumPy restatement of the synthetic test card (csrc/core/io.cpp `b2i_synth_image`).

The reference arm of bench.py must not load this repo's native library (the driver records which
`.so` files each arm maps), so it generates the very same image bytes with NumPy and writes the
PPM itself. `tests/test_reference_arm.py` checks that the two generators agree byte for byte.
"""
import numpy as np


def synth_image(W: int, H: int, C: int = 3, seed: int = 0) -> np.ndarray:
    """HxWxC uint8 (channel order as in memory: the reference treats it as B, G, R)."""
    u32 = np.uint32
    x = np.arange(W, dtype=np.uint64)[None, :]
    y = np.arange(H, dtype=np.uint64)[:, None]
    m = np.uint64(0xFFFFFFFF)
    h = ((x * np.uint64(0x9E3779B1)) & m) ^ ((y * np.uint64(0x85EBCA77)) & m) ^ np.uint64((seed * 0xC2B2AE3D) & 0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & m
    h ^= h >> np.uint64(12)
    h = (h * np.uint64(0x297A2D39)) & m
    h ^= h >> np.uint64(15)
    h = h.astype(u32)
    noise = (h & u32(63)).astype(np.int32) - 32
    xi = np.arange(W, dtype=np.int64)[None, :]
    yi = np.arange(H, dtype=np.int64)[:, None]
    grad = ((xi * 255) // (W - 1 if W > 1 else 1)).astype(np.int32)
    vgrad = ((yi * 255) // (H - 1 if H > 1 else 1)).astype(np.int32)
    checker = np.where((((xi >> 5) ^ (yi >> 5)) & 1) != 0, 40, -40).astype(np.int32)
    dx, dy = xi - W // 2, yi - H // 2
    ring = np.where((((dx * dx + dy * dy) >> 12) & 1) != 0, 25, -25).astype(np.int32)
    salt = ((h >> u32(20)) % u32(97)) == 0
    saltv = ((h >> u32(24)) & u32(255)).astype(np.int32)
    out = np.empty((H, W, C), np.uint8)
    for c in range(C):
        if c == 0:
            v = grad + checker + noise
        elif c == 1:
            v = vgrad + ring + (noise >> 1)
        else:
            # C integer division truncates toward zero; grad + vgrad is never negative
            v = (grad + vgrad) // 2 - checker + ((h >> u32(8)) & u32(31)).astype(np.int32) - 16
        v = np.where(salt, saltv, v)
        out[:, :, c] = np.clip(v, 0, 255).astype(np.uint8)
    return out


def write_ppm_from_bgr(path: str, img: np.ndarray) -> None:
    """Binary P6 (RGB on disk) from an HxWx3 array in B, G, R memory order."""
    H, W, _ = img.shape
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (W, H))
        f.write(np.ascontiguousarray(img[:, :, ::-1]).tobytes())
