"""Independent NumPy / OpenCV restatements of the reference numerics.

These deliberately share no code with csrc/host_ref: the C++ host reference is tested against
them, and the CUDA kernels are tested against the C++ host reference.
"""
import numpy as np

EMBOSS3 = np.array([[-2, -1, 0], [-1, 1, 1], [0, 1, 2]], np.int64)          # kernel.cu:71-75
EMBOSS5 = np.diag([4, 4, 1, -4, -4]).astype(np.int64)                       # kernel.cu:76-82
SOBEL_X = np.array([[-1, 0, 1], [-2, 0, 2], [-1, 0, 1]], np.int64)
GAUSS5 = np.outer([1, 4, 6, 4, 1], [1, 4, 6, 4, 1]).astype(np.int64)
GAUSS3 = np.outer([1, 2, 1], [1, 2, 1]).astype(np.int64)


def gray_ref(img, order="bgr"):
    """kernel.cu:39-42 — every term truncated to uchar on its own, FP64 constants."""
    b, g, r = (img[..., 0], img[..., 1], img[..., 2]) if order == "bgr" else (
        img[..., 2], img[..., 1], img[..., 0])
    t = lambda ch, k: (ch.astype(np.float32).astype(np.float64) * k).astype(np.uint8).astype(np.int32)
    return (t(b, 0.11) + t(g, 0.59) + t(r, 0.3)).astype(np.uint8)


def gray_cv(img, order="bgr"):
    import cv2
    return cv2.cvtColor(img, cv2.COLOR_BGR2GRAY if order == "bgr" else cv2.COLOR_RGB2GRAY)


def contrast_ref(g, factor=3.5):
    """kernel.cu:56 — fp32 FMA, clamp, truncating store."""
    f = np.float32(factor)
    v = (g.astype(np.int32) - 128).astype(np.float32)
    # float32 fma: exact product in float64, one rounding
    y = (v.astype(np.float64) * np.float64(f) + 128.0).astype(np.float32)
    return np.clip(y, 0, 255).astype(np.uint8)


def contrast_cv(g, factor=3):
    return np.clip(np.rint(g.astype(np.float32) * np.float32(factor)
                           + np.float32(128 - 128 * factor)), 0, 255).astype(np.uint8)


def _pad(img, r, border):
    mode = {"reflect101": "reflect", "replicate": "edge", "zero": "constant"}[border]
    pw = ((r, r), (r, r)) + (((0, 0),) if img.ndim == 3 else ())
    return np.pad(img.astype(np.int64), pw, mode=mode)


def correlate(img, K, border):
    """Out-of-place correlation, anchor centre, with full-image padding (not REF)."""
    k = K.shape[0]
    r = k // 2
    H, W = img.shape[:2]
    p = _pad(img, r, border)
    s = np.zeros(img.shape, np.int64)
    for j in range(k):
        for i in range(k):
            if K[j, i]:
                s += K[j, i] * p[j:j + H, i:i + W]
    return s


def stencil(img, K, border="ref", shift=0, bias=0):
    k = K.shape[0]
    r = k // 2
    H, W = img.shape[:2]
    if border != "ref":
        return np.clip((correlate(img, K, border) + bias) >> shift, 0, 255).astype(np.uint8)
    # REF: x in [r+1, W-r-1], y in [r+1, H-r-1] computed, rest passes through
    out = img.copy()
    if W - r - 1 >= r + 1 and H - r - 1 >= r + 1:
        s = np.clip((correlate(img, K, "zero") + bias) >> shift, 0, 255).astype(np.uint8)
        out[r + 1:H - r, r + 1:W - r] = s[r + 1:H - r, r + 1:W - r]
    return out


def sobel(img, border="reflect101"):
    H, W = img.shape[:2]
    if border == "ref":
        out = img.copy()
        if W >= 4 and H >= 4:
            gx, gy = correlate(img, SOBEL_X, "zero"), correlate(img, SOBEL_X.T, "zero")
            s = np.clip(np.abs(gx) + np.abs(gy), 0, 255).astype(np.uint8)
            out[2:H - 1, 2:W - 1] = s[2:H - 1, 2:W - 1]
        return out
    gx, gy = correlate(img, SOBEL_X, border), correlate(img, SOBEL_X.T, border)
    return np.clip(np.abs(gx) + np.abs(gy), 0, 255).astype(np.uint8)


def expand(g):
    return np.repeat(g[..., None] if g.ndim == 2 else g, 3, axis=2)


def pipeline_kernel_cu(img, order="bgr"):
    return expand(stencil(contrast_ref(gray_ref(img, order)), EMBOSS3, "ref"))


def pipeline_kern_cpp(img):
    """The exact OpenCV calls of kern.cpp:62-77 through the Python binding of the same library."""
    import cv2
    kern = EMBOSS3.astype(np.int8)
    t = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
    t = cv2.addWeighted(t, 3.0, t, 0, -256.0)  # MatExpr 3*(m-128)+128 -> convertTo(3, -256)
    e = cv2.filter2D(t, -1, kern)
    return cv2.cvtColor(e, cv2.COLOR_GRAY2BGR)
