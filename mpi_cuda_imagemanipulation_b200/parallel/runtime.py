"""In-process multi-GPU runtime (wraps ``b2i_ctx_t``, ``csrc/runtime/context.cpp``).

One Python process drives N GPUs — the stand-in for ``mpirun -np N`` (``kernel.cu:104-107``).
With fewer physical GPUs than requested the bands share devices ("virtual GPUs"), which keeps
the whole multi-GPU code path testable on a single B200.
"""
from __future__ import annotations

import ctypes as C
import json
from typing import Optional, Sequence

import numpy as np

from .. import _lib
from .._lib import check, lib
from ..ops.chain import as_chain, image_geometry


def topology() -> dict:
    """Devices and the peer-to-peer matrix as the C++ runtime sees them."""
    buf = C.create_string_buffer(1 << 16)
    check(lib.b2i_topology_json(buf, len(buf)))
    return json.loads(buf.value.decode())


class Runtime:
    """N GPUs of one node, one root GPU that owns input and output (like rank 0)."""

    def __init__(self, gpus: int = 1, devices: Optional[Sequence[int]] = None, root: int = 0):
        self._c = C.c_void_p()
        dev = (C.c_int * gpus)(*devices) if devices is not None else None
        check(lib.b2i_ctx_create(gpus, dev, root, C.byref(self._c)))
        self.gpus, self.root = gpus, root
        self._plans = {}
        self.last_ms = None
        self.last_per_gpu_ms = None

    def close(self):
        c, self._c = getattr(self, "_c", None), None
        if c and lib is not None:
            lib.b2i_ctx_destroy(c)

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _plan(self, chain, cin, **kw):
        from ..ops.cuda import Plan
        ch = as_chain(chain, **kw)
        key = (ch.describe(), bytes(ch.c), cin)
        if key not in self._plans:
            self._plans[key] = Plan(ch, cin)
        return self._plans[key]

    def run(self, img: np.ndarray, chain, *, mode: str = "root", iters: int = 1, passes: int = 1,
            sync: str = "events", no_tma: bool = False, root_share=None, replicate: bool = False,
            **kw) -> np.ndarray:
        """Host image in, host image out.

        mode="root":     scatter/gather fused into the kernels through the root GPU's memory.
                         root_share=f keeps f of the rows on the root GPU, "auto" measures the
                         rates with equal bands first and balances (link-aware band plan).
        mode="resident": bands stay on their GPUs for ``passes`` applications of the chain,
                         halos pushed by the kernels; ``sync`` is "events" or "flags".
        mode="direct":   each GPU moves its band over its own PCIe link.

        replicate=True (mode="root"): the result ends on every GPU, not on the root only — each
        kernel stores its band once to an NVLS multicast address and the NVSwitch writes all
        copies. ``self.replicas`` then holds every GPU's copy (host arrays), ``self.nvls`` says
        whether a multicast object (True) or the one-GPU stand-in (False) carried it.
        """
        img = np.ascontiguousarray(img, dtype=np.uint8)
        H, W, Cn = image_geometry(img)
        plan = self._plan(chain, Cn, **kw)
        oc = plan.cout
        out = np.empty((H, W, oc), np.uint8)
        flags = _lib.LAUNCH_NO_TMA if no_tma else 0
        ms = C.c_double(0)
        per = (C.c_double * self.gpus)()
        if mode == "direct":
            check(lib.b2i_ctx_run_direct(self._c, plan.handle, img.ctypes.data, W, H, Cn, W * Cn,
                                         out.ctypes.data, W * oc, flags, C.byref(ms)))
        else:
            if replicate and mode != "root":
                raise ValueError("replicate=True needs mode='root'")
            check(lib.b2i_ctx_set_replicated(self._c, int(bool(replicate))))
            share = 0.0 if root_share in (None, "auto") else float(root_share)
            check(lib.b2i_ctx_set_root_share(self._c, share if mode == "root" else 0.0))
            check(lib.b2i_ctx_load_root(self._c, img.ctypes.data, W, H, Cn, W * Cn, oc))
            if mode == "root":
                check(lib.b2i_ctx_run_root(self._c, plan.handle, iters, flags, C.byref(ms), per))
                if root_share == "auto" and self.gpus > 1:
                    # measured rates with equal bands -> the share at which root and the rest finish together
                    t = list(per)
                    rows = [H * (i + 1) // self.gpus - H * i // self.gpus for i in range(self.gpus)]
                    t_root = max(t[self.root], 1e-6)
                    t_rem = max(max(x for i, x in enumerate(t) if i != self.root), 1e-6)
                    r_root = rows[self.root] / t_root
                    r_rem = sum(r for i, r in enumerate(rows) if i != self.root) / t_rem
                    model = r_root / (r_root + r_rem)
                    # the model ignores the slowdown peer traffic causes on the root and the per-band
                    # sync cost: time its share and a few larger ones (up to 1.0 = the link cannot pay
                    # for itself, the root keeps the image) and keep the fastest
                    best = None
                    for cand in sorted({min(1.0, round(model + d, 3)) for d in (0.0, 0.04, 0.08, 0.12)} | {1.0}):
                        check(lib.b2i_ctx_set_root_share(self._c, cand))
                        check(lib.b2i_ctx_load_root(self._c, img.ctypes.data, W, H, Cn, W * Cn, oc))
                        check(lib.b2i_ctx_run_root(self._c, plan.handle, iters, flags, C.byref(ms), per))
                        if best is None or ms.value < best[1] * 0.98:
                            best = (cand, ms.value)
                    self.root_share = best[0]
                    check(lib.b2i_ctx_set_root_share(self._c, self.root_share))
                    check(lib.b2i_ctx_load_root(self._c, img.ctypes.data, W, H, Cn, W * Cn, oc))
                    check(lib.b2i_ctx_run_root(self._c, plan.handle, iters, flags, C.byref(ms), per))
            elif mode == "resident":
                check(lib.b2i_ctx_run_resident(self._c, plan.handle, passes, flags,
                                               int(sync == "flags"), C.byref(ms), per))
            else:
                raise ValueError(f"unknown mode {mode!r}")
            check(lib.b2i_ctx_fetch_root(self._c, out.ctypes.data, W * oc))
            self.replicas = None
            if replicate:
                self.nvls = lib.b2i_ctx_replicated(self._c) == 2
                self.replicas = []
                for g in range(self.gpus):
                    r = np.empty((H, W, oc), np.uint8)
                    check(lib.b2i_ctx_fetch_replica(self._c, g, r.ctypes.data, W * oc))
                    self.replicas.append(r)
        self.last_ms = ms.value
        self.last_per_gpu_ms = list(per)
        return out if img.ndim == 3 or oc == 3 else out[:, :, 0]
