"""One process per GPU (``torchrun``), the launch shape of ``mpirun -np N`` (``kernel.cu:104-107``).

``torch.distributed`` is used for plumbing only: exchanging 64-byte CUDA IPC handles, barriers
outside the timed regions and the MAX reduction of timings. Image bytes never travel through it:
every rank's filter kernel reads its row band (+halo) directly from the root GPU's image and
writes its rows directly into the root GPU's output through IPC-mapped peer memory over
NVLink/NVSwitch (scatter and gather fused into the filter pass), and resident multi-pass chains
push halo rows into the neighbours' ghost rows from inside the kernel and order the passes with
device-side release/acquire epoch flags instead of ``MPI_Barrier`` (``kernel.cu:127,133``).
"""
from __future__ import annotations

import ctypes as C
import json
import mmap
import os
import statistics
import time
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from .. import _lib
from .._lib import Launch, View, check, lib
from ..ops.chain import band_rows

G = 8                     # ghost rows above/below a resident band (kGhostRows)
W_COUNTERS, W_STATUS, W_FLAG_UP, W_FLAG_DN, W_DONE = 0, 16, 32, 36, 64   # byte offsets in `words`
W_SF_UP, W_SF_DN = 1024, 2048   # per-strip flags (256 u32 each) written by the upper / lower neighbour


class SymBuf:
    """The same-sized device buffer on every rank; ptrs[r] addresses rank r's copy from here."""

    def __init__(self, ptrs: List[int], nbytes: int, rank: int):
        self.ptrs, self.nbytes, self.rank = ptrs, nbytes, rank

    @property
    def local(self) -> int:
        return self.ptrs[self.rank]


class DistRuntime:
    def __init__(self, root: int = 0):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.root = root
        torch.cuda.set_device(self.local)
        self.device = torch.device("cuda", self.local)
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=self.device)
        self._bufs: List[SymBuf] = []
        self._shm = []
        self.words = self.symmetric(4096)
        self.epoch = 0
        self.timeout_ns = 0       # device-side waits give up after this long (0 = 2 s)
        self.geom = None

    # ---- plumbing ---------------------------------------------------------------------------------
    def symmetric(self, nbytes: int) -> SymBuf:
        ptr = C.c_uint64(0)
        handle = C.create_string_buffer(64)
        check(lib.b2i_ipc_alloc(nbytes, C.byref(ptr), handle))
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw))
        ptrs = []
        for r, h in enumerate(handles):
            if r == self.rank:
                ptrs.append(ptr.value)
            else:
                p = C.c_uint64(0)
                check(lib.b2i_ipc_open(h, C.byref(p)))
                ptrs.append(p.value)
        buf = SymBuf(ptrs, nbytes, self.rank)
        self._bufs.append(buf)
        return buf

    def shared_host(self, name: str, nbytes: int) -> np.ndarray:
        """A pinned host array every rank maps (POSIX shared memory): the root-owned host image."""
        path = [None]
        if self.rank == self.root:
            path[0] = f"/dev/shm/b2i_{os.getpid()}_{name}"
            with open(path[0], "wb") as f:
                f.truncate(nbytes)
        dist.broadcast_object_list(path, src=self.root)
        f = open(path[0], "r+b")
        mm = mmap.mmap(f.fileno(), nbytes)
        arr = np.frombuffer(mm, dtype=np.uint8)
        check(lib.b2i_host_register(arr.ctypes.data, nbytes))
        self._shm.append((path[0], f, mm, arr))
        dist.barrier()
        return arr

    def stream(self) -> int:
        return torch.cuda.current_stream().cuda_stream

    def barrier(self):
        torch.cuda.synchronize()
        dist.barrier()

    def close(self):
        torch.cuda.synchronize()
        dist.barrier()
        for path, f, mm, arr in self._shm:
            lib.b2i_host_unregister(arr.ctypes.data)
            del arr
            if self.rank == self.root:
                try:
                    os.unlink(path)
                except OSError:
                    pass
        self._shm = []
        for b in self._bufs:
            for r, p in enumerate(b.ptrs):
                if r != self.rank:
                    lib.b2i_ipc_close(p)
        dist.barrier()
        for b in self._bufs:
            lib.b2i_dev_free(b.local)
        self._bufs = []

    def status(self) -> int:
        st = torch.empty(1, dtype=torch.int32)
        check(lib.b2i_cuda_memcpy2d(st.data_ptr(), 4, self.words.local + W_STATUS, 4, 4, 1, 1, self.stream()))
        torch.cuda.synchronize()
        return int(st.item())

    def check_status(self) -> None:
        """Collective. Device-side waits never hang: a wait that times out (or meets a flag its writer
        poisoned when it gave up) records a code in this rank's status word and the kernel carries on to
        its end. Here every rank learns whether ANY rank recorded one: if so the flags, epochs and the
        status words are reset on every rank, so the runtime stays usable, and every rank raises
        ``B200ImgError`` with ``B2I_ERR_TIMEOUT``. (The reference's ranks hang in ``MPI_Barrier`` when
        one of them fails, ``kernel.cu:111-114,127``.)"""
        st = self.status()
        t = torch.tensor([st], device=self.device, dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        worst = int(t.item())
        if worst == 0:
            return
        self.reset_flags()
        raise _lib.B200ImgError(_lib.B2I_ERR_TIMEOUT,
                                f"a device-side wait gave up (this rank's status {st}, worst over ranks {worst}): "
                                f"1/2 = upper/lower neighbour's halo flag, 3 = a done flag, 4 = a task flag, 5 = a TMA load")

    def reset_flags(self) -> None:
        """Collective: zeroes every rank's flag words and task flags and restarts the epochs."""
        self.barrier()
        z = torch.zeros(self.words.nbytes, dtype=torch.uint8, device=self.device)
        check(lib.b2i_cuda_memcpy2d(self.words.local, self.words.nbytes, z.data_ptr(), self.words.nbytes,
                                    self.words.nbytes, 1, 2, self.stream()))
        if getattr(self, "_task_flags", None) is not None:
            self._task_flags.zero_()
        self.epoch = 0
        self.barrier()

    # ---- ROOT mode: root-owned device images, scatter/gather fused into the kernels ------------------
    def setup_root(self, W: int, H: int, Cn: int, out_c: int):
        sp, dp = (W * Cn + 15) // 16 * 16, (W * out_c + 15) // 16 * 16
        self.geom = dict(W=W, H=H, C=Cn, out_c=out_c, sp=sp, dp=dp)
        self.src = self.symmetric(sp * H)      # only the root's copy is used
        self.dst = self.symmetric(dp * H)
        self.set_root_share(None)

    def set_root_share(self, share):
        """Band plan of ROOT mode. share=None: equal bands (the reference's plan, kernel.cu:117, with
        the remainder kept). share=f in (0,1): the root GPU takes the first f*H rows, the other
        ranks split the rest evenly. The root reads and writes its own HBM while every other rank
        moves its band through the root's one NVLink port, so equal bands leave the root idle
        most of the step; :meth:`calibrate_root` measures both rates and picks f."""
        H = self.geom["H"]
        if share is None or self.world == 1:
            self.bands = [band_rows(H, self.world, r) for r in range(self.world)]
        else:
            cut = min(H, max(0, int(round(H * share))))
            self.bands = [(0, cut)] + [tuple(cut + b for b in band_rows(H - cut, self.world - 1, r))
                                       for r in range(self.world - 1)]
            if self.root != 0:
                raise NotImplementedError("weighted bands assume root 0")
        self.root_share = share
        self.y0, self.y1 = self.bands[self.rank]

    def load_root(self, img: np.ndarray):
        g = self.geom
        if self.rank == self.root:
            a = np.ascontiguousarray(img, dtype=np.uint8)
            check(lib.b2i_cuda_memcpy2d(self.src.local, g["sp"], a.ctypes.data, g["W"] * g["C"],
                                        g["W"] * g["C"], g["H"], 0, self.stream()))
        self.barrier()

    def fetch_root(self) -> Optional[np.ndarray]:
        g = self.geom
        self.barrier()
        if self.rank != self.root:
            return None
        out = np.empty((g["H"], g["W"], g["out_c"]), np.uint8)
        check(lib.b2i_cuda_memcpy2d(out.ctypes.data, g["W"] * g["out_c"], self.dst.local, g["dp"],
                                    g["W"] * g["out_c"], g["H"], 1, self.stream()))
        torch.cuda.synchronize()
        return out

    def _attach_tmp(self, L: Launch, plan, W: int, H: int):
        """Per-rank scratch images for plans of several segments (rows of the band plus its halo)."""
        if plan is None or plan.n_segments <= 1 or L.y1 <= L.y0:
            return
        ra, rb = max(0, L.y0 - plan.halo), min(H, L.y1 + plan.halo)
        pitch = (W * 3 + 15) // 16 * 16
        need = pitch * (rb - ra)
        if getattr(self, "_tmp", None) is None or self._tmp[0].numel() < need:
            self._tmp = [torch.empty(need, dtype=torch.uint8, device=f"cuda:{torch.cuda.current_device()}")
                         for _ in range(2)]
        for k in range(2):
            L.tmp[k] = View(self._tmp[k].data_ptr(), pitch, ra, rb)

    def _root_launch(self, no_tma=False, plan=None) -> Launch:
        g = self.geom
        L = Launch()
        L.W, L.H, L.y0, L.y1 = g["W"], g["H"], self.y0, self.y1
        self._attach_tmp(L, plan, g["W"], g["H"])
        L.src = View(self.src.ptrs[self.root], g["sp"], 0, g["H"])    # peer loads  = the scatter
        L.dst = View(self.dst.ptrs[self.root], g["dp"], 0, g["H"])    # peer stores = the gather
        L.counters = self.words.local + W_COUNTERS
        L.status = self.words.local + W_STATUS
        L.stream = self.stream()
        L.timeout_ns = self.timeout_ns
        L.flags = _lib.LAUNCH_NO_TMA if no_tma else 0
        return L

    def run_root(self, plan, steps: int = 1, no_tma: bool = False):
        """`steps` times: every rank produces its band of chain(root image) into the root's output.
        On the root, each step ends with a device-side wait for the other ranks' done flags."""
        L = self._root_launch(no_tma, plan)
        n_wait, flags = 0, (C.c_uint64 * _lib.B2I_MAX_GPUS)()
        if self.rank == self.root:
            for r in range(self.world):
                if r != self.root and self.bands[r][1] > self.bands[r][0]:
                    flags[n_wait] = self.words.local + W_DONE + 4 * r
                    n_wait += 1
        signal = 0 if self.rank == self.root else self.words.ptrs[self.root] + W_DONE + 4 * self.rank
        if self.y1 > self.y0 or n_wait:
            check(lib.b2i_cuda_run_steps(plan.handle, C.byref(L), steps, self.epoch, flags, n_wait, signal))
        self.epoch += steps

    def calibrate_root(self, plan, steps: int = 8) -> float:
        """Measures the root's local rate and the other ranks' combined (link-bound) rate with equal
        bands, then sets the root's share so that both finish together. Collective."""
        if self.world == 1:
            return 1.0
        self.set_root_share(None)
        self.run_root(plan, 3)
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # every rank times only its own band kernels (no waiting for the others)
        L = self._root_launch()
        dist.barrier()
        e0.record()
        for _ in range(steps):
            check(lib.b2i_cuda_run(plan.handle, C.byref(L)))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        rows = self.y1 - self.y0
        t = torch.tensor([ms if self.rank == self.root else 0.0, 0.0 if self.rank == self.root else ms,
                          float(rows if self.rank == self.root else 0), float(0 if self.rank == self.root else rows)],
                         device=self.device, dtype=torch.float64)
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)      # times: root's, slowest remote's
        dist.all_reduce(t, op=dist.ReduceOp.SUM)         # rows: root's, all remotes'
        r_root = t[2].item() / max(tmax[0].item(), 1e-6)
        r_rem = t[3].item() / max(tmax[1].item(), 1e-6)  # remotes run concurrently on one shared link
        share = r_root / (r_root + r_rem)
        # The model ignores what the two sides do to each other: peer traffic into the root's HBM
        # slows the root's own kernel (measured at N = 2: 25 %), and every remote band costs a flag
        # round trip. So the model's share and a few larger ones are timed as whole steps, and the
        # fastest wins -- up to 1.0, where the root keeps everything because the link cannot pay
        # for itself. All ranks see the same all-reduced times, hence pick the same share.
        best, best_ms = share, None
        for cand in sorted({min(1.0, round(share + d, 3)) for d in (0.0, 0.04, 0.08, 0.12)} | {1.0}):
            self.set_root_share(cand)
            self.run_root(plan, 2)
            self.barrier()
            ms_c = min(_timed(lambda: self.run_root(plan, steps)) / steps for _ in range(2))
            if best_ms is None or ms_c < best_ms * 0.98:   # a larger share must win clearly
                best, best_ms = cand, ms_c
        self.set_root_share(best)
        self.barrier()
        return best

    # ---- RESIDENT mode: bands (+ghost rows) stay on their GPUs, kernels push halos ---------------------
    def setup_resident(self, W: int, H: int, Cn: int):
        bp = (W * 3 + 15) // 16 * 16
        rows = max(band_rows(H, self.world, r)[1] - band_rows(H, self.world, r)[0] for r in range(self.world))
        self.rgeom = dict(W=W, H=H, C=Cn, bp=bp)
        # two band buffers ping-pong; run_resident_cycle rotates through four (see there)
        self.band = [self.symmetric(bp * (rows + 2 * G)) for _ in range(4)]
        self._last_out = None
        self.y0, self.y1 = band_rows(H, self.world, self.rank)   # resident bands are equal: no root link involved

    def scatter_resident(self, src_ptr_on_root: int, src_pitch: int):
        """Band + ghost rows from the root image into buffer 0 (a peer read, once)."""
        g = self.rgeom
        ra, rb = max(0, self.y0 - G), min(g["H"], self.y1 + G)
        check(lib.b2i_cuda_memcpy2d(self.band[0].local + g["bp"] * (ra - (self.y0 - G)), g["bp"],
                                    src_ptr_on_root + src_pitch * ra, src_pitch, g["W"] * g["C"], rb - ra, 2,
                                    self.stream()))
        self.barrier()

    def _resident_launch(self, inb: int, halo: int, no_tma=False, plan=None, strip_flags=False,
                         ext: int = 0, push: bool = True, outb=None, wait: bool = True, signal: bool = True) -> Launch:
        """One pass over this rank's band: reads band buffer `inb`, writes `outb` (default: the other one
        of the pair). The rest is for run_resident_cycle:
        ext > 0: the pass also computes `ext` ghost rows above and below the band (towards neighbours
        only; the image ends where it ends) from the deeper ghost rows of its input.
        push=False: no halo rows leave. wait / signal: whether the tasks near the band's ends wait for the
        neighbours' flags / release the neighbours' flags when they are done."""
        g = self.rgeom
        outb = inb ^ 1 if outb is None else outb
        L = Launch()
        up = self.rank - 1 if self.rank > 0 else None
        dn = self.rank + 1 if self.rank + 1 < self.world else None
        ya = self.y0 - (ext if up is not None else 0)
        yb = self.y1 + (ext if dn is not None else 0)
        L.W, L.H, L.y0, L.y1 = g["W"], g["H"], ya, yb
        self._attach_tmp(L, plan, g["W"], g["H"])
        va, vb = max(0, self.y0 - G), min(g["H"], self.y1 + G)
        L.src = View(self.band[inb].local + g["bp"] * (va - (self.y0 - G)), g["bp"], va, vb)
        L.dst = View(self.band[outb].local + g["bp"] * (G - (self.y0 - ya)), g["bp"], ya, yb)
        L.counters = self.words.local + W_COUNTERS
        L.status = self.words.local + W_STATUS
        if halo > 0:
            # rows from each end of the launch whose tasks wait and signal: with push, the rows that leave;
            # without, the rows whose stencil reads ghost rows (ext of them are ghost rows themselves)
            L.push_rows = halo if push else ext + halo
            if up is not None:
                uy0, uy1 = band_rows(g["H"], self.world, up)
                if push:
                    L.push_up = View(self.band[outb].ptrs[up] + g["bp"] * (G + (uy1 - uy0)), g["bp"], self.y0, self.y0 + G)
                if wait:
                    L.wait_up_flag = self.words.local + (W_SF_UP if strip_flags else W_FLAG_UP)
                if signal:
                    L.signal_up_flag = self.words.ptrs[up] + (W_SF_DN if strip_flags else W_FLAG_DN)
            if dn is not None:
                if push:
                    L.push_dn = View(self.band[outb].ptrs[dn] + g["bp"] * (G - halo), g["bp"], self.y1 - halo, self.y1)
                if wait:
                    L.wait_dn_flag = self.words.local + (W_SF_DN if strip_flags else W_FLAG_DN)
                if signal:
                    L.signal_dn_flag = self.words.ptrs[dn] + (W_SF_UP if strip_flags else W_FLAG_UP)
        L.stream = self.stream()
        L.timeout_ns = self.timeout_ns
        L.flags = (_lib.LAUNCH_NO_TMA if no_tma else 0) | (_lib.LAUNCH_STRIP_FLAGS if strip_flags else 0)
        return L

    N_TASK_FLAGS = 16384

    def run_resident_cycle(self, plan, passes: int, cycle: int) -> None:
        """`passes` applications with halo rows exchanged only every `cycle` passes (2 or 4; `passes` must
        be a multiple; cycle * halo <= the G = 8 ghost rows). Same result, bit for bit.

        Pass 0 of a cycle reads cycle * halo ghost rows the neighbours pushed and also computes
        (cycle - 1) * halo of them for itself, the next pass (cycle - 2) * halo, ...; the last pass computes
        the band only and pushes its first and last cycle * halo rows. Only pass 0 waits (for the rows) and
        only the last pass signals (that they are there); the passes between are ordinary single-GPU
        launches. No flag says "I am done reading": the cycles alternate between two PAIRS of band buffers
        (0, 1) and (2, 3), the last pass of a cycle writing (and pushing) into the other pair. A GPU pushes
        into a neighbour's buffer after it has seen, in its own pass 0, the neighbour's signal of the
        cycle before, which the neighbour sent after it had finished every pass that touched the ghost rows
        of that buffer. One launch per pass (`b2i_cuda_run_resident_cycle`)."""
        halo = plan.halo
        if cycle not in (2, 4) or passes % cycle or cycle * halo > G:
            raise ValueError(f"cycle {cycle} with halo {halo}: need cycle in (2, 4), passes % cycle == 0, cycle * halo <= {G}")
        if self.y1 > self.y0:
            descs = (Launch * (2 * cycle))()
            for c in range(2):
                pair, other = ((0, 1), 2) if c == 0 else ((2, 3), 0)
                for j in range(cycle):
                    last = j == cycle - 1
                    inb = pair[j & 1]
                    if last:      # the band only; its ends go to the neighbours' ghost rows of the other pair
                        L = self._resident_launch(inb, cycle * halo, False, plan, outb=other, wait=False)
                    elif j == 0:  # reads what the neighbours pushed
                        L = self._resident_launch(inb, halo, False, plan, ext=(cycle - 1) * halo, push=False,
                                                  outb=pair[1], signal=False)
                    else:         # no neighbour involved
                        L = self._resident_launch(inb, 0, False, plan, ext=(cycle - 1 - j) * halo, outb=pair[(j & 1) ^ 1])
                    descs[c * cycle + j] = L
            check(lib.b2i_cuda_run_resident_cycle(plan.handle, descs, 2 * cycle, passes, self.epoch))
        self._last_out = 2 if (passes // cycle) % 2 == 1 else 0
        self.epoch += passes + 1

    def run_resident(self, plan, passes: int, no_tma: bool = False, fused: bool = True) -> bool:
        """`passes` applications of the plan's chain to the resident bands, halo rows pushed between
        the GPUs from inside the kernels. fused=True: ONE cooperative launch that walks the passes
        itself (``b2i_cuda_run_resident_fused``: one separable 3-channel stencil, e.g. gauss5), else,
        or when the plan does not fit, one launch per pass. Returns whether the fused launch ran."""
        halo = plan.halo
        if halo > G:
            raise ValueError(f"resident passes support a halo of at most {G} rows per pass")
        even, odd = self._resident_launch(0, halo, no_tma, plan), self._resident_launch(1, halo, no_tma, plan)
        self._last_out = None
        ran_fused = False
        if self.y1 > self.y0:
            if fused and not no_tma:
                if getattr(self, "_task_flags", None) is None:
                    self._task_flags = torch.zeros(self.N_TASK_FLAGS, dtype=torch.int32, device=self.device)
                sf = not os.environ.get("B200IMG_NO_STRIP_FLAGS")
                ev, od = self._resident_launch(0, halo, no_tma, plan, sf), self._resident_launch(1, halo, no_tma, plan, sf)
                rc = lib.b2i_cuda_run_resident_fused(plan.handle, C.byref(ev), C.byref(od), passes, self.epoch,
                                                     self._task_flags.data_ptr(), self.N_TASK_FLAGS)
                if rc != _lib.B2I_ERR_UNSUPPORTED:
                    check(rc)
                    ran_fused = True
            if not ran_fused:
                check(lib.b2i_cuda_run_resident_steps(plan.handle, C.byref(even), C.byref(odd), passes, self.epoch))
        self.epoch += passes + 1
        return ran_fused

    def gather_resident(self, passes: int, dst_ptr_on_root: int, dst_pitch: int, out_c: int):
        g = self.rgeom
        check(lib.b2i_cuda_memcpy2d(dst_ptr_on_root + dst_pitch * self.y0, dst_pitch,
                                    self.band[passes & 1 if self._last_out is None else self._last_out].local + g["bp"] * G,
                                    g["bp"], g["W"] * out_c,
                                    self.y1 - self.y0, 2, self.stream()))
        self.barrier()


def _timed(fn, stream_sync=True):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dist.barrier()
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)          # max over ranks
    dist.barrier()
    return float(ms.item())


def _config_block(W, H, n):
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    spec = importlib.util.spec_from_file_location("_b2i_bench_common", os.path.join(root, "bench_common.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.config_block(W, H, n)


def _boundary_rows(H: int, bands, reach: int = 8):
    """Row ranges to verify: the top, the bottom and `reach` rows either side of every band boundary."""
    cuts = sorted({b[0] for b in bands if b[1] > b[0]} | {b[1] for b in bands if b[1] > b[0]})
    spans = [(0, min(H, 2 * reach)), (max(0, H - 2 * reach), H)]
    spans += [(max(0, c - reach), min(H, c + reach)) for c in cuts if 0 < c < H]
    return spans


def _check_rows(chain, img: np.ndarray, got: np.ndarray, spans) -> dict:
    """Bit-exact comparison of `got` with the host reference on the given row spans (rank 0)."""
    import mpi_cuda_imagemanipulation_b200 as b2
    bad, rows = 0, 0
    for a, b in spans:
        want = b2.host_run(chain, img, rows=(a, b))
        bad += int((got[a:b].reshape(b - a, -1) != want.reshape(b - a, -1)).sum())
        rows += b - a
    return {"ok": bad == 0, "rows_checked": rows, "spans": len(spans), "mismatching_bytes": bad}


def _resident_extra(rt, chain: str, W: int, H: int, img, checked: dict, key: str, p_check: int, p_time: int,
                    reuse_root: bool = False) -> dict:
    """Resident bands of a WxH image: p_check passes checked bit for bit around every band boundary
    (gathered to the root), then p_time passes per timed call: one cooperative launch that walks the
    passes where the chain allows it, and one launch per pass for comparison. Collective."""
    import mpi_cuda_imagemanipulation_b200 as b2
    from ..ops.cuda import Plan
    rank, world = rt.rank, rt.world
    try:
        plan = Plan(chain, 3)
        if not reuse_root:
            img = b2.synth_image(W, H, 3, seed=2) if rank == 0 else None
            rt.setup_root(W, H, 3, 3)
            rt.load_root(img)
        rt.setup_resident(W, H, 3)
        rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
        fused_ran = rt.run_resident(plan, p_check)
        rt.barrier()
        rt.gather_resident(p_check, rt.dst.ptrs[0], rt.geom["dp"], 3)
        got = rt.fetch_root()
        ok = True
        if rank == 0:
            bands = [band_rows(H, world, r) for r in range(world)]
            checked[key] = _check_rows(",".join([chain] * p_check), img, got, _boundary_rows(H, bands))
            checked[key]["fused_launch"] = bool(fused_ran)
            ok = checked[key]["ok"]
        del got
        out = {"passes_per_call": p_time, "halo_rows": plan.halo, "image": f"{W}x{H}", "chain": chain}
        for mode in (("fused", "per_pass_launches") if fused_ran else ("per_pass_launches",)):
            rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
            rt.run_resident(plan, p_time, fused=mode == "fused")
            rt.barrier()
            rr = [_timed(lambda: rt.run_resident(plan, p_time, fused=mode == "fused")) / p_time for _ in range(5)]
            out[mode + "_ms_per_pass"] = round(statistics.median(rr), 5)
        # halo rows exchanged every 2 / 4 passes only, the ghost rows between computed locally (run_resident_cycle):
        # checked like the others, then timed
        for cyc in (2, 4):
            if cyc * plan.halo > G or p_check % cyc or p_time % cyc or "," in chain:
                continue
            rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
            rt.run_resident_cycle(plan, p_check, cyc)
            rt.barrier()
            rt.gather_resident(p_check, rt.dst.ptrs[0], rt.geom["dp"], 3)
            got = rt.fetch_root()
            if rank == 0:
                c = _check_rows(",".join([chain] * p_check), img, got, _boundary_rows(H, bands))
                checked[key][f"exchange_every_{cyc}_ok"] = c["ok"]
                checked[key]["ok"] = ok = ok and c["ok"]
            del got
            rt.scatter_resident(rt.src.ptrs[0], rt.geom["sp"])
            rt.run_resident_cycle(plan, p_time, cyc)
            rt.barrier()
            rr = [_timed(lambda: rt.run_resident_cycle(plan, p_time, cyc)) / p_time for _ in range(5)]
            out[f"exchange_every_{cyc}_ms_per_pass"] = round(statistics.median(rr), 5)
        best = min(v for k, v in out.items() if k.endswith("_ms_per_pass"))
        out.update(ms_per_pass=best, mpix_per_s=round(W * H / 1e6 / (best / 1e3), 1), status=rt.status(), checked_ok=ok)
        if rank == 0 and not ok:
            out["error"] = f"result mismatch: {checked[key]}"
        return out
    except Exception as e:  # extras must not take the headline down
        return {"error": f"{type(e).__name__}: {e}"[:200]}


def bench_distributed(args, rank: int, world: int, local: int) -> Optional[dict]:
    """bench.py for N > 1: BASELINE.json's 8K scatter -> filter -> gather config, one rank per GPU."""
    import mpi_cuda_imagemanipulation_b200 as b2
    from ..ops.cuda import Plan
    from ..utils.clocks import ClockSampler

    rt = DistRuntime(root=0)
    W, H, K, Wm = args.width, args.height, args.steps, max(args.warmup, 3)
    chain = "kernel_cu"
    plan = Plan(chain, 3)
    img = b2.synth_image(W, H, 3, seed=1) if rank == 0 else None
    rt.setup_root(W, H, 3, 3)
    rt.load_root(img)
    # equal bands first (the reference's plan), for the record; then bands sized to the link
    rt.run_root(plan, Wm)
    rt.barrier()
    ms_equal = statistics.median([_timed(lambda: rt.run_root(plan, K)) / K for _ in range(3)])
    # what the equal-band steps (every rank's kernel loading from and storing into GPU 0) produced:
    # checked on rows around EVERY band boundary, before the root share changes the plan
    checked = {}
    equal_bands = list(rt.bands)
    got = rt.fetch_root()
    if rank == 0:
        checked["root_equal_bands"] = _check_rows(chain, img, got, _boundary_rows(H, equal_bands))
    share = rt.calibrate_root(plan)
    rt.run_root(plan, Wm)
    rt.barrier()
    got = rt.fetch_root()
    if rank == 0:
        checked["root_chosen_share"] = _check_rows(chain, img, got, _boundary_rows(H, rt.bands) + [(0, 64)])
        assert checked["root_chosen_share"]["ok"], f"bench output mismatch: {checked['root_chosen_share']}"
        assert checked["root_equal_bands"]["ok"], f"equal-band output mismatch: {checked['root_equal_bands']}"
    info = plan.info()
    # kernels launched per timed step, summed over ranks: a rank without rows launches nothing, and
    # the completion flags ride in the sweep kernel (no separate flag kernels)
    nl = torch.tensor([float(info["launches"]) if rt.y1 > rt.y0 else 0.0], device=rt.device)
    dist.all_reduce(nl, op=dist.ReduceOp.SUM)
    launches_per_step = int(nl.item())
    with ClockSampler(local) as clk:
        reps = [_timed(lambda: rt.run_root(plan, K)) / K for _ in range(args.repeats)]
        # keep the GPUs busy ~0.5 s so nvidia-smi samples clocks under this load. The repeat count
        # comes from the all-reduced timing, so every rank runs the same number of steps (the
        # epochs of the device flags must stay in lock-step across ranks).
        for _ in range(max(1, min(2000, int(500.0 / max(reps[-1] * K, 1e-3))))):
            rt.run_root(plan, K)
        rt.barrier()
    ms = statistics.median(reps)
    st = rt.status()
    assert st == 0, f"rank {rank}: a flag wait timed out (status {st})"

    # ---- end to end: root-owned pinned host image -> N GPUs over N PCIe links -> pinned host image ----
    rt.set_root_share(None)   # every rank has its own PCIe link: equal bands
    h_in = rt.shared_host("in", W * H * 3)
    h_out = rt.shared_host("out", W * H * 3)
    if rank == 0:
        h_in[:] = img.reshape(-1)
    rt.barrier()
    halo = plan.halo
    ra, rb = max(0, rt.y0 - halo), min(H, rt.y1 + halo)
    bp = W * 3
    # three frames in flight per rank: H2D, filter and D2H run on their own streams, ordered by events
    NS = 3
    d_in = [torch.empty((rb - ra) * bp + 64, dtype=torch.uint8, device=rt.device) for _ in range(NS)]
    d_out = [torch.empty((rt.y1 - rt.y0) * bp + 64, dtype=torch.uint8, device=rt.device) for _ in range(NS)]
    s_in, s_run, s_out = torch.cuda.Stream(rt.device), torch.cuda.Stream(rt.device), torch.cuda.Stream(rt.device)
    ev_in = [torch.cuda.Event() for _ in range(NS)]
    ev_run = [torch.cuda.Event() for _ in range(NS)]
    ev_out = [torch.cuda.Event() for _ in range(NS)]
    launches = []
    for k in range(NS):
        L = Launch()
        L.W, L.H, L.y0, L.y1 = W, H, rt.y0, rt.y1
        L.src = View(d_in[k].data_ptr(), bp, ra, rb)
        L.dst = View(d_out[k].data_ptr(), bp, rt.y0, rt.y1)
        L.stream = s_run.cuda_stream
        launches.append(L)

    def e2e_steps():
        main = torch.cuda.current_stream()
        fork = torch.cuda.Event()
        fork.record(main)
        for st in (s_in, s_run, s_out):
            st.wait_event(fork)
        for i in range(K):
            k = i % NS
            if i >= NS:
                s_in.wait_event(ev_run[k])       # the slot's previous frame has been filtered
                s_run.wait_event(ev_out[k])      # ... and its result copied out
            check(lib.b2i_cuda_memcpy2d(d_in[k].data_ptr(), bp, h_in.ctypes.data + ra * bp, bp, bp, rb - ra, 0,
                                        s_in.cuda_stream))
            ev_in[k].record(s_in)
            s_run.wait_event(ev_in[k])
            check(lib.b2i_cuda_run(plan.handle, C.byref(launches[k])))
            ev_run[k].record(s_run)
            s_out.wait_event(ev_run[k])
            check(lib.b2i_cuda_memcpy2d(h_out.ctypes.data + rt.y0 * bp, bp, d_out[k].data_ptr(), bp, bp,
                                        rt.y1 - rt.y0, 1, s_out.cuda_stream))
            ev_out[k].record(s_out)
        for st in (s_in, s_run, s_out):
            join = torch.cuda.Event()
            join.record(st)
            main.wait_event(join)

    e2e_steps()
    e2e_ms = statistics.median([_timed(e2e_steps) / K for _ in range(3)])
    rt.barrier()
    if rank == 0:   # every rank wrote its own rows of the shared host image: all boundaries, and rows inside every band
        spans = _boundary_rows(H, rt.bands) + [((b[0] + b[1]) // 2, (b[0] + b[1]) // 2 + 4) for b in rt.bands if b[1] - b[0] > 8]
        checked["e2e_every_rank"] = _check_rows(chain, img, h_out.reshape(H, W, 3), spans)
        assert checked["e2e_every_rank"]["ok"], f"e2e output mismatch: {checked['e2e_every_rank']}"

    # ---- extras: resident bands, halo rows pushed from inside the kernels (BASELINE.json configs 3 and 4) ----
    extras = {}
    extras["gauss5_resident"] = _resident_extra(rt, "gauss5", W, H, img, checked, "gauss5_resident",
                                                p_check=8, p_time=64, reuse_root=True)
    if os.environ.get("B200IMG_BENCH_CONFIG4", "1") != "0":
        extras["blur_sobel_16k_resident"] = _resident_extra(rt, "gauss5,sobel", 16384, 16384, None, checked,
                                                            "blur_sobel_16k_resident", p_check=2, p_time=8)

    mp = W * H / 1e6
    res = None
    if rank == 0:
        res = {
            "metric": "megapixels/sec (device-timed, max over ranks)",
            "value": round(mp / (ms / 1e3), 1), "unit": "MP/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": round(ms, 6), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "uint8", "data": "synthetic",
            "config": _config_block(W, H, world),
            "detail": {"mode": "root-owned input and output on GPU 0; scatter and gather fused into the filter "
                               "kernel over NVLink peer memory (IPC), done flags acquired on the root each step",
                       "bands": f"root GPU takes {share:.3f} of the rows, the other {world - 1} split the rest: the share is "
                                f"the fastest of the link model's (root's local rate vs the others' combined rate through "
                                f"the root's one NVLink port) and a few larger ones, timed as whole steps; 1.000 means the "
                                f"link cannot pay for itself on this filter and the root keeps the image",
                       "equal_bands_ms_per_step": round(ms_equal, 6),
                       "equal_bands_of_link_roofline": round((world - 1) / world * W * H * 3 / 770e9 * 1e3 / ms_equal, 3),
                       "link_roofline": "(N-1)/N of the 99.5 MB image leaves GPU 0 and as much enters it, per direction at the "
                                        "770 GB/s of a peer copy; the fused scatter / gather (TMA over NVLink) reaches the "
                                        "fraction above, which is why the calibrated plan keeps the image on the root",
                       "l2": "input 99.5 MB + output 99.5 MB per step; remote ranks bypass the root's L2 for peer traffic",
                       "timed_as": f"{K} steps per replay from C, CUDA events per rank, max over ranks, median of {args.repeats}"},
            "clocks": clk.summary(),
            "e2e": {"value": round(mp / (e2e_ms / 1e3), 1), "unit": "MP/s", "ms_per_step": round(e2e_ms, 4),
                    "h2d_bytes_per_step": W * H * 3 + (world - 1) * 2 * halo * W * 3, "d2h_bytes_per_step": W * H * 3,
                    "how": "root-owned pinned host image in POSIX shared memory; every rank copies its band (+halo) "
                           "over its own PCIe link, filters it and copies its rows back; 3 frames in flight per "
                           "rank on copy/compute/copy streams; CUDA events, max over ranks"},
            "gpu_launches": K * launches_per_step,
            "kernel": {"tile_h": info["tile_h"], "grid": info["grid"], "tma_load": info["tma_load"],
                       "tma_store": info["tma_store"]},
            "extras": extras, "impl": "ours", "checked": checked,
            # machine-readable form of what `config.bands` says in words: with root_share 1.0 the
            # other ranks only publish their done flags and the headline is a one-GPU number
            "root_share": round(float(share), 3),
            "effective_gpus": 1 if share >= 0.9995 else world,
        }
    rt.close()
    return res
