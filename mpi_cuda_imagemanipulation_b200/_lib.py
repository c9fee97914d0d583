"""ctypes binding of libb200img.so (the C ABI in include/b200img.h).

The library is built in-tree by ``make`` (or ``__graft_entry__.build()``) and loaded from
``mpi_cuda_imagemanipulation_b200/lib``. There is no pure-Python fallback for the compute path:
if the shared object is missing the import fails loudly, so a GPU box can never silently run a
slower stand-in.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(_HERE)
# B200IMG_LIB points at another build of the library (A/B measurements, system-wide installs)
LIB_PATH = os.environ.get("B200IMG_LIB") or os.path.join(_HERE, "lib", "libb200img.so")

B2I_MAX_OPS = 16
B2I_MAX_GPUS = 16
B2I_ERR_UNSUPPORTED = -4
B2I_ERR_TIMEOUT = -6

# b2i_status
OK, ERR_ARG, ERR_IO, ERR_CUDA, ERR_UNSUPPORTED, ERR_NO_DEVICE, ERR_TIMEOUT, ERR_NOMEM = (
    0, -1, -2, -3, -4, -5, -6, -7)
# b2i_opkind
OP_GRAY_REF, OP_GRAY_CV, OP_LUT, OP_STENCIL, OP_SOBEL, OP_EXPAND = 1, 2, 3, 4, 5, 6
# b2i_border
BORDER_REF, BORDER_REFLECT101, BORDER_REPLICATE, BORDER_ZERO = 0, 1, 2, 3
BORDERS = {"ref": 0, "reflect101": 1, "reflect": 1, "replicate": 2, "zero": 3, None: -1, "": -1,
           "default": -1}
# b2i_order
ORDER_BGR, ORDER_RGB = 0, 1
ORDERS = {"bgr": 0, "rgb": 1}


class Op(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("border", C.c_int32), ("ksize", C.c_int32), ("shift", C.c_int32),
        ("bias", C.c_int32), ("reserved", C.c_int32),
        ("weights", C.c_int8 * 32), ("lut", C.c_uint8 * 256), ("name", C.c_char * 24),
    ]


class ChainStruct(C.Structure):
    _fields_ = [("n_ops", C.c_int32), ("order", C.c_int32), ("ops", Op * B2I_MAX_OPS)]


class B200ImgError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b200img error {code}: {msg}")
        self.code = code


def build(verbose: bool = False) -> None:
    """Runs ``make`` at the repo root (library + CLI, sm_100a)."""
    res = subprocess.run(["make", "-C", _REPO, "-j8"], capture_output=not verbose, text=True)
    if res.returncode != 0:
        raise RuntimeError("building libb200img.so failed:\n" + (res.stdout or "") + (res.stderr or ""))


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        if os.environ.get("B200IMG_NO_AUTOBUILD"):
            raise ImportError(f"{LIB_PATH} is missing; run `make` at the repo root")
        build()
    return C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)


lib = _load()
lib.b2i_last_error.restype = C.c_char_p
lib.b2i_version.restype = C.c_char_p


def check(rc: int) -> int:
    if rc < 0:
        raise B200ImgError(rc, lib.b2i_last_error().decode("utf-8", "replace"))
    return rc


def _sig(name, restype, *argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


_P = C.POINTER
u8p = C.c_void_p  # raw addresses: numpy .ctypes.data / torch .data_ptr()

_sig("b2i_chain_parse", C.c_int, C.c_char_p, C.c_int, C.c_int, _P(ChainStruct))
_sig("b2i_chain_preset", C.c_int, C.c_char_p, C.c_int, _P(ChainStruct))
_sig("b2i_chain_out_channels", C.c_int, _P(ChainStruct), C.c_int)
_sig("b2i_chain_halo", C.c_int, _P(ChainStruct))
_sig("b2i_chain_describe", C.c_int, _P(ChainStruct), C.c_char_p, C.c_size_t)
_sig("b2i_lut_contrast_ref", None, C.c_float, u8p)
_sig("b2i_lut_contrast_cv", None, C.c_double, u8p)
_sig("b2i_band_rows", C.c_int, C.c_int, C.c_int, C.c_int, _P(C.c_int), _P(C.c_int))
_sig("b2i_host_run", C.c_int, _P(ChainStruct), u8p, C.c_int, C.c_int, C.c_int, C.c_size_t, u8p,
     C.c_size_t)
_sig("b2i_host_run_rows", C.c_int, _P(ChainStruct), u8p, C.c_int, C.c_int, C.c_int, C.c_size_t,
     C.c_int, C.c_int, u8p, C.c_size_t)
_sig("b2i_host_run_ranks", C.c_int, _P(ChainStruct), u8p, C.c_int, C.c_int, C.c_int, C.c_size_t,
     u8p, C.c_size_t, C.c_int, C.c_int, _P(C.c_double))
_sig("b2i_read_pnm", C.c_int, C.c_char_p, C.c_int, _P(C.c_void_p), _P(C.c_int), _P(C.c_int),
     _P(C.c_int))
_sig("b2i_write_pnm", C.c_int, C.c_char_p, C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_size_t)
_sig("b2i_jpeg_available", C.c_int)
_sig("b2i_read_image", C.c_int, C.c_char_p, C.c_int, _P(C.c_void_p), _P(C.c_int), _P(C.c_int), _P(C.c_int))
_sig("b2i_write_image", C.c_int, C.c_char_p, C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_int)
_sig("b2i_read_raw", C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int, _P(C.c_void_p))
_sig("b2i_write_raw", C.c_int, C.c_char_p, u8p, C.c_int, C.c_int, C.c_int, C.c_size_t)
_sig("b2i_free", None, C.c_void_p)
_sig("b2i_synth_image", C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_uint32)


def has(symbol: str) -> bool:
    return hasattr(lib, symbol)


# ---- CUDA path (present once the kernels are linked in) ----------------------------------------
LAUNCH_NO_TMA, LAUNCH_NAIVE, LAUNCH_DST_MCAST, LAUNCH_STRIP_FLAGS = 1, 2, 4, 8


class View(C.Structure):
    _fields_ = [("ptr", C.c_uint64), ("pitch", C.c_int64), ("y_begin", C.c_int32),
                ("y_end", C.c_int32)]


class Launch(C.Structure):
    _fields_ = [
        ("W", C.c_int32), ("H", C.c_int32), ("y0", C.c_int32), ("y1", C.c_int32),
        ("src", View), ("dst", View), ("tmp", View * 2),
        ("push_up", View), ("push_dn", View), ("push_rows", C.c_int32),
        ("wait_epoch", C.c_uint32), ("signal_epoch", C.c_uint32),
        ("wait_up_flag", C.c_uint64), ("wait_dn_flag", C.c_uint64),
        ("signal_up_flag", C.c_uint64), ("signal_dn_flag", C.c_uint64),
        ("done_flag", C.c_uint64), ("counters", C.c_uint64), ("status", C.c_uint64),
        ("timeout_ns", C.c_uint64), ("stream", C.c_void_p),
        ("tile_h", C.c_int32), ("flags", C.c_int32),
        ("wait_done_flags", C.c_uint64), ("wait_done_mask", C.c_uint32), ("reserved0", C.c_uint32),
    ]


class PlanInfo(C.Structure):
    _fields_ = [
        ("cin", C.c_int32), ("cout", C.c_int32), ("halo", C.c_int32), ("n_segments", C.c_int32),
        ("seg_pointwise", C.c_int32 * B2I_MAX_OPS), ("seg_halo", C.c_int32 * B2I_MAX_OPS),
        ("seg_phases", C.c_int32 * B2I_MAX_OPS), ("seg_cout", C.c_int32 * B2I_MAX_OPS),
        ("last_tile_h", C.c_int32), ("last_smem_bytes", C.c_int32), ("last_grid", C.c_int32),
        ("last_blocks_per_sm", C.c_int32), ("last_tma_load", C.c_int32),
        ("last_tma_store", C.c_int32), ("last_launches", C.c_int32),
    ]


_sig("b2i_plan_create", C.c_int, _P(ChainStruct), C.c_int, _P(C.c_void_p))
_sig("b2i_plan_destroy", None, C.c_void_p)
_sig("b2i_plan_halo", C.c_int, C.c_void_p)
_sig("b2i_plan_info", C.c_int, C.c_void_p, _P(PlanInfo))
_sig("b2i_cuda_run", C.c_int, C.c_void_p, _P(Launch))
_sig("b2i_cuda_run_image", C.c_int, C.c_void_p, C.c_uint64, C.c_int64, C.c_uint64, C.c_int64,
     C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int)
_sig("b2i_cuda_device_count", C.c_int)

# ---- multi-GPU runtime ---------------------------------------------------------------------------
_sig("b2i_topology_json", C.c_int, C.c_char_p, C.c_size_t)
_sig("b2i_ctx_create", C.c_int, C.c_int, _P(C.c_int), C.c_int, _P(C.c_void_p))
_sig("b2i_ctx_destroy", None, C.c_void_p)
_sig("b2i_ctx_gpus", C.c_int, C.c_void_p)
_sig("b2i_ctx_load_root", C.c_int, C.c_void_p, u8p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_int)
_sig("b2i_ctx_set_root_share", C.c_int, C.c_void_p, C.c_double)
_sig("b2i_ctx_fetch_root", C.c_int, C.c_void_p, u8p, C.c_size_t)
_sig("b2i_ctx_set_replicated", C.c_int, C.c_void_p, C.c_int)
_sig("b2i_ctx_replicated", C.c_int, C.c_void_p)
_sig("b2i_ctx_fetch_replica", C.c_int, C.c_void_p, C.c_int, u8p, C.c_size_t)
_sig("b2i_mcast_create", C.c_int, C.c_int, _P(C.c_int), C.c_size_t, _P(C.c_void_p))
_sig("b2i_mcast_destroy", None, C.c_void_p)
_sig("b2i_mcast_ptr", C.c_uint64, C.c_void_p)
_sig("b2i_mcast_replica", C.c_uint64, C.c_void_p, C.c_int)
_sig("b2i_mcast_bytes", C.c_size_t, C.c_void_p)
_sig("b2i_mcast_is_multicast", C.c_int, C.c_void_p)
_sig("b2i_ctx_run_root", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, _P(C.c_double),
     _P(C.c_double))
_sig("b2i_ctx_run_resident", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
     _P(C.c_double), _P(C.c_double))
_sig("b2i_ctx_run_direct", C.c_int, C.c_void_p, C.c_void_p, u8p, C.c_int, C.c_int, C.c_int,
     C.c_size_t, u8p, C.c_size_t, C.c_int, _P(C.c_double))
_sig("b2i_cuda_run_host_image", C.c_int, _P(ChainStruct), u8p, C.c_int, C.c_int, C.c_int,
     C.c_size_t, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, _P(C.c_double), _P(C.c_double))

NOALLOC_MAX_PIXELS = 2 * 1024 * 1024
_sig("b2i_cuda_cluster_probe", C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float),
     C.POINTER(C.c_float), C.POINTER(C.c_longlong))
_sig("b2i_cuda_run_image_noalloc", C.c_int, C.c_void_p, C.c_uint64, C.c_int64, C.c_uint64, C.c_int64,
     C.c_int, C.c_int, C.c_void_p, C.c_int)

# ---- multi-process plumbing ------------------------------------------------------------------------
_sig("b2i_ipc_alloc", C.c_int, C.c_size_t, _P(C.c_uint64), C.c_char_p)
_sig("b2i_ipc_open", C.c_int, C.c_char_p, _P(C.c_uint64))
_sig("b2i_ipc_close", C.c_int, C.c_uint64)
_sig("b2i_dev_free", C.c_int, C.c_uint64)
_sig("b2i_host_register", C.c_int, C.c_void_p, C.c_size_t)
_sig("b2i_host_unregister", C.c_int, C.c_void_p)
_sig("b2i_cuda_wait_flags", C.c_int, _P(C.c_uint64), C.c_int, C.c_uint32, C.c_uint64, C.c_uint64,
     C.c_void_p)
_sig("b2i_cuda_set_flags", C.c_int, _P(C.c_uint64), C.c_int, C.c_uint32, C.c_void_p)
_sig("b2i_cuda_run_steps", C.c_int, C.c_void_p, _P(Launch), C.c_int, C.c_uint32, _P(C.c_uint64), C.c_int,
     C.c_uint64)
_sig("b2i_cuda_run_resident_steps", C.c_int, C.c_void_p, _P(Launch), _P(Launch), C.c_int, C.c_uint32)
_sig("b2i_cuda_run_resident_cycle", C.c_int, C.c_void_p, _P(Launch), C.c_int, C.c_int, C.c_uint32)
_sig("b2i_cuda_run_resident_fused", C.c_int, C.c_void_p, _P(Launch), _P(Launch), C.c_int, C.c_uint32, C.c_uint64,
     C.c_int)
_sig("b2i_cuda_memcpy2d", C.c_int, C.c_uint64, C.c_size_t, C.c_uint64, C.c_size_t, C.c_size_t, C.c_size_t,
     C.c_int, C.c_void_p)
