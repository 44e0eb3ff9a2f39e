"""ctypes binding of libdgs.so (declarations mirror include/dgs.h one to one).

There is no fallback: if the shared object is missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import pathlib
import threading
from typing import Any, Optional

LIB_PATH = pathlib.Path(__file__).resolve().parent / "csrc" / "libdgs.so"

# element types (include/dgs.h)
F32, BF16, F16, F64, I64 = 0, 1, 2, 3, 4
READ_ACC, WRITE_WIRE = 1, 2
OPT_MAXIMIZE, ADAM_DECOUPLED_WD, SGD_NESTEROV, SGD_FIRST_STEP = 1, 2, 4, 8
REDUCE_AVG, REDUCE_SUM, REDUCE_MAX, REDUCE_MIN = 1, 2, 3, 4
MAX_RANGES = 24
MAX_METRICS = 64
ALIGN_BYTES = 128
NORM_MAX_PARTIALS = 9472
NOALLOC_MAX_PARTIALS = 4096
NOALLOC_MAX_ELEMS = 4096 * 16384
DIRECT_MAX_SEGS = 512
PEER_MAX, PEER_SLOTS, PEER_HANDLE_BYTES, PEER_SYNC_BYTES, PEER_MAX_BUCKETS = 8, 4, 64, 512, 128


class Seg(ctypes.Structure):
    _fields_ = [("src", ctypes.c_void_p), ("dst_off", ctypes.c_int64), ("numel", ctypes.c_int64)]


class AdamRange(ctypes.Structure):
    _fields_ = [
        ("begin", ctypes.c_int64),
        ("end", ctypes.c_int64),
        ("lr", ctypes.c_double),
        ("beta1", ctypes.c_double),
        ("beta2", ctypes.c_double),
        ("eps", ctypes.c_double),
        ("weight_decay", ctypes.c_double),
        ("step_size", ctypes.c_double),
        ("bias_correction2_sqrt", ctypes.c_double),
        ("flags", ctypes.c_int32),
        ("_pad", ctypes.c_int32),
    ]


class SgdRange(ctypes.Structure):
    _fields_ = [
        ("begin", ctypes.c_int64),
        ("end", ctypes.c_int64),
        ("lr", ctypes.c_double),
        ("momentum", ctypes.c_double),
        ("dampening", ctypes.c_double),
        ("weight_decay", ctypes.c_double),
        ("flags", ctypes.c_int32),
        ("_pad", ctypes.c_int32),
    ]


class ScalarRef(ctypes.Structure):
    _fields_ = [("ptr", ctypes.c_void_p), ("dtype", ctypes.c_int32), ("_pad", ctypes.c_int32)]


_vp, _i, _i64, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float

# name -> argtypes; every function returns int status unless listed in _RESTYPES
SIGNATURES = {
    "dgs_version": [],
    "dgs_last_error": [],
    "dgs_launch_count": [],
    "dgs_flatten_accumulate": [ctypes.POINTER(Seg), _i, _i, _vp, _vp, _i, _i, _f, _vp],
    "dgs_scale_cast": [_vp, _vp, _i, _i64, _f, _vp],
    "dgs_l2norm_grid": [_i64],
    "dgs_l2norm_partial": [_vp, _i, _i64, _vp, _vp],
    "dgs_l2norm_final": [_vp, _i, _vp, _vp],
    "dgs_l2norm_final_sumsq": [_vp, _i, _vp, _vp],
    "dgs_sqrt_inplace": [_vp, _vp],
    "dgs_l2norm": [_vp, _i, _i64, _vp, _vp, _vp],
    "dgs_l2norm_noalloc": [_vp, _i, _i64, _vp, _vp],
    "dgs_clip_scale": [_vp, _i, _i64, _vp, _f, _vp],
    "dgs_step_adam": [_vp, _vp, _vp, _i, _vp, _i, _i64, ctypes.POINTER(AdamRange), _i, _vp, _f, _vp],
    "dgs_step_sgd": [_vp, _vp, _i, _vp, _i, _i64, ctypes.POINTER(SgdRange), _i, _vp, _f, _vp],
    "dgs_step_adam_scaled": [_vp, _vp, _vp, _i, _vp, _i, _i64, ctypes.POINTER(AdamRange), _i, _vp, _f, _vp, _vp],
    "dgs_step_sgd_scaled": [_vp, _vp, _i, _vp, _i, _i64, ctypes.POINTER(SgdRange), _i, _vp, _f, _vp, _vp],
    "dgs_direct_step_adam": [ctypes.POINTER(Seg), ctypes.POINTER(ctypes.c_int32), _i, _i, _vp, _f, _vp, _vp, _vp, _i,
                             ctypes.POINTER(AdamRange), _i, _i, _f, _vp, _i, _vp, _vp],
    "dgs_direct_step_sgd": [ctypes.POINTER(Seg), ctypes.POINTER(ctypes.c_int32), _i, _i, _vp, _f, _vp, _vp, _i,
                            ctypes.POINTER(SgdRange), _i, _i, _f, _vp, _i, _vp, _vp],
    "dgs_timer_enable": [_i],
    "dgs_timer_read": [_vp, _i, ctypes.POINTER(_i)],
    "dgs_metric_pack": [ctypes.POINTER(ScalarRef), _i, _vp, _vp, _vp],
    "dgs_metric_average": [_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "dgs_metric_average_packed": [_vp, _i64, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    "dgs_metric_reduce": [_vp, _vp, _i, _i, _i, _vp, _vp],
    "dgs_comm_unique_id": [_vp],
    "dgs_comm_init": [_vp, _i, _i, _i, ctypes.POINTER(_vp)],
    "dgs_comm_destroy": [_vp],
    "dgs_comm_rank": [_vp, ctypes.POINTER(_i), ctypes.POINTER(_i)],
    "dgs_comm_allreduce_async": [_vp, _vp, _i64, _i, _vp],
    "dgs_comm_reduce_scatter_async": [_vp, _vp, _i64, _i, _vp],
    "dgs_comm_broadcast_async": [_vp, _vp, _i64, _i, _vp],
    "dgs_comm_allgather_async": [_vp, _vp, _vp, _i64, _vp],
    "dgs_comm_l2norm_partial_async": [_vp, _vp, _i, _i64, _vp],
    "dgs_comm_wait": [_vp, _vp],
    "dgs_nccl_version": [ctypes.POINTER(_i)],
    "dgs_peer_export": [_vp, _vp, ctypes.POINTER(_i64)],
    "dgs_peer_open": [_vp, _i64, ctypes.POINTER(_vp)],
    "dgs_peer_close": [_vp],
    "dgs_peer_close_all": [],
    "dgs_peer_error": [ctypes.POINTER(_i)],
    "dgs_peer_error_clear": [],
    "dgs_peer_barrier": [ctypes.POINTER(_vp), _i, _i, _i, ctypes.c_uint32, ctypes.c_double, _vp],
    "dgs_flatten_push": [ctypes.POINTER(Seg), _i, _i, _vp, ctypes.POINTER(_vp), _i, _i, _f, _i64, _i64, _i, _i, _vp],
    "dgs_peer_reduce": [_vp, _i, _i64, _i, _i, _vp, _vp],
    "dgs_peer_reduce_multi": [_vp, _i, ctypes.POINTER(_i64), ctypes.POINTER(_i64), _i, _i, _i, _vp, _i, _vp],
    "dgs_peer_scatter": [ctypes.POINTER(_vp), ctypes.POINTER(_i64), ctypes.POINTER(_i64), _i, _i, _i, _vp],
    "dgs_peer_norm_exchange": [ctypes.POINTER(_vp), _i, _i, _i, ctypes.c_uint32, _vp, _i, _vp, ctypes.c_double, _vp],
    "dgs_step_adam_push": [_vp, _vp, _vp, _i, _vp, _i, _i64, ctypes.POINTER(AdamRange), _i, _vp, _f, ctypes.POINTER(_vp), _i, _i, _vp],
    "dgs_step_sgd_push": [_vp, _vp, _i, _vp, _i, _i64, ctypes.POINTER(SgdRange), _i, _vp, _f, ctypes.POINTER(_vp), _i, _i, _vp],
}
_RESTYPES = {"dgs_last_error": ctypes.c_char_p, "dgs_launch_count": ctypes.c_int64}
# functions whose int return is a value, not a status
_NOT_STATUS = {"dgs_version", "dgs_last_error", "dgs_launch_count", "dgs_l2norm_grid"}

_lock = threading.Lock()
_lib: Optional[ctypes.CDLL] = None


class DgsError(RuntimeError):
    pass


def load() -> ctypes.CDLL:
    """Load libdgs.so (once) and attach argtypes.  Raises if it was not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not LIB_PATH.exists():
            raise DgsError(
                f"{LIB_PATH} is missing: build it with `python -m determined_clone_b200._build` "
                "(there is no CPU or torch fallback for the grad-sync path)"
            )
        lib = ctypes.CDLL(str(LIB_PATH))
        for name, argtypes in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, ctypes.c_int)
        _lib = lib
        return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().dgs_last_error().decode("utf-8", "replace")
        raise DgsError(f"libdgs {what} failed with status {rc}: {msg}")


class _Calls:
    """`calls.dgs_xxx(...)` == load().dgs_xxx(...) with the status checked."""

    def __getattr__(self, name: str) -> Any:
        fn = getattr(load(), name)
        if name in _NOT_STATUS:
            return fn

        def checked(*args: Any) -> None:
            rc = fn(*args)
            if rc != 0:
                check(rc, name)

        checked.__name__ = name
        setattr(self, name, checked)
        return checked


calls = _Calls()


def dtype_code(torch_dtype: Any) -> int:
    import torch

    table = {
        torch.float32: F32,
        torch.bfloat16: BF16,
        torch.float16: F16,
        torch.float64: F64,
        torch.int64: I64,
    }
    try:
        return table[torch_dtype]
    except KeyError:
        raise DgsError(f"dtype {torch_dtype} is not supported by libdgs") from None
