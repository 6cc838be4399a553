"""ctypes binding of libadlm_head.so (include/adlm_head.h).

The library is the product: if it is missing or does not load, importing the compute path fails
loudly — there is no CPU or stock-torch fallback (BASELINE.json north_star).
"""
import ctypes
import os
from ctypes import c_char_p, c_double, c_float, c_int, c_longlong, c_size_t, c_ulonglong, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libadlm_head.so")

ACT_LOG, ACT_LINEAR = 0, 1
PREC_FP32, PREC_TF32, PREC_TF32X3 = 0, 1, 2
PREC = {"fp32": PREC_FP32, "tf32": PREC_TF32, "tf32x3": PREC_TF32X3}
E_INVALID, E_UNSUPPORTED, E_CUDA, E_WORKSPACE = -1, -2, -3, -4
MAX_PROTOS_PER_CLASS = 16
MAX_PUSH_CLASSES = 32
STEP_NSCALARS = 8
PUSH_NONE = 0x7FFFFFFFFFFFFFFF
ABI_VERSION = 9
EVAL_TILE_X, EVAL_TILE_Y = 64, 16
DP_MAX_WORLD = 16

# name -> (restype, argtypes); kept next to include/adlm_head.h (tests check they agree)
SIGNATURES = {
    "adlm_last_error": (c_char_p, []),
    "adlm_abi_version": (c_int, []),
    "adlm_launch_count": (c_ulonglong, []),
    "adlm_timing_begin": (c_int, [c_int, c_int]),
    "adlm_timing_end": (c_int, [c_void_p]),
    "adlm_head_fwd": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                              c_float, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "adlm_head_fwd_big_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "adlm_head_fwd_big": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                  c_float, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "adlm_head_bwd": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                              c_float, c_int, c_int, c_void_p, c_void_p, c_void_p,
                              c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "adlm_head_bwd_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "adlm_step_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "adlm_step_fwd_bwd": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                  c_float, c_int, c_void_p, c_void_p, c_int,
                                  c_float, c_float, c_float, c_float, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_size_t, c_void_p]),
    "adlm_step_counts": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_int,
                                 c_void_p, c_void_p, c_size_t, c_void_p]),
    "adlm_step_counts_noalloc": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_int,
                                         c_void_p, c_void_p]),
    "adlm_patch_class_mask": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                      c_void_p, c_void_p]),
    "adlm_push_scan": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p,
                               c_longlong, c_int, c_void_p, c_void_p]),
    "adlm_push_gather": (c_int, [c_void_p, c_int, c_int, c_int, c_longlong, c_void_p, c_int,
                                 c_void_p, c_void_p]),
}



class DpComm(ctypes.Structure):
    """adlm_dp_comm (include/adlm_head.h)"""
    _fields_ = [("rank", c_int), ("world", c_int), ("seq", c_ulonglong), ("peers", ctypes.POINTER(c_void_p)),
                ("timeout_ms", ctypes.c_uint)]


SIGNATURES["adlm_dp_exchange_bytes"] = (c_size_t, [c_int, c_int, c_int, c_int])
SIGNATURES["adlm_dp_exchange_layout"] = (c_int, [c_int, c_int, c_int, c_int, c_void_p])
SIGNATURES["adlm_step_fwd_bwd_dp"] = (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                              c_float, c_int, c_void_p, c_void_p, c_int,
                                              c_float, c_float, c_float, c_float,
                                              c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_size_t, ctypes.POINTER(DpComm), c_void_p])

SIGNATURES["adlm_head_fwd_pre"] = (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                           c_float, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                           c_void_p])
SIGNATURES["adlm_step_fwd_bwd_pre"] = (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_int,
                                               c_float, c_int, c_void_p, c_void_p, c_int,
                                               c_float, c_float, c_float, c_float, c_void_p,
                                               c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                               c_void_p, c_void_p, c_size_t, c_void_p])

SIGNATURES["adlm_eval_upsample"] = (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                            c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                            c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p])

SIGNATURES["adlm_push_scan_images"] = (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p,
                                               c_int, c_void_p, c_void_p])
SIGNATURES["adlm_min_pool_fwd"] = (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int,
                                           c_void_p, c_void_p, c_void_p, c_void_p])
SIGNATURES["adlm_min_pool_bwd"] = (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int,
                                           c_void_p, c_void_p, c_void_p, c_void_p, c_void_p])

_lib = None


class AdlmError(RuntimeError):
    pass


def load():
    """Load libadlm_head.so once.  Raises if it is not built — never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is not built. Run `python -m adlm_inherently_interpretable_segmentation_b200.build` "
            "(needs nvcc); the prototype head has no CPU / stock-torch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    got = lib.adlm_abi_version()
    if got != ABI_VERSION:
        raise ImportError(f"libadlm_head.so ABI version {got}, host expects {ABI_VERSION}: rebuild")
    _lib = lib
    return lib


def last_error():
    return load().adlm_last_error().decode("utf-8", "replace")


def check(rc, what):
    if rc == 0:
        return
    msg = f"{what}: {last_error()} (code {rc})"
    if rc == E_UNSUPPORTED:
        raise NotImplementedError(msg)
    if rc in (E_INVALID, E_WORKSPACE):
        raise ValueError(msg)
    raise AdlmError(msg)


K_HEAD_FWD, K_HEAD_BWD, K_PUSH_SCAN, K_PATCH_MASK, K_KLD, K_STEP_PREPARE = 1, 2, 3, 4, 5, 6


class kernel_timing:
    """with kernel_timing(tag, max_samples) as t: ...; t.ms -> list of per-launch milliseconds"""

    def __init__(self, tag, max_samples):
        self.tag, self.cap, self.ms = tag, max_samples, []

    def __enter__(self):
        check(load().adlm_timing_begin(self.tag, self.cap), "adlm_timing_begin")
        return self

    def __exit__(self, *exc):
        buf = (c_float * self.cap)()
        n = load().adlm_timing_end(ctypes.cast(buf, c_void_p))
        if n < 0:
            check(n, "adlm_timing_end")
        self.ms = [buf[i] for i in range(n)]
        return False


def launch_count():
    return int(load().adlm_launch_count())


def ptr(t):
    """device pointer of a tensor, or NULL"""
    return None if t is None else c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return c_void_p(torch.cuda.current_stream().cuda_stream)
