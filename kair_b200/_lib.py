"""ctypes binding of the C-ABI library (include/kair_b200.h).

The library is built in-tree by kair_b200/build.py (nvcc, sm_100a).  There is no fallback: if the shared object is
missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkair_b200.so")

c_i32 = ctypes.c_int32
c_i64 = ctypes.c_int64
c_f32 = ctypes.c_float
c_ptr = ctypes.c_void_p


class GemmTnArgs(ctypes.Structure):
    _fields_ = [
        ("A", c_ptr), ("a_pitch", c_i64),
        ("B", c_ptr), ("b_pitch", c_i64),
        ("M", c_i32), ("N", c_i32), ("K", c_i32), ("bn", c_i32),
        ("mode", c_i32), ("act", c_i32), ("act_slope", c_f32), ("ln_eps", c_f32),
        ("bias", c_ptr), ("resid", c_ptr), ("resid_pitch", c_i64),
        ("aux", c_ptr), ("aux_pitch", c_i64),
        ("rowstat", c_ptr),
        ("outB", c_ptr), ("outB_pitch", c_i64),
        ("outB2", c_ptr), ("outB2_pitch", c_i64),
        ("outF", c_ptr), ("outF_pitch", c_i64),
        ("n_valid", c_i32), ("ones_col", c_i32), ("conv", c_i32),
        ("batch", c_i32), ("H", c_i32), ("W", c_i32),
    ]


class GemmNtArgs(ctypes.Structure):
    _fields_ = [
        ("A", c_ptr), ("a_pitch", c_i64),
        ("B", c_ptr), ("b_pitch", c_i64),
        ("D", c_ptr), ("d_pitch", c_i64),
        ("T", c_i32), ("n1", c_i32), ("n2", c_i32), ("bn", c_i32), ("splits", c_i32),
        ("conv", c_i32), ("batch", c_i32), ("H", c_i32), ("W", c_i32),
    ]


class WindowAttnArgs(ctypes.Structure):
    _fields_ = [
        ("qkv", c_ptr), ("qkv_pitch", c_i64),
        ("out", c_ptr), ("out_pitch", c_i64),
        ("table", c_ptr),
        ("dout", c_ptr), ("dout_pitch", c_i64),
        ("dqkv", c_ptr), ("dqkv_pitch", c_i64),
        ("dtable", c_ptr),
        ("batch", c_i32), ("H", c_i32), ("W", c_i32),
        ("num_heads", c_i32), ("window", c_i32), ("shift", c_i32),
    ]


# name -> (restype, argtypes); every symbol of include/kair_b200.h must be listed here (tests check it)
SIGNATURES = {
    "kair_last_error": (ctypes.c_char_p, []),
    "kair_version": (ctypes.c_int, []),
    "kair_check_device": (ctypes.c_int, []),
    "kair_gemm_tn": (ctypes.c_int, [ctypes.POINTER(GemmTnArgs), c_ptr]),
    "kair_gemm_nt": (ctypes.c_int, [ctypes.POINTER(GemmNtArgs), c_ptr]),
    "kair_window_attn_fwd": (ctypes.c_int, [ctypes.POINTER(WindowAttnArgs), c_ptr]),
    "kair_window_attn_bwd": (ctypes.c_int, [ctypes.POINTER(WindowAttnArgs), c_ptr]),
}

_lib = None
_lock = threading.Lock()


def load():
    """Load (once) and return the ctypes handle; raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "kair_b200: %s not found. Build it with `python -m kair_b200.build` (needs nvcc); "
                "there is no CPU / PyTorch fallback." % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc, what):
    if rc != 0:
        msg = load().kair_last_error()
        raise RuntimeError("kair_b200.%s failed (rc=%d): %s" % (what, rc, msg.decode() if msg else "?"))
