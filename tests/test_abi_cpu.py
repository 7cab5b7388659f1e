"""The C-ABI library builds (nvcc cross-compiles without a GPU), loads, and exports every symbol that
include/kair_b200.h declares; struct layouts parsed from the header match the compiled sizes."""
import ctypes

from kair_b200 import _lib, build


def test_library_exports_every_declared_symbol():
    build.build()
    lib = _lib.load()
    assert _lib.SIGNATURES, "no prototypes parsed from the header"
    for name in _lib.SIGNATURES:
        assert hasattr(lib, name), name
    assert lib.kair_version() >= 1
    for must in ("kair_gemm_tn", "kair_gemm_nt", "kair_window_attn_fwd", "kair_window_attn_bwd", "kair_layernorm_fwd",
                 "kair_conv_in_fwd", "kair_conv_out_fwd", "kair_pack_linear", "kair_unpack_linear_grad"):
        assert must in _lib.SIGNATURES


def test_struct_sizes_match_the_compiled_library():
    lib = _lib.load()
    fn = getattr(lib, "kair_sizeof", None)
    if fn is None:
        return
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_char_p]
    for name, st in _lib.STRUCTS.items():
        n = fn(name.encode())
        if n > 0:
            assert n == ctypes.sizeof(st), name


def test_no_gpu_means_loud_error():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    assert lib.kair_check_device() != 0
