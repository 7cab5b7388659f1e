"""Thin tensor-level wrappers over the C-ABI kernels.  PyTorch supplies device memory and streams only."""
import ctypes

import torch

from . import _lib

EPI_LINEAR, EPI_GELU, EPI_RESID_LN, EPI_DGELU, EPI_LNBWD = range(5)
ACT_NONE, ACT_LRELU = 0, 1


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _req(t, dtype, name):
    if t is None:
        return
    if not t.is_cuda:
        raise RuntimeError("kair_b200: %s must be a CUDA tensor (no CPU fallback)" % name)
    if t.dtype != dtype:
        raise RuntimeError("kair_b200: %s must be %s, got %s" % (name, dtype, t.dtype))
    if t.stride(-1) != 1:
        raise RuntimeError("kair_b200: %s must be contiguous in its last dimension" % name)


def _pitch(t):
    """row pitch in elements of a [rows, cols] (or NHWC) tensor whose leading dims are densely packed"""
    return 0 if t is None else t.stride(-2)


def gemm_tn(A, B, *, N, K, bn, mode=EPI_LINEAR, act=ACT_NONE, act_slope=0.0, ln_eps=1e-5, bias=None, resid=None,
            aux=None, rowstat=None, outB=None, outB2=None, outF=None, n_valid=0, ones_col=-1, conv=None):
    """D = epilogue(A @ B^T).  A: bf16 [M, >=K] (or NHWC [B,H,W,>=K] with conv=(B,H,W)); B: bf16 [N, taps*K]."""
    _req(A, torch.bfloat16, "A"); _req(B, torch.bfloat16, "B")
    _req(bias, torch.float32, "bias"); _req(resid, torch.float32, "resid")
    _req(aux, torch.bfloat16, "aux"); _req(rowstat, torch.float32, "rowstat")
    _req(outB, torch.bfloat16, "outB"); _req(outB2, torch.bfloat16, "outB2"); _req(outF, torch.float32, "outF")
    a = _lib.GemmTnArgs()
    a.A = _ptr(A); a.a_pitch = _pitch(A)
    a.B = _ptr(B); a.b_pitch = _pitch(B)
    if conv is not None:
        a.conv = 1
        a.batch, a.H, a.W = conv
        a.M = conv[0] * conv[1] * conv[2]
    else:
        a.conv = 0
        a.M = A.shape[0]
    a.N = N; a.K = K; a.bn = bn
    a.mode = mode; a.act = act; a.act_slope = act_slope; a.ln_eps = ln_eps
    a.bias = _ptr(bias)
    a.resid = _ptr(resid); a.resid_pitch = _pitch(resid)
    a.aux = _ptr(aux); a.aux_pitch = _pitch(aux)
    a.rowstat = _ptr(rowstat)
    a.outB = _ptr(outB); a.outB_pitch = _pitch(outB)
    a.outB2 = _ptr(outB2); a.outB2_pitch = _pitch(outB2)
    a.outF = _ptr(outF); a.outF_pitch = _pitch(outF)
    a.n_valid = n_valid; a.ones_col = ones_col
    _lib.check(_lib.load().kair_gemm_tn(ctypes.byref(a), _stream()), "gemm_tn")


def gemm_nt(A, B, D, *, n1, n2, bn, splits=0, conv=None):
    """D[n1, taps*n2] += A[:, :n1]^T @ B[:, :n2] (token contraction).  D is fp32 and must be pre-zeroed."""
    _req(A, torch.bfloat16, "A"); _req(B, torch.bfloat16, "B"); _req(D, torch.float32, "D")
    a = _lib.GemmNtArgs()
    a.A = _ptr(A); a.a_pitch = _pitch(A)
    a.B = _ptr(B); a.b_pitch = _pitch(B)
    a.D = _ptr(D); a.d_pitch = _pitch(D)
    a.n1 = n1; a.n2 = n2; a.bn = bn; a.splits = splits
    if conv is not None:
        a.conv = 1
        a.batch, a.H, a.W = conv
        a.T = conv[0] * conv[1] * conv[2]
    else:
        a.conv = 0
        a.T = A.shape[0]
    _lib.check(_lib.load().kair_gemm_nt(ctypes.byref(a), _stream()), "gemm_nt")


def _attn_args(qkv, table, batch, H, W, num_heads, window, shift):
    _req(qkv, torch.bfloat16, "qkv"); _req(table, torch.float32, "table")
    if not table.is_contiguous():
        raise RuntimeError("kair_b200: relative_position_bias_table must be contiguous")
    a = _lib.WindowAttnArgs()
    a.qkv = _ptr(qkv); a.qkv_pitch = _pitch(qkv)
    a.table = _ptr(table)
    a.batch, a.H, a.W = batch, H, W
    a.num_heads, a.window, a.shift = num_heads, window, shift
    return a


def window_attn_fwd(qkv, table, out, *, batch, H, W, num_heads, window=8, shift=0):
    """out[t, h*32:(h+1)*32] = windowed softmax attention; qkv: bf16 [B*H*W, >=3*nH*32] in [3][nH][32] layout."""
    _req(out, torch.bfloat16, "out")
    a = _attn_args(qkv, table, batch, H, W, num_heads, window, shift)
    a.out = _ptr(out); a.out_pitch = _pitch(out)
    _lib.check(_lib.load().kair_window_attn_fwd(ctypes.byref(a), _stream()), "window_attn_fwd")


def window_attn_bwd(qkv, table, dout, dqkv, dtable, *, batch, H, W, num_heads, window=8, shift=0):
    """Backward of window_attn_fwd: writes dqkv (bf16) and accumulates into dtable (fp32, caller zeroes)."""
    _req(dout, torch.bfloat16, "dout"); _req(dqkv, torch.bfloat16, "dqkv"); _req(dtable, torch.float32, "dtable")
    a = _attn_args(qkv, table, batch, H, W, num_heads, window, shift)
    a.dout = _ptr(dout); a.dout_pitch = _pitch(dout)
    a.dqkv = _ptr(dqkv); a.dqkv_pitch = _pitch(dqkv)
    a.dtable = _ptr(dtable)
    _lib.check(_lib.load().kair_window_attn_bwd(ctypes.byref(a), _stream()), "window_attn_bwd")
