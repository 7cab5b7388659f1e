"""Thin tensor-level wrappers over the C-ABI kernels.  PyTorch supplies device memory and streams only."""
import ctypes
import os

import torch

from . import _lib

EPI_LINEAR, EPI_GELU, EPI_RESID_LN, EPI_DGELU, EPI_LNBWD = range(5)
ACT_NONE, ACT_LRELU = 0, 1

_S = _lib.STRUCTS


# KAIR_B200_TORCH_OPS=1: launch lists are replayed through torch.ops.kair_b200.run_list (kair_b200/torch_ext.py) instead
# of the ctypes call -- the second binding of the same C ABI
USE_TORCH_OPS = os.environ.get("KAIR_B200_TORCH_OPS") == "1"


class Launch:
    """One recorded kernel launch: a bound C entry point plus its filled argument struct.

    Engines build these once per plan (all pointers are static workspace / parameter addresses) and replay them
    every step, so the per-step host cost is one ctypes call per kernel."""

    __slots__ = ("name", "fn", "args", "keep", "meta")

    def __init__(self, name, args, keep=()):
        self.name = name
        self.fn = getattr(_lib.load(), name)
        self.args = args
        self.keep = keep  # tensors whose storage the struct points into
        self.meta = None  # optional accounting hints for kair_b200.profiling (e.g. {"terms": 3} on split-operand GEMMs)

    def run(self, stream):
        rc = self.fn(ctypes.byref(self.args), stream)
        if rc != 0:
            _lib.check(rc, self.name)


class _RawLaunch:
    """Launch of an entry point that takes scalars instead of an argument struct."""

    __slots__ = ("name", "call", "keep")

    def __init__(self, name, call, keep=()):
        self.name, self.call, self.keep = name, call, keep

    def run(self, stream):
        rc = self.call(stream)
        if rc != 0:
            _lib.check(rc, self.name)


class LaunchList:
    """A whole pass (forward or backward list of a plan) replayed by ONE FFI call: kair_run_list walks a C array of
    (entry point id, argument struct) pairs, so the per-pass host cost is one ctypes crossing instead of one per
    kernel.  The Launch objects stay referenced (they own the argument structs and the tensors those point into)."""

    __slots__ = ("launches", "array", "n", "fn")

    def __init__(self, launches):
        lib = _lib.load()
        self.launches = list(launches)
        self.n = len(self.launches)
        self.array = (_S["kair_launch"] * max(self.n, 1))()
        for i, l in enumerate(self.launches):
            op = lib.kair_op_id(l.name.encode())
            if op < 0:
                raise RuntimeError("kair_b200: %s cannot be part of a launch list" % l.name)
            self.array[i].op = op
            self.array[i].args = ctypes.cast(ctypes.pointer(l.args), ctypes.c_void_p)
        self.fn = lib.kair_run_list

    def run(self, stream):
        if USE_TORCH_OPS:   # same C entry point, reached through the torch C++ extension (current device / stream)
            return self.run_torch(torch.empty(0, device="cuda"))
        rc = self.fn(self.array, self.n, stream)
        if rc != 0:
            _lib.check(rc, "kair_run_list")

    def run_torch(self, anchor):
        """the same replay through the torch C++ extension (`torch.ops.kair_b200.run_list`, kair_b200/torch_ext.py) on
        the current stream of `anchor`'s device"""
        from . import torch_ext
        torch_ext.load().run_list(anchor, ctypes.addressof(self.array), self.n)

    def __len__(self):
        return self.n

    def __iter__(self):
        return iter(self.launches)


def make_list(launches):
    return LaunchList(launches)


def stream_handle(device=None):
    """the current stream of `device` (default: the current device) as the `void* stream` argument of the C ABI"""
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def device_guard(device):
    """context manager that makes `device` current: the launchers use the current device (cudaGetDevice) and stream"""
    return torch.cuda.device(device)


def zeros(shape, dtype, device):
    return torch.zeros(shape, dtype=dtype, device=device)


def _ptr(t):
    return None if t is None else t.data_ptr()


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _req(t, dtype, name):
    if t is None:
        return
    if not t.is_cuda:
        raise RuntimeError("kair_b200: %s must be a CUDA tensor (no CPU fallback)" % name)
    if t.dtype != dtype:
        raise RuntimeError("kair_b200: %s must be %s, got %s" % (name, dtype, t.dtype))
    if t.dim() > 0 and t.stride(-1) != 1:
        raise RuntimeError("kair_b200: %s must be contiguous in its last dimension" % name)


def _pitch(t):
    """row pitch in elements of a [rows, cols] (or NHWC) tensor whose leading dims are densely packed"""
    return 0 if t is None else t.stride(-2)


def ptr_array(tensors, device):
    """int64 device tensor holding the data pointers of `tensors` (a `const void* const*` argument)."""
    return torch.tensor([t.data_ptr() for t in tensors], dtype=torch.int64, device=device)


# ------------------------------------------------------------------------------------------------ GEMMs

def build_gemm_tn(A, B, *, N, K, bn, mode=EPI_LINEAR, act=ACT_NONE, act_slope=0.0, ln_eps=1e-5, bias=None, resid=None,
            aux=None, rowstat=None, outB=None, outB2=None, outF=None, n_valid=0, ones_col=-1, conv=None,
            acc_scale=None, outB_scale=None, tokens_per_sample=0, dilation=1):
    """D = epilogue(A @ B^T).  A: bf16 [M, >=K] (or NHWC [B,H,W,>=K] with conv=(B,H,W)); B: bf16 [N, taps*K]."""
    _req(A, torch.bfloat16, "A"); _req(B, torch.bfloat16, "B")
    _req(bias, torch.float32, "bias"); _req(resid, torch.float32, "resid")
    _req(aux, torch.bfloat16, "aux"); _req(rowstat, torch.float32, "rowstat")
    _req(outB, torch.bfloat16, "outB"); _req(outB2, torch.bfloat16, "outB2"); _req(outF, torch.float32, "outF")
    a = _S["kair_gemm_tn_args"]()
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
    a.acc_scale = _ptr(acc_scale); a.outB_scale = _ptr(outB_scale); a.tokens_per_sample = tokens_per_sample
    a.dilation = dilation
    return Launch("kair_gemm_tn", a)


def build_mlp_fused(X, W1, W2, *, N, F, bias1=None, bias2=None, resid=None, outF=None, outB=None, outU=None, outG=None,
                    rowstat=None, acc_scale=None, tokens_per_sample=0, ln=True, n_valid=0, ones_col=-1, ones_col_g=-1,
                    ln_eps=1e-5):
    """fc1 + GELU + fc2 + DropPath + residual (+ LayerNorm) of a Swin block's MLP half in one launch; the hidden
    activations stay on chip (outU / outG: optional bf16 copies for a backward pass).  See kair_b200.h."""
    _req(X, torch.bfloat16, "X"); _req(W1, torch.bfloat16, "W1"); _req(W2, torch.bfloat16, "W2")
    _req(bias1, torch.float32, "bias1"); _req(bias2, torch.float32, "bias2"); _req(resid, torch.float32, "resid")
    _req(outF, torch.float32, "outF"); _req(outB, torch.bfloat16, "outB"); _req(outU, torch.bfloat16, "outU")
    _req(outG, torch.bfloat16, "outG"); _req(rowstat, torch.float32, "rowstat"); _req(acc_scale, torch.float32, "acc_scale")
    a = _S["kair_mlp_fused_args"]()
    a.X = _ptr(X); a.x_pitch = _pitch(X)
    a.W1 = _ptr(W1); a.w1_pitch = _pitch(W1)
    a.W2 = _ptr(W2); a.w2_pitch = _pitch(W2)
    a.bias1 = _ptr(bias1); a.bias2 = _ptr(bias2)
    a.resid = _ptr(resid); a.resid_pitch = _pitch(resid)
    a.outF = _ptr(outF); a.outF_pitch = _pitch(outF)
    a.outB = _ptr(outB); a.outB_pitch = _pitch(outB)
    a.outU = _ptr(outU); a.outU_pitch = _pitch(outU)
    a.outG = _ptr(outG); a.outG_pitch = _pitch(outG)
    a.rowstat = _ptr(rowstat); a.acc_scale = _ptr(acc_scale); a.tokens_per_sample = tokens_per_sample
    a.M = X.shape[0]; a.N = N; a.F = F
    a.ln = 1 if ln else 0; a.n_valid = n_valid; a.ones_col = ones_col; a.ones_col_g = ones_col_g; a.ln_eps = ln_eps
    return Launch("kair_mlp_fused", a)


def build_gemm_nt(A, B, D, *, n1, n2, bn, splits=0, conv=None, dilation=1, swap=False):
    """D[n1, taps*n2] += A[:, :n1]^T @ B[:, :n2] (token contraction).  D is fp32 and must be pre-zeroed.
    swap (conv, n2 == 64): the same products written transposed, D[9*n2, n1] (rows tap*n2 + input channel)."""
    _req(A, torch.bfloat16, "A"); _req(B, torch.bfloat16, "B"); _req(D, torch.float32, "D")
    a = _S["kair_gemm_nt_args"]()
    a.A = _ptr(A); a.a_pitch = _pitch(A)
    a.B = _ptr(B); a.b_pitch = _pitch(B)
    a.D = _ptr(D); a.d_pitch = _pitch(D)
    a.n1 = n1; a.n2 = n2; a.bn = bn; a.splits = splits; a.dilation = dilation; a.swap = 1 if swap else 0
    if swap and (conv is None or n2 % 64 or D.shape[-2] < 9 * n2 or D.shape[-1] < n1):
        raise RuntimeError("kair_b200: gemm_nt swap needs a 3x3 convolution, n2 a multiple of 64 and D [9*n2, >= n1]")
    if conv is not None:
        a.conv = 1
        a.batch, a.H, a.W = conv
        a.T = conv[0] * conv[1] * conv[2]
    else:
        a.conv = 0
        a.T = A.shape[0]
    return Launch("kair_gemm_nt", a)


# ------------------------------------------------------------------------------------------------ window attention

def _attn_args(qkv, table, batch, H, W, num_heads, window, shift):
    _req(qkv, qkv.dtype, "qkv"); _req(table, torch.float32, "table")
    if qkv.dtype not in (torch.bfloat16, torch.float32):
        raise RuntimeError("kair_b200: qkv must be bf16 (tensor-core mode) or fp32 (fp32-accurate mode)")
    if not table.is_contiguous():
        raise RuntimeError("kair_b200: relative_position_bias_table must be contiguous")
    a = _S["kair_window_attn_args"]()
    a.qkv = _ptr(qkv); a.qkv_pitch = _pitch(qkv)
    a.table = _ptr(table)
    a.batch, a.H, a.W = batch, H, W
    a.num_heads, a.window, a.shift = num_heads, window, shift
    a.fp32 = 1 if qkv.dtype == torch.float32 else 0
    return a


def build_window_attn_fwd(qkv, table, out, *, batch, H, W, num_heads, window=8, shift=0):
    """out[t, h*32:(h+1)*32] = windowed softmax attention; qkv: bf16 [B*H*W, >=3*nH*32] in [3][nH][32] layout."""
    _req(out, qkv.dtype, "out")
    a = _attn_args(qkv, table, batch, H, W, num_heads, window, shift)
    a.out = _ptr(out); a.out_pitch = _pitch(out)
    return Launch("kair_window_attn_fwd", a)


def build_window_attn_bwd(qkv, table, dout, dqkv, dtable, *, batch, H, W, num_heads, window=8, shift=0):
    """Backward of window_attn_fwd: writes dqkv (bf16) and accumulates into dtable (fp32, caller zeroes)."""
    _req(dout, qkv.dtype, "dout"); _req(dqkv, qkv.dtype, "dqkv"); _req(dtable, torch.float32, "dtable")
    a = _attn_args(qkv, table, batch, H, W, num_heads, window, shift)
    a.dout = _ptr(dout); a.dout_pitch = _pitch(dout)
    a.dqkv = _ptr(dqkv); a.dqkv_pitch = _pitch(dqkv)
    a.dtable = _ptr(dtable)
    return Launch("kair_window_attn_bwd", a)


# ------------------------------------------------------------------------------------------------ packing

def _plist(tensors, like, keep):
    """device array of data pointers (`const void* const*`) for a list of same-shaped fp32 tensors, or None"""
    if tensors is None:
        return None
    for t in tensors:
        _req(t, torch.float32, "parameter")
        if not t.is_contiguous():
            raise RuntimeError("kair_b200: parameters / gradient buffers must be contiguous")
    arr = ptr_array(tensors, like.device)
    keep.append(arr)
    keep.extend(tensors)
    return arr.data_ptr()


def build_pack_linear(W, b, gamma, beta, row_map, col_map, row_scale, bias_const, Wp, WpT, bp, *, out_f, in_f, terms=1):
    """Pack L same-shaped nn.Linear layers (lists of fp32 tensors) into Wp [L,N_pad,terms*K_pad] (+ WpT
    [L,K_pad,terms*N_pad], bp).  terms > 1: weight-side segments of the fp32-accurate split (kair_b200.h)."""
    keep = []
    a = _S["kair_pack_linear_args"]()
    a.W = _plist(W, Wp, keep); a.b = _plist(b, Wp, keep)
    a.gamma = _plist(gamma, Wp, keep); a.beta = _plist(beta, Wp, keep)
    a.row_map = _ptr(row_map); a.col_map = _ptr(col_map)
    a.row_scale = _ptr(row_scale); a.bias_const = _ptr(bias_const)
    a.Wp = _ptr(Wp); a.WpT = _ptr(WpT); a.bp = _ptr(bp)
    a.L, a.N_pad = Wp.shape[0], Wp.shape[1]
    a.K_pad = Wp.shape[2] // terms
    a.terms = terms
    a.out_f, a.in_f = out_f, in_f
    return Launch("kair_pack_linear", a, keep)


def build_unpack_linear_grad(dWp, W, gamma, beta, dW, db, dgamma, dbeta, row_map, col_map, row_scale, *, out_f, in_f,
                             ones_col):
    """Map packed weight gradients dWp [L,N_pad,K_pad] back onto dW / db / dgamma / dbeta (lists of fp32 tensors)."""
    keep = []
    a = _S["kair_unpack_linear_args"]()
    a.dWp = _ptr(dWp)
    a.W = _plist(W, dWp, keep); a.gamma = _plist(gamma, dWp, keep); a.beta = _plist(beta, dWp, keep)
    a.dW = _plist(dW, dWp, keep); a.db = _plist(db, dWp, keep)
    a.dgamma = _plist(dgamma, dWp, keep); a.dbeta = _plist(dbeta, dWp, keep)
    a.row_map = _ptr(row_map); a.col_map = _ptr(col_map); a.row_scale = _ptr(row_scale)
    a.L, a.N_pad, a.K_pad = dWp.shape
    a.out_f, a.in_f, a.ones_col = out_f, in_f, ones_col
    return Launch("kair_unpack_linear_grad", a, keep)


def build_pack_conv(W, b, Wp, WpT, bp, *, O, I, I_pad, taps=9, oscale=None, oshift=None, terms=1):
    """Pack L same-shaped conv weights (O,I,kh,kw) into Wp [L,O_pad,taps*I_pad] (+ flipped WpT [L,I_pad,taps*O_pad]).
    oscale / oshift: optional per-layer (O,) fp32 tensors folded in (Wp = W*oscale[o], bp = b*oscale[o] + oshift[o])."""
    keep = []
    a = _S["kair_pack_conv_args"]()
    a.W = _plist(W, Wp, keep); a.b = _plist(b, Wp, keep)
    a.oscale = _plist(oscale, Wp, keep); a.oshift = _plist(oshift, Wp, keep)
    a.Wp = _ptr(Wp); a.WpT = _ptr(WpT); a.bp = _ptr(bp)
    a.L, a.O_pad = Wp.shape[0], Wp.shape[1]
    a.O, a.I, a.I_pad, a.taps = O, I, I_pad, taps
    a.terms = terms
    return Launch("kair_pack_conv", a, keep)


def build_unpack_conv_grad(dWp, dW, db, *, O, I, I_pad, taps=9, ones_col=-1, center_tap=4, dwp_pitch=0, transposed=False):
    keep = []
    a = _S["kair_unpack_conv_args"]()
    a.dWp = _ptr(dWp); a.dW = _plist(dW, dWp, keep); a.db = _plist(db, dWp, keep)
    a.L, a.O_pad = dWp.shape[0], (dWp.shape[2] if transposed else dWp.shape[1])
    a.O, a.I, a.I_pad, a.taps, a.ones_col, a.center_tap = O, I, I_pad, taps, ones_col, center_tap
    a.dwp_pitch = dwp_pitch
    a.transposed = 1 if transposed else 0
    if transposed and (db is not None or dWp.shape[1] != taps * I_pad):
        raise RuntimeError("kair_b200: transposed packed gradients are [L, taps*I_pad, O_pad] and carry no bias column")
    return Launch("kair_unpack_conv_grad", a, keep)


# ------------------------------------------------------------------------------------------------ LayerNorm

def build_layernorm_fwd(x, gamma, beta, *, C, eps=1e-5, y_f32=None, y_bf16=None, xhat_bf16=None, mean=None, rstd=None,
                  ones_col=-1, terms=1):
    _req(x, torch.float32, "x")
    a = _S["kair_layernorm_args"]()
    a.x = _ptr(x); a.gamma = _ptr(gamma); a.beta = _ptr(beta)
    a.y_f32 = _ptr(y_f32); a.y_bf16 = _ptr(y_bf16); a.xhat_bf16 = _ptr(xhat_bf16)
    a.mean = _ptr(mean); a.rstd = _ptr(rstd)
    a.T = x.shape[0]; a.C = C; a.pitch = x.stride(0); a.ones_col = ones_col; a.eps = eps
    a.terms = terms
    for t_ in (y_bf16, xhat_bf16):
        if t_ is not None and (t_.stride(0) != terms * x.stride(0)):
            raise RuntimeError("kair_b200: layernorm_fwd bf16 outputs must have pitch terms*pitch(x)")
    return Launch("kair_layernorm_fwd", a)


def build_layernorm_bwd(x, mean, rstd, gamma, dgamma, dbeta, *, C, dy_f32=None, dy_bf16=None, dx_add=None, dx_f32=None,
                  dx_bf16=None):
    a = _S["kair_layernorm_bwd_args"]()
    a.dy_f32 = _ptr(dy_f32); a.dy_bf16 = _ptr(dy_bf16)
    a.x = _ptr(x); a.mean = _ptr(mean); a.rstd = _ptr(rstd); a.gamma = _ptr(gamma)
    a.dx_add = _ptr(dx_add); a.dx_f32 = _ptr(dx_f32); a.dx_bf16 = _ptr(dx_bf16)
    a.dgamma = _ptr(dgamma); a.dbeta = _ptr(dbeta)
    a.T = x.shape[0]; a.C = C; a.pitch = x.stride(0)
    return Launch("kair_layernorm_bwd", a)


# ------------------------------------------------------------------------------------------------ thin convolutions

def _conv_in_args(img, W, b, mean, rng, O, O_pad, pitch):
    _req(img, torch.float32, "img"); _req(W, torch.float32, "W"); _req(mean, torch.float32, "mean")
    if not img.is_contiguous() or not W.is_contiguous():
        raise RuntimeError("kair_b200: conv_in needs contiguous NCHW image and weights")
    a = _S["kair_conv_in_args"]()
    a.img = _ptr(img); a.W = _ptr(W); a.b = _ptr(b); a.mean = _ptr(mean); a.range = rng
    a.batch, a.ci, a.H, a.Wd = img.shape
    a.O, a.O_pad, a.pitch = O, O_pad, pitch
    a.ones_col = -1
    return a


def build_conv_in_fwd(img, W, b, mean, rng, *, O, O_pad, pitch, out_f32=None, out_bf16=None, act=ACT_NONE, act_slope=0.0,
                ones_col=-1):
    a = _conv_in_args(img, W, b, mean, rng, O, O_pad, pitch)
    a.out_f32 = _ptr(out_f32); a.out_bf16 = _ptr(out_bf16)
    a.act, a.act_slope, a.ones_col = act, act_slope, ones_col
    return Launch("kair_conv_in_fwd", a)


def build_conv_in_wgrad(img, W, mean, rng, dW, db, *, O, O_pad, pitch, dy_f32=None, dy_bf16=None):
    a = _conv_in_args(img, W, None, mean, rng, O, O_pad, pitch)
    a.dy_f32 = _ptr(dy_f32); a.dy_bf16 = _ptr(dy_bf16); a.dW = _ptr(dW); a.db = _ptr(db)
    return Launch("kair_conv_in_wgrad", a)


def _conv_out_args(W, shape, C, pitch, out_scale):
    _req(W, torch.float32, "W")
    a = _S["kair_conv_out_args"]()
    a.W = _ptr(W)
    a.batch, a.co, a.H, a.Wd = shape
    a.C, a.pitch, a.out_scale = C, pitch, out_scale
    return a


def build_conv_out_fwd(feat, W, b, out, out_mean, *, C, pitch, out_scale=1.0, skip=None, skip_mean=None, skip_scale=0.0):
    _req(feat, feat.dtype, "feat"); _req(out, torch.float32, "out")
    a = _conv_out_args(W, out.shape, C, pitch, out_scale)
    if feat.dtype == torch.float32:   # fp32-accurate mode
        a.feat_f32 = _ptr(feat)
    elif feat.dtype == torch.bfloat16:
        a.feat = _ptr(feat)
    else:
        raise RuntimeError("kair_b200: conv_out features must be bf16 or fp32")
    a.b = _ptr(b); a.out = _ptr(out); a.out_mean = _ptr(out_mean)
    a.skip = _ptr(skip); a.skip_mean = _ptr(skip_mean); a.skip_scale = skip_scale
    return Launch("kair_conv_out_fwd", a)


def build_conv_out_dgrad(dout, W, *, C, pitch, out_scale=1.0, dfeat_f32=None, dfeat_bf16=None, mask=None,
                         mask_slope=0.0):
    _req(dout, torch.float32, "dout"); _req(mask, torch.bfloat16, "mask")
    a = _conv_out_args(W, dout.shape, C, pitch, out_scale)
    a.dout = _ptr(dout); a.dfeat_f32 = _ptr(dfeat_f32); a.dfeat_bf16 = _ptr(dfeat_bf16)
    a.mask = _ptr(mask); a.mask_slope = mask_slope
    return Launch("kair_conv_out_dgrad", a)


def build_conv_out_wgrad(dout, feat, W, dW, db, *, C, pitch, out_scale=1.0):
    _req(dout, torch.float32, "dout"); _req(feat, feat.dtype, "feat")
    a = _conv_out_args(W, dout.shape, C, pitch, out_scale)
    if feat.dtype == torch.float32:
        a.feat_f32 = _ptr(feat)
    else:
        _req(feat, torch.bfloat16, "feat")
        a.feat = _ptr(feat)
    a.dout = _ptr(dout); a.dW = _ptr(dW); a.db = _ptr(db)
    return Launch("kair_conv_out_wgrad", a)


def build_im2col3x3(img, out, *, cpad, direction, mean=None, scale=1.0, ones_col=-1):
    """out[t, tap*cpad + ch] = (img[ch, p(t) + direction*off(tap)] - mean[ch]) * scale as bf16 (see kair_b200.h)."""
    _req(img, torch.float32, "img"); _req(out, torch.bfloat16, "out"); _req(mean, torch.float32, "mean")
    if not img.is_contiguous():
        raise RuntimeError("kair_b200: im2col3x3 needs a contiguous NCHW image")
    a = _S["kair_im2col_args"]()
    a.img = _ptr(img); a.mean = _ptr(mean); a.out = _ptr(out); a.scale = scale
    a.batch, a.c, a.H, a.Wd = img.shape
    a.cpad, a.pitch, a.dir, a.ones_col = cpad, out.stride(0), direction, ones_col
    return Launch("kair_im2col3x3", a)


def build_conv_out_pack(*, mode, co, C, cpad, W=None, Wt=None, D=None, dW=None, db=None, ones_row=-1):
    """mode 0: Wt[k, tap*cpad + c] = W[c, k, tap] (bf16); mode 1: dW[c, k, tap] = D[k, tap*cpad + c], db from ones_row."""
    _req(W, torch.float32, "W"); _req(Wt, torch.bfloat16, "Wt"); _req(D, torch.float32, "D")
    _req(dW, torch.float32, "dW"); _req(db, torch.float32, "db")
    a = _S["kair_conv_out_pack_args"]()
    a.W = _ptr(W); a.Wt = _ptr(Wt); a.D = _ptr(D); a.dW = _ptr(dW); a.db = _ptr(db)
    ref = Wt if mode == 0 else D
    a.mode, a.co, a.C, a.C_pad, a.cpad, a.pitch, a.ones_row = mode, co, C, ref.shape[0], cpad, ref.stride(0), ones_row
    return Launch("kair_conv_out_pack", a)


SPLIT_COPY, SPLIT_GELU, SPLIT_DGELU, SPLIT_ACT_BWD, SPLIT_LRELU = range(5)


def build_split_terms(x, out, *, n, terms, op=SPLIT_COPY, aux=None, out_f32=None, row_scale=None, rows_per_sample=0,
                      ones_col=-1, slope=0.0):
    """fp32 rows -> split bf16 GEMM operand [T, terms*n] (fp32-accurate mode, kair_b200.h 'SPLIT PRECISION'), fused with
    the element-wise stage `op`; out_f32 optionally receives the transformed fp32 values."""
    _req(x, torch.float32, "x"); _req(out, torch.bfloat16, "out"); _req(aux, torch.float32, "aux")
    _req(out_f32, torch.float32, "out_f32"); _req(row_scale, torch.float32, "row_scale")
    a = _S["kair_split_args"]()
    a.x = _ptr(x); a.x_pitch = _pitch(x)
    a.aux = _ptr(aux); a.aux_pitch = _pitch(aux)
    a.out = _ptr(out); a.out_pitch = _pitch(out)
    a.out_f32 = _ptr(out_f32); a.out_f32_pitch = _pitch(out_f32)
    a.row_scale = _ptr(row_scale)
    a.T = x.shape[0]; a.n = n; a.terms = terms; a.op = op; a.ones_col = ones_col
    a.rows_per_sample = rows_per_sample; a.slope = slope
    return Launch("kair_split_terms", a)


def build_ln_bwd_split(g, xhat, rstd, outF, *, C, terms, resid=None, out=None, row_scale=None, rows_per_sample=0):
    """LayerNorm backward + residual gradient on split x-hat (fp32-accurate mode; counterpart of EPI_LNBWD)."""
    _req(g, torch.float32, "g"); _req(xhat, torch.bfloat16, "xhat"); _req(rstd, torch.float32, "rstd")
    _req(outF, torch.float32, "outF"); _req(resid, torch.float32, "resid"); _req(out, torch.bfloat16, "out")
    _req(row_scale, torch.float32, "row_scale")
    if _pitch(outF) != _pitch(g) or (resid is not None and _pitch(resid) != _pitch(g)):
        raise RuntimeError("kair_b200: ln_bwd_split fp32 operands must share one pitch")
    a = _S["kair_ln_bwd_split_args"]()
    a.g = _ptr(g); a.xhat = _ptr(xhat); a.xhat_pitch = _pitch(xhat); a.rstd = _ptr(rstd); a.resid = _ptr(resid)
    a.outF = _ptr(outF); a.out = _ptr(out); a.out_pitch = _pitch(out); a.row_scale = _ptr(row_scale)
    a.T = g.shape[0]; a.C = C; a.pitch = _pitch(g); a.terms = terms; a.rows_per_sample = rows_per_sample
    return Launch("kair_ln_bwd_split", a)


SPECT_METHODS = {"log": 0, "anscombe": 1, "linear": 2, "none": 3}


def build_spect_patches(frames_L, frames_H, sample, out_L, out_H, *, patch, method, max_pixel):
    """One SPECT training batch from device-resident frame stacks (see kair_b200.h)."""
    ref = frames_L if frames_L is not None else frames_H
    for t, n in ((frames_L, "frames_L"), (frames_H, "frames_H"), (out_L, "out_L"), (out_H, "out_H")):
        _req(t, torch.float32, n)
        if t is not None and not t.is_contiguous():
            raise RuntimeError("kair_b200: spect_patches tensors must be contiguous")
    _req(sample, torch.int32, "sample")
    if ref.dim() != 4 or ref.shape[1] != 2 or sample.dim() != 2 or sample.shape[1] != 4 or not sample.is_contiguous():
        raise RuntimeError("kair_b200: frames must be [n, 2, H, W] and sample [B, 4] int32")
    a = _S["kair_spect_patch_args"]()
    a.frames_L = _ptr(frames_L); a.frames_H = _ptr(frames_H); a.sample = _ptr(sample)
    a.out_L = _ptr(out_L); a.out_H = _ptr(out_H)
    a.B, a.P, a.FH, a.FW = sample.shape[0], patch, ref.shape[2], ref.shape[3]
    a.method = SPECT_METHODS[method]; a.max_pixel = float(max_pixel)
    return Launch("kair_spect_patches", a)


def build_spect_normalize(x, y, *, method, max_pixel, inverse=False):
    _req(x, torch.float32, "x"); _req(y, torch.float32, "y")
    if not (x.is_contiguous() and y.is_contiguous()) or x.numel() != y.numel():
        raise RuntimeError("kair_b200: spect_normalize needs contiguous tensors of one size")
    a = _S["kair_spect_norm_args"]()
    a.x = _ptr(x); a.y = _ptr(y); a.n = x.numel()
    a.method = SPECT_METHODS[method]; a.inverse = 1 if inverse else 0; a.max_pixel = float(max_pixel)
    return Launch("kair_spect_normalize", a)


def build_spect_metrics(E, H, out, *, max_val):
    """E, H: (C, h, w) de-normalised counts; out: fp64 [C, 4] (accumulated; zero it first)."""
    _req(E, torch.float32, "E"); _req(H, torch.float32, "H"); _req(out, torch.float64, "out")
    if not (E.is_contiguous() and H.is_contiguous() and out.is_contiguous()) or E.shape != H.shape or E.dim() != 3:
        raise RuntimeError("kair_b200: spect_metrics needs contiguous (C, h, w) images of one shape")
    a = _S["kair_spect_metric_args"]()
    a.E = _ptr(E); a.H = _ptr(H); a.out = _ptr(out)
    a.C, a.h, a.w = E.shape
    a.max_val = float(max_val)
    return Launch("kair_spect_metrics", a)


def _tiles_args(tiles, origins, C, H, W):
    _req(tiles, torch.float32, "tiles"); _req(origins, torch.int32, "origins")
    if tiles.dim() != 4 or not tiles.is_contiguous() or origins.dim() != 2 or origins.shape[1] != 3 or \
            not origins.is_contiguous() or origins.shape[0] != tiles.shape[0] or tiles.shape[1] != C:
        raise RuntimeError("kair_b200: tiles must be contiguous (n, C, th, tw) and origins [n, 3] int32")
    a = _S["kair_tiles_args"]()
    a.tiles = _ptr(tiles); a.origins = _ptr(origins)
    a.n, a.C, a.th, a.tw = tiles.shape
    a.H, a.W = H, W
    return a


def build_tiles_gather(img, origins, tiles):
    """tiles[n] = D_mode(img[:, y:y+.., x:x+..]) for origins[n] = (y, x, mode); img (C, H, W) fp32."""
    _req(img, torch.float32, "img")
    if img.dim() != 3 or not img.is_contiguous():
        raise RuntimeError("kair_b200: tiles_gather needs a contiguous (C, H, W) image")
    a = _tiles_args(tiles, origins, img.shape[0], img.shape[1], img.shape[2])
    a.img = _ptr(img)
    return Launch("kair_tiles_gather", a)


def build_tiles_accumulate(tiles, origins, E, Wsum=None, *, weight=1.0):
    """E[:, y.., x..] += weight * D_mode^-1(tiles[n]); Wsum[y.., x..] += weight.  E (C, H, W), Wsum (H, W) fp32."""
    _req(E, torch.float32, "E"); _req(Wsum, torch.float32, "Wsum")
    if E.dim() != 3 or not E.is_contiguous() or (Wsum is not None and (not Wsum.is_contiguous() or Wsum.shape != E.shape[1:])):
        raise RuntimeError("kair_b200: tiles_accumulate needs contiguous E (C, H, W) and Wsum (H, W)")
    a = _tiles_args(tiles, origins, E.shape[0], E.shape[1], E.shape[2])
    a.E = _ptr(E); a.Wsum = _ptr(Wsum); a.weight = float(weight)
    return Launch("kair_tiles_accumulate", a)


def build_axpby(x, *, n, a=1.0, y=None, b=0.0, out_f32=None, out_bf16=None, mask=None, mask_slope=0.0, n_pad=None):
    """out = a*x + b*y on fp32 rows ([T, >= n]); bf16 copy optionally masked by a saved activation and zero-padded to
    n_pad columns (kair_b200.h)."""
    _req(x, torch.float32, "x"); _req(y, torch.float32, "y"); _req(out_f32, torch.float32, "out_f32")
    _req(out_bf16, torch.bfloat16, "out_bf16"); _req(mask, torch.bfloat16, "mask")
    g = _S["kair_axpby_args"]()
    g.x = _ptr(x); g.x_pitch = _pitch(x); g.y = _ptr(y); g.y_pitch = _pitch(y)
    g.out_f32 = _ptr(out_f32); g.out_f32_pitch = _pitch(out_f32)
    g.out_bf16 = _ptr(out_bf16); g.out_bf16_pitch = _pitch(out_bf16)
    g.mask = _ptr(mask); g.mask_pitch = _pitch(mask)
    g.T = x.shape[0]; g.n = n; g.n_pad = n_pad if n_pad is not None else n
    g.a, g.b, g.mask_slope = float(a), float(b), float(mask_slope)
    return Launch("kair_axpby", g)


def build_ffd_pack(img, sigma, out):
    """FFDNet input: replicate-pad to even size, PixelUnShuffle(2), noise-level map as an extra channel -> bf16 token
    rows [B*H2*W2, pitch] (kair_b200.h)."""
    _req(img, torch.float32, "img"); _req(sigma, torch.float32, "sigma"); _req(out, torch.bfloat16, "out")
    if not img.is_contiguous() or img.dim() != 4 or sigma.numel() != img.shape[0] or not sigma.is_contiguous():
        raise RuntimeError("kair_b200: ffd_pack needs a contiguous (B, C, h, w) image and B noise levels")
    a = _S["kair_ffd_pack_args"]()
    a.img = _ptr(img); a.sigma = _ptr(sigma); a.out = _ptr(out); a.pitch = _pitch(out)
    a.batch, a.C, a.h, a.w = img.shape
    return Launch("kair_ffd_pack", a)


def build_add_inplace(y, x, yb=None):
    """y += x (fp32, in place); yb (bf16, optional) receives the updated y."""
    _req(y, torch.float32, "y"); _req(x, torch.float32, "x"); _req(yb, torch.bfloat16, "yb")
    if not (y.is_contiguous() and x.is_contiguous() and (yb is None or yb.is_contiguous())):
        raise RuntimeError("kair_b200: add_inplace operands must be contiguous")
    a = _S["kair_add_args"]()
    a.y, a.x, a.yb, a.n = _ptr(y), _ptr(x), _ptr(yb), y.numel()
    return Launch("kair_add_inplace_s", a)


# ------------------------------------------------------------------------------------------------ conv-net helpers

def _regroup_args(batch, H, W, C, fine_pitch, coarse_pitch):
    a = _S["kair_regroup_args"]()
    a.batch, a.H, a.W, a.C = batch, H, W, C
    a.fine_pitch, a.coarse_pitch = fine_pitch, coarse_pitch
    return a


def build_space_to_depth(fine, coarse, *, batch, H, W, C):
    """coarse[(b,h,w)][tap*C + c] = fine[(b,2h+dy,2w+dx)][c], tap = dy*2+dx; (H, W) is the fine resolution."""
    _req(fine, torch.bfloat16, "fine"); _req(coarse, torch.bfloat16, "coarse")
    a = _regroup_args(batch, H, W, C, _pitch(fine), _pitch(coarse))
    a.fine_bf16 = _ptr(fine); a.coarse_bf16 = _ptr(coarse)
    return Launch("kair_space_to_depth", a)


def build_depth_to_space(coarse, *, batch, H, W, C, out_f32=None, out_bf16=None):
    """inverse regrouping; coarse is fp32 or bf16 [T/4, >=4C], outputs fp32 and/or bf16 [T, >=C]."""
    if coarse.dtype not in (torch.float32, torch.bfloat16):
        raise RuntimeError("kair_b200: depth_to_space input must be fp32 or bf16")
    _req(coarse, coarse.dtype, "coarse"); _req(out_f32, torch.float32, "out_f32"); _req(out_bf16, torch.bfloat16, "out_bf16")
    ref = out_f32 if out_f32 is not None else out_bf16
    a = _regroup_args(batch, H, W, C, _pitch(ref), _pitch(coarse))
    if out_f32 is not None and out_bf16 is not None and _pitch(out_f32) != _pitch(out_bf16):
        raise RuntimeError("kair_b200: depth_to_space outputs must share one pitch")
    if coarse.dtype == torch.float32:
        a.coarse_in_f32 = _ptr(coarse)
    else:
        a.coarse_in_bf16 = _ptr(coarse)
    a.fine_out_f32 = _ptr(out_f32); a.fine_out_bf16 = _ptr(out_bf16)
    return Launch("kair_depth_to_space", a)


def build_colsum(x, out, *, C, scale=1.0):
    """out[c] += scale * sum_t x[t, c]; x fp32 or bf16 [T, >=C], out fp32 (caller zeroes)."""
    _req(out, torch.float32, "out")
    a = _S["kair_colsum_args"]()
    if x.dtype == torch.bfloat16:
        a.x_bf16 = _ptr(x)
    elif x.dtype == torch.float32:
        a.x_f32 = _ptr(x)
    else:
        raise RuntimeError("kair_b200: colsum input must be fp32 or bf16")
    _req(x, x.dtype, "x")
    a.out = _ptr(out); a.T = x.shape[0]; a.pitch = _pitch(x); a.C = C; a.scale = scale
    return Launch("kair_colsum", a)


def _bn_args(C, T, pitch, eps, momentum, act, act_slope):
    a = _S["kair_bn_args"]()
    a.C, a.T, a.pitch, a.eps, a.momentum, a.act, a.act_slope = C, T, pitch, eps, momentum, act, act_slope
    return a


def _set_dy(a, dy):
    """dy of the BatchNorm backward kernels: bf16 or fp32 [T, pitch] with the pitch of z"""
    if dy.dtype == torch.float32:
        a.dy_f32 = _ptr(dy)
    elif dy.dtype == torch.bfloat16:
        a.dy = _ptr(dy)
    else:
        raise RuntimeError("kair_b200: BatchNorm dy must be fp32 or bf16")
    if _pitch(dy) != a.pitch:
        raise RuntimeError("kair_b200: BatchNorm dy must share the pitch of z")


def build_bn_stats(z, acc, *, C):
    _req(z, torch.float32, "z"); _req(acc, torch.float64, "acc")
    a = _bn_args(C, z.shape[0], _pitch(z), 0.0, 0.0, 0, 0.0)
    a.z = _ptr(z); a.acc = _ptr(acc)
    return Launch("kair_bn_stats", a)


def build_bn_finalize(acc, mean, rstd, running_mean, running_var, num_batches_tracked, *, C, T, eps, momentum):
    _req(acc, torch.float64, "acc"); _req(mean, torch.float32, "mean"); _req(rstd, torch.float32, "rstd")
    _req(running_mean, torch.float32, "running_mean"); _req(running_var, torch.float32, "running_var")
    _req(num_batches_tracked, torch.int64, "num_batches_tracked")
    a = _bn_args(C, T, C, eps, momentum, 0, 0.0)
    a.acc = _ptr(acc); a.mean = _ptr(mean); a.rstd = _ptr(rstd)
    a.running_mean = _ptr(running_mean); a.running_var = _ptr(running_var)
    a.num_batches_tracked = _ptr(num_batches_tracked)
    return Launch("kair_bn_finalize", a)


def build_bn_apply(z, mean, rstd, gamma, beta, *, C, act=ACT_NONE, act_slope=0.0, out_bf16=None, out_f32=None):
    _req(z, torch.float32, "z"); _req(out_bf16, torch.bfloat16, "out_bf16"); _req(out_f32, torch.float32, "out_f32")
    a = _bn_args(C, z.shape[0], _pitch(z), 0.0, 0.0, act, act_slope)
    a.z = _ptr(z); a.mean = _ptr(mean); a.rstd = _ptr(rstd); a.gamma = _ptr(gamma); a.beta = _ptr(beta)
    a.out_bf16 = _ptr(out_bf16); a.out_f32 = _ptr(out_f32)
    return Launch("kair_bn_apply", a)


def build_bn_bwd_reduce(z, y, dy, mean, rstd, dgamma, dbeta, *, C, act=ACT_NONE, act_slope=0.0):
    _req(z, torch.float32, "z"); _req(y, torch.bfloat16, "y"); _req(dy, dy.dtype, "dy")
    _req(dgamma, torch.float32, "dgamma"); _req(dbeta, torch.float32, "dbeta")
    a = _bn_args(C, z.shape[0], _pitch(z), 0.0, 0.0, act, act_slope)
    a.z = _ptr(z); a.y = _ptr(y); _set_dy(a, dy); a.mean = _ptr(mean); a.rstd = _ptr(rstd)
    a.dgamma = _ptr(dgamma); a.dbeta = _ptr(dbeta)
    return Launch("kair_bn_bwd_reduce", a)


def build_bn_bwd_apply(z, y, dy, mean, rstd, gamma, dgamma, dbeta, dz, *, C, act=ACT_NONE, act_slope=0.0, dbias=None):
    _req(z, torch.float32, "z"); _req(y, torch.bfloat16, "y"); _req(dy, dy.dtype, "dy")
    _req(dz, torch.bfloat16, "dz"); _req(dbias, torch.float32, "dbias")
    a = _bn_args(C, z.shape[0], _pitch(z), 0.0, 0.0, act, act_slope)
    a.z = _ptr(z); a.y = _ptr(y); _set_dy(a, dy); a.mean = _ptr(mean); a.rstd = _ptr(rstd); a.gamma = _ptr(gamma)
    a.dgamma = _ptr(dgamma); a.dbeta = _ptr(dbeta); a.out_bf16 = _ptr(dz); a.dbias = _ptr(dbias)
    return Launch("kair_bn_bwd_apply", a)


def build_bn_fold(gamma, beta, running_mean, running_var, scale, shift, *, C, eps):
    for t, n in ((gamma, "gamma"), (beta, "beta"), (running_mean, "running_mean"), (running_var, "running_var"),
                 (scale, "scale"), (shift, "shift")):
        _req(t, torch.float32, n)
    a = _bn_args(C, 1, C, eps, 0.0, 0, 0.0)
    a.gamma = _ptr(gamma); a.beta = _ptr(beta); a.running_mean = _ptr(running_mean); a.running_var = _ptr(running_var)
    a.scale = _ptr(scale); a.shift = _ptr(shift)
    return Launch("kair_bn_fold", a)


def _shuffle_args(x, out, batch, H, W, C, r, allow_fp32=False):
    if allow_fp32 and x.dtype == torch.float32:
        _req(x, torch.float32, "in"); _req(out, torch.float32, "out")
    else:
        _req(x, torch.bfloat16, "in"); _req(out, torch.bfloat16, "out")
    a = _S["kair_shuffle_args"]()
    a.fp32 = 1 if x.dtype == torch.float32 else 0
    setattr(a, "in", _ptr(x)); a.out = _ptr(out)
    a.in_pitch, a.out_pitch = _pitch(x), _pitch(out)
    a.batch, a.H, a.W, a.C, a.r = batch, H, W, C, r
    return a


def build_pixel_shuffle(x, out, *, batch, H, W, C, r):
    """nn.PixelShuffle(r) on token features: x [B*H*W, >= r*r*C] -> out [B*rH*rW, >= C] (bf16)."""
    return Launch("kair_pixel_shuffle", _shuffle_args(x, out, batch, H, W, C, r, allow_fp32=True))


def build_upsample_nearest2(x, out, *, batch, H, W, C):
    """nearest-neighbour x2 of token features: x [B*H*W, >= C] -> out [B*2H*2W, >= C] (bf16)."""
    return Launch("kair_upsample_nearest2", _shuffle_args(x, out, batch, H, W, C, 2, allow_fp32=True))


def build_pixel_shuffle_bwd(dy, dx, *, batch, H, W, C, r):
    """adjoint of build_pixel_shuffle: dy [B*rH*rW, >= C] -> dx [B*H*W, >= r*r*C] (bf16)."""
    return Launch("kair_pixel_shuffle_bwd", _shuffle_args(dy, dx, batch, H, W, C, r))


def build_upsample_nearest2_bwd(dy, dx, *, batch, H, W, C, mask=None, mask_slope=0.0):
    """adjoint of build_upsample_nearest2 (2x2 sums), times the LeakyReLU mask of the low-resolution activation."""
    a = _shuffle_args(dy, dx, batch, H, W, C, 2)
    if mask is not None:
        _req(mask, torch.bfloat16, "mask")
        a.mask = _ptr(mask); a.mask_pitch = _pitch(mask); a.mask_slope = mask_slope
    return Launch("kair_upsample_nearest2_bwd", a)


def build_pixel_shuffle_out_bwd(dout, dx, *, r, co, scale=1.0):
    """adjoint of build_pixel_shuffle_out: dout (B, co, r*H, r*W) fp32 -> dx bf16 [B*H*W, >= co*r*r] (* scale)."""
    _req(dx, torch.bfloat16, "in"); _req(dout, torch.float32, "out")
    if not dout.is_contiguous():
        raise RuntimeError("kair_b200: pixel_shuffle_out_bwd needs a contiguous NCHW gradient")
    a = _S["kair_shuffle_out_args"]()
    setattr(a, "in", _ptr(dx)); a.out = _ptr(dout); a.mean = None
    a.in_pitch = _pitch(dx); a.scale = scale
    a.batch, a.co = dout.shape[0], co
    a.H, a.W, a.r = dout.shape[2] // r, dout.shape[3] // r, r
    return Launch("kair_pixel_shuffle_out_bwd", a)


def build_pixel_shuffle_out(x, out, mean, *, r, scale=1.0):
    """out (B, co, r*H, r*W) fp32 = PixelShuffle(r)(x) * scale + mean[c]; x bf16 [B*H*W, >= co*r*r]."""
    _req(x, x.dtype, "in"); _req(out, torch.float32, "out"); _req(mean, torch.float32, "mean")
    if x.dtype not in (torch.bfloat16, torch.float32):
        raise RuntimeError("kair_b200: pixel_shuffle_out input must be bf16 or fp32")
    if not out.is_contiguous():
        raise RuntimeError("kair_b200: pixel_shuffle_out needs a contiguous NCHW output")
    a = _S["kair_shuffle_out_args"]()
    a.fp32 = 1 if x.dtype == torch.float32 else 0
    setattr(a, "in", _ptr(x)); a.out = _ptr(out); a.mean = _ptr(mean)
    a.in_pitch = _pitch(x); a.scale = scale
    a.batch, a.co = out.shape[0], out.shape[1]
    a.H, a.W, a.r = out.shape[2] // r, out.shape[3] // r, r
    return Launch("kair_pixel_shuffle_out", a)


def build_conv_chain(act0, act1, Wp, bias, slope, sync, *, L, conv):
    """L 3x3 convolutions 64 -> 64 (+ bias, leaky-ReLU slope[l]) in one persistent launch; layer l reads act(l & 1) and
    writes act((l + 1) & 1).  Wp bf16 [L, 64, 576], bias f32 [L, 64] or None, slope f32 [L], sync int32 [2] (zero)."""
    _req(act0, torch.bfloat16, "act0"); _req(act1, torch.bfloat16, "act1"); _req(Wp, torch.bfloat16, "Wp")
    _req(bias, torch.float32, "bias"); _req(slope, torch.float32, "slope"); _req(sync, torch.int32, "sync")
    B, H, W = conv
    if Wp.dim() != 3 or tuple(Wp.shape[1:]) != (64, 576) or Wp.shape[0] < L or not Wp.is_contiguous():
        raise RuntimeError("kair_b200: conv_chain needs contiguous packed weights [L, 64, 576]")
    if (bias is not None and (not bias.is_contiguous() or bias.numel() < L * 64)) or slope.numel() < L or sync.numel() < 2:
        raise RuntimeError("kair_b200: conv_chain needs bias [L, 64], slope [L] and two sync words")
    if _pitch(act0) != _pitch(act1) or act0.numel() < B * H * W * 64 or act1.numel() < B * H * W * 64:
        raise RuntimeError("kair_b200: conv_chain needs two [B*H*W, >= 64] activation buffers of one pitch")
    a = _S["kair_conv_chain_args"]()
    a.act0, a.act1, a.Wp, a.bias, a.slope, a.sync = _ptr(act0), _ptr(act1), _ptr(Wp), _ptr(bias), _ptr(slope), _ptr(sync)
    a.pitch, a.L, a.batch, a.H, a.W = _pitch(act0), L, B, H, W
    return Launch("kair_conv_chain", a, keep=(act0, act1, Wp, bias, slope, sync))


def build_pixel_loss(x, y, dx, loss, scratch, *, eps=1e-9, mode=0, max_pixel=0.0):
    """loss[0] = mean(sqrt((x-y)^2 + eps)) (mode 0, CharbonnierLoss) or mean|x-y| (mode 1) and dx = d(loss)/dx."""
    _req(x, torch.float32, "x"); _req(y, torch.float32, "y"); _req(dx, torch.float32, "dx")
    _req(loss, torch.float32, "loss"); _req(scratch, torch.float64, "scratch")
    if not (x.is_contiguous() and y.is_contiguous() and (dx is None or dx.is_contiguous())) or x.shape != y.shape:
        raise RuntimeError("kair_b200: pixel_loss needs contiguous x, y (and dx) of the same shape")
    if scratch.numel() < 2:
        raise RuntimeError("kair_b200: pixel_loss scratch must hold two 8-byte words")
    a = _S["kair_pixel_loss_args"]()
    a.x, a.y, a.dx, a.loss, a.scratch = _ptr(x), _ptr(y), _ptr(dx), _ptr(loss), _ptr(scratch)
    a.n, a.eps, a.mode, a.max_pixel = x.numel(), eps, mode, max_pixel
    return Launch("kair_pixel_loss", a)


ADAM_CHUNK = 8192


def build_adam_ema(p, g, m, v, e, *, beta1, beta2, eps, weight_decay=0.0, ema_decay=0.0):
    """One launch of Adam (+ EMA when `e` is given) over the tensor lists.  The per-step scalars (step_size,
    bc2_sqrt) are set on the returned launch's `args` before every run (see kair_b200/optim.py)."""
    keep = []
    dev = p[0].device
    a = _S["kair_adam_ema_args"]()
    a.g = _plist(g, p[0], keep)
    a.p = _plist(p, p[0], keep)
    a.m = _plist(m, p[0], keep); a.v = _plist(v, p[0], keep)
    a.e = _plist(e, p[0], keep)
    for lst in (g, m, v) + ((e,) if e is not None else ()):
        if [t.numel() for t in lst] != [t.numel() for t in p]:
            raise RuntimeError("kair_b200: adam_ema tensor lists must match the parameters element for element")
    sizes = torch.tensor([t.numel() for t in p], dtype=torch.int64, device=dev)
    ct, co = [], []
    for i, t in enumerate(p):
        for off in range(0, t.numel(), ADAM_CHUNK):
            ct.append(i); co.append(off)
    chunk_tensor = torch.tensor(ct, dtype=torch.int32, device=dev)
    chunk_offset = torch.tensor(co, dtype=torch.int64, device=dev)
    keep += [sizes, chunk_tensor, chunk_offset]
    a.sizes, a.chunk_tensor, a.chunk_offset = sizes.data_ptr(), chunk_tensor.data_ptr(), chunk_offset.data_ptr()
    a.n_chunks, a.chunk_elems = len(ct), ADAM_CHUNK
    a.beta1, a.beta2, a.eps, a.weight_decay, a.ema_decay = beta1, beta2, eps, weight_decay, ema_decay
    a.step_size, a.bc2_sqrt = 0.0, 1.0
    return Launch("kair_adam_ema", a, keep)


def adam_ema_set_grads(launch, g, tables):
    """Point a recorded kair_adam_ema launch at other gradient tensors.  autograd hands out new gradient tensors every
    step, but the caching allocator cycles through a few blocks: `tables` (owned by the caller) maps each set of
    addresses seen so far to its device pointer table, so in steady state this only swaps one pointer in the argument
    struct -- no copy, no host synchronisation."""
    key = tuple(t.data_ptr() for t in g)
    arr = tables.get(key)
    if arr is None:
        for t in g:
            _req(t, torch.float32, "gradient")
            if not t.is_contiguous():
                raise RuntimeError("kair_b200: gradients must be contiguous")
        if len(tables) >= 8:
            tables.clear()
        arr = tables[key] = ptr_array(g, g[0].device)
    launch.args.g = arr.data_ptr()


def build_copy(dst, src):
    """dst <- src on the stream (cudaMemcpyAsync device-to-device: a handful of values, e.g. a bias gradient picked out of
    a column-sum vector); listable like every other launch."""
    if not (dst.is_contiguous() and src.is_contiguous()) or dst.dtype != src.dtype or dst.numel() != src.numel():
        raise RuntimeError("kair_b200: copy needs contiguous tensors of one dtype and size")
    a = _S["kair_memcpy_args"]()
    a.dst, a.src, a.bytes = _ptr(dst), _ptr(src), dst.numel() * dst.element_size()
    return Launch("kair_memcpy_d2d", a)


# ------------------------------------------------------------------------------------------------ lifetimes
# A Launch stores raw device addresses, so it must own a reference to every tensor it points into: the plan's
# workspace tensors are otherwise free to be garbage-collected (and their memory reused) once the builder returns.

def _tensors_of(args, kwargs):
    out = []
    for v in list(args) + list(kwargs.values()):
        if isinstance(v, torch.Tensor):
            out.append(v)
        elif isinstance(v, (list, tuple)):
            out.extend(t for t in v if isinstance(t, torch.Tensor))
    return out


def _owning(builder):
    def build(*args, **kwargs):
        launch = builder(*args, **kwargs)
        launch.keep = tuple(launch.keep) + tuple(_tensors_of(args, kwargs))
        return launch
    build.__name__ = builder.__name__
    build.__doc__ = builder.__doc__
    return build


for _n in [k for k in list(globals()) if k.startswith("build_")]:
    globals()[_n] = _owning(globals()[_n])
del _n


# ------------------------------------------------------------------------------------------------ eager forms
# `name(...)` == `build_name(...).run(current stream)`: used by the unit tests and one-off calls.

def _eager(builder):
    def run(*args, **kwargs):
        builder(*args, **kwargs).run(_stream())
    run.__name__ = builder.__name__[len("build_"):]
    run.__doc__ = builder.__doc__
    return run


for _n in [k for k in list(globals()) if k.startswith("build_")]:
    globals()[_n[len("build_"):]] = _eager(globals()[_n])
del _n
