"""Each memory-bound helper kernel vs its PyTorch statement in tests/emu_ops.py on the same seeded inputs."""
import pytest
import torch

import emu_ops

pytestmark = pytest.mark.gpu


def _g(seed):
    return torch.Generator().manual_seed(seed)


def _run_both(name, cpu_kwargs, out_keys, tol=1e-4, list_keys=()):
    """Run build_<name> through the emulator (CPU tensors) and the CUDA kernel (GPU copies); compare out_keys."""
    from kair_b200 import ops

    def to_gpu(v):
        if isinstance(v, torch.Tensor):
            return v.clone().cuda()
        if isinstance(v, list):
            return [to_gpu(t) for t in v]
        return v
    gpu_kwargs = {k: to_gpu(v) for k, v in cpu_kwargs.items()}
    getattr(emu_ops, "build_" + name)(**cpu_kwargs).run(None)
    getattr(ops, "build_" + name)(**gpu_kwargs).run(ops.stream_handle())
    torch.cuda.synchronize()
    for k in out_keys:
        a, b = cpu_kwargs[k], gpu_kwargs[k]
        pairs = zip(a, b) if isinstance(a, list) else [(a, b)]
        for i, (x, y) in enumerate(pairs):
            x, y = x.float(), y.float().cpu()
            err = (x - y).abs().max().item() / (x.abs().max().item() + 1e-9)
            assert err < tol, (name, k, i, err)


@pytest.mark.parametrize("T", [64, 1000])
def test_layernorm_fwd_bwd(T):
    C, Cp = 180, 192
    g = _g(T)
    x = torch.zeros(T, Cp); x[:, :C] = torch.randn(T, C, generator=g) * 2 + 0.5
    gamma, beta = 1 + 0.2 * torch.randn(C, generator=g), 0.1 * torch.randn(C, generator=g)
    kw = dict(x=x, gamma=gamma, beta=beta, C=C, y_f32=torch.zeros(T, Cp), y_bf16=torch.zeros(T, Cp).bfloat16(),
              xhat_bf16=torch.zeros(T, Cp).bfloat16(), mean=torch.zeros(T), rstd=torch.zeros(T), ones_col=C)
    _run_both("layernorm_fwd", kw, ["y_f32", "mean", "rstd"], 1e-5)
    _run_both("layernorm_fwd", kw, ["y_bf16", "xhat_bf16"], 1e-2)
    dy = torch.zeros(T, Cp); dy[:, :C] = torch.randn(T, C, generator=g)
    dyb = torch.zeros(T, Cp).bfloat16(); dyb[:, :C] = torch.randn(T, C, generator=g).bfloat16()
    add = torch.zeros(T, Cp); add[:, :C] = torch.randn(T, C, generator=g)
    kw2 = dict(x=x, mean=kw["mean"], rstd=kw["rstd"], gamma=gamma, dgamma=torch.zeros(C), dbeta=torch.zeros(C), C=C,
               dy_f32=dy, dy_bf16=dyb, dx_add=add, dx_f32=torch.zeros(T, Cp), dx_bf16=torch.zeros(T, Cp).bfloat16())
    _run_both("layernorm_bwd", kw2, ["dx_f32", "dgamma", "dbeta"], 1e-4)
    _run_both("layernorm_bwd", dict(kw2, dgamma=torch.zeros(C), dbeta=torch.zeros(C)), ["dx_bf16"], 1e-2)


@pytest.mark.parametrize("ci,B,H,W", [(1, 2, 16, 24), (3, 1, 9, 13), (4, 1, 8, 8)])
def test_conv_in(ci, B, H, W):
    O, Op = 180, 192
    g = _g(ci * 7 + H)
    T = B * H * W
    img = torch.rand(B, ci, H, W, generator=g)
    Wt = torch.randn(O, ci, 3, 3, generator=g) * 0.3
    b = torch.randn(O, generator=g) * 0.1
    mean = torch.rand(ci, generator=g) * 0.5
    kw = dict(img=img, W=Wt, b=b, mean=mean, rng=255.0, O=O, O_pad=Op, pitch=Op, out_f32=torch.zeros(T, Op),
              out_bf16=torch.zeros(T, Op).bfloat16(), act=emu_ops.ACT_LRELU, act_slope=0.0, ones_col=O)
    _run_both("conv_in_fwd", kw, ["out_f32"], 1e-5)
    _run_both("conv_in_fwd", kw, ["out_bf16"], 1e-2)
    dy = torch.zeros(T, Op); dy[:, :O] = torch.randn(T, O, generator=g)
    kw2 = dict(img=img, W=Wt, mean=mean, rng=2.0, dW=torch.zeros(O, ci, 3, 3), db=torch.zeros(O), O=O, O_pad=Op,
               pitch=Op, dy_f32=dy)
    _run_both("conv_in_wgrad", kw2, ["dW", "db"], 1e-4)


@pytest.mark.parametrize("co,B,H,W,C,Cp", [(1, 2, 16, 24, 180, 192), (3, 1, 9, 13, 180, 192), (1, 1, 40, 70, 64, 64),
                                           (3, 2, 33, 37, 60, 64), (4, 1, 8, 8, 48, 64), (2, 1, 5, 3, 200, 256)])
def test_conv_out(co, B, H, W, C, Cp):
    g = _g(co * 5 + W)
    T = B * H * W
    feat = torch.zeros(T, Cp).bfloat16(); feat[:, :C] = torch.randn(T, C, generator=g).bfloat16()
    if Cp > C:
        feat[:, C] = 1.0
    Wt = torch.randn(co, C, 3, 3, generator=g) * 0.1
    b = torch.randn(co, generator=g)
    skip = torch.rand(B, co, H, W, generator=g)
    mean = torch.rand(co, generator=g)
    kw = dict(feat=feat, W=Wt, b=b, out=torch.zeros(B, co, H, W), out_mean=mean, C=C, pitch=Cp, out_scale=0.5,
              skip=skip, skip_mean=mean, skip_scale=2.0)
    # tensor-core path: fp32 weights as two bf16 terms (~16 mantissa bits), fp32 accumulation in MMA order
    _run_both("conv_out_fwd", kw, ["out"], 3e-5)
    dout = torch.randn(B, co, H, W, generator=g)
    kw2 = dict(dout=dout, W=Wt, C=C, pitch=Cp, out_scale=0.5, dfeat_f32=torch.zeros(T, Cp),
               dfeat_bf16=torch.zeros(T, Cp).bfloat16())
    _run_both("conv_out_dgrad", kw2, ["dfeat_f32"], 1e-5)
    _run_both("conv_out_dgrad", kw2, ["dfeat_bf16"], 1e-2)
    kw3 = dict(dout=dout, feat=feat, W=Wt, dW=torch.zeros(co, C, 3, 3), db=torch.zeros(co), C=C, pitch=Cp,
               out_scale=0.5)
    _run_both("conv_out_wgrad", kw3, ["dW", "db"], 1e-4)


def test_pack_unpack_linear_qkv_layout():
    from kair_b200.swinir_engine import SwinIREngine, SwinIRSpec
    C, L = 180, 3
    g = _g(3)
    spec = SwinIRSpec(in_chans=1, embed_dim=C, depths=[2], num_heads=[6], window_size=8, mlp_ratio=2, qk_scale=None,
                      img_range=1.0, upscale=1, upsampler="", resi_connection="1conv", patch_norm=True, qkv_bias=True)
    eng = SwinIREngine.__new__(SwinIREngine)
    eng.spec, eng.device = spec, torch.device("cpu")
    eng.C, eng.nH, eng.hd, eng.Cp, eng.HP, eng.QP, eng.Fp, eng.scale = C, 6, 30, 192, 192, 576, 384, 30 ** -0.5
    eng._build_maps()
    W = [torch.randn(3 * C, C, generator=g) for _ in range(L)]
    b = [torch.randn(3 * C, generator=g) for _ in range(L)]
    gam = [1 + 0.3 * torch.randn(C, generator=g) for _ in range(L)]
    bet = [0.3 * torch.randn(C, generator=g) for _ in range(L)]
    kw = dict(W=W, b=b, gamma=gam, beta=bet, row_map=eng.map_qkv, col_map=eng.map_c, row_scale=eng.scale_qkv,
              bias_const=eng.const_qkv, Wp=torch.zeros(L, 576, 192).bfloat16(), WpT=torch.zeros(L, 192, 576).bfloat16(),
              bp=torch.zeros(L, 576), out_f=3 * C, in_f=C)
    _run_both("pack_linear", kw, ["bp"], 1e-5)
    _run_both("pack_linear", kw, ["Wp", "WpT"], 1e-2)
    dWp = torch.randn(L, 576, 192, generator=g)
    z = lambda *s: [torch.zeros(*s) for _ in range(L)]
    kw2 = dict(dWp=dWp, W=W, gamma=gam, beta=bet, dW=z(3 * C, C), db=z(3 * C), dgamma=z(C), dbeta=z(C),
               row_map=eng.map_qkv, col_map=eng.map_c, row_scale=eng.scale_qkv, out_f=3 * C, in_f=C, ones_col=C)
    _run_both("unpack_linear_grad", kw2, ["dW", "db", "dgamma", "dbeta"], 1e-4)
    # proj-style (no LayerNorm fold, head-padded K)
    Wq = [torch.randn(C, C, generator=g) for _ in range(L)]
    kw3 = dict(dWp=torch.randn(L, 192, 192, generator=g), W=Wq, gamma=None, beta=None, dW=z(C, C), db=z(C), dgamma=None,
               dbeta=None, row_map=eng.map_c, col_map=eng.map_h, row_scale=None, out_f=C, in_f=C, ones_col=30)
    _run_both("unpack_linear_grad", kw3, ["dW", "db"], 1e-5)


def test_pack_unpack_conv():
    L, O, I, Ip = 2, 180, 180, 192
    g = _g(4)
    W = [torch.randn(O, I, 3, 3, generator=g) for _ in range(L)]
    b = [torch.randn(O, generator=g) for _ in range(L)]
    kw = dict(W=W, b=b, Wp=torch.zeros(L, 192, 9 * Ip).bfloat16(), WpT=torch.zeros(L, Ip, 9 * 192).bfloat16(),
              bp=torch.zeros(L, 192), O=O, I=I, I_pad=Ip)
    _run_both("pack_conv", kw, ["Wp", "WpT", "bp"], 1e-2)
    kw2 = dict(dWp=torch.randn(L, 192, 9 * Ip, generator=g), dW=[torch.zeros(O, I, 3, 3) for _ in range(L)],
               db=[torch.zeros(O) for _ in range(L)], O=O, I=I, I_pad=Ip, ones_col=180, center_tap=4)
    _run_both("unpack_conv_grad", kw2, ["dW", "db"], 1e-6)


def test_add_inplace():
    g = _g(5)
    y, x = torch.randn(1000, 192, generator=g), torch.randn(1000, 192, generator=g)
    kw = dict(y=y, x=x, yb=torch.zeros(1000, 192).bfloat16())
    _run_both("add_inplace", kw, ["y"], 1e-6)


@pytest.mark.parametrize("c,direction", [(1, 1), (3, -1), (4, 1)])
def test_im2col3x3(c, direction):
    g = _g(6)
    img = torch.randn(2, c, 24, 40, generator=g)
    kw = dict(img=img, out=torch.zeros(2 * 24 * 40, 64).bfloat16(), cpad=7, direction=direction,
              mean=torch.randn(c, generator=g), scale=1.7, ones_col=4 * 7 + c)
    _run_both("im2col3x3", kw, ["out"], 1e-2)


def test_conv_out_pack_roundtrip():
    g = _g(7)
    co, C = 3, 180
    W = torch.randn(co, C, 3, 3, generator=g)
    _run_both("conv_out_pack", dict(mode=0, co=co, C=C, cpad=7, W=W, Wt=torch.zeros(192, 64).bfloat16()), ["Wt"], 1e-2)
    D = torch.randn(192, 64, generator=g)
    _run_both("conv_out_pack", dict(mode=1, co=co, C=C, cpad=7, D=D, dW=torch.zeros(co, C, 3, 3), db=torch.zeros(co),
                                    ones_row=C), ["dW", "db"], 1e-6)
