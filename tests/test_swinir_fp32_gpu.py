"""fp32-accurate mode on the CUDA path (through the C ABI): every kernel the mode adds or re-parameterises against its
statement in tests/emu_ops.py, then the drop-in module against the oracle at north_star's fp32-mode bounds --
forward <= 1e-4 max-abs, every parameter gradient <= 1e-3 relative (relative L2 and max-abs error over max-abs
reference).  Full-depth cfg2 / cfg3 / cfg5 in this mode: tests/test_parity_full_gpu.py."""
import pytest
import torch

import emu_ops
import golden_cases as gc
from test_elementwise_gpu import _run_both

pytestmark = pytest.mark.gpu


def _g(seed):
    return torch.Generator().manual_seed(seed)


@pytest.mark.parametrize("terms", [1, 3, 6])
@pytest.mark.parametrize("op", [emu_ops.SPLIT_COPY, emu_ops.SPLIT_GELU, emu_ops.SPLIT_DGELU, emu_ops.SPLIT_ACT_BWD,
                                emu_ops.SPLIT_LRELU])
def test_split_terms(terms, op):
    T, n, pitch = 300, 192, 200
    g = _g(10 * terms + op)
    x = torch.randn(T, pitch, generator=g) * 3
    aux = torch.randn(T, pitch, generator=g) * 2
    rs = torch.rand(3, generator=g) + 0.5
    kw = dict(x=x, out=torch.zeros(T, terms * n + 8).bfloat16(), n=n, terms=terms, op=op, aux=aux,
              out_f32=torch.zeros(T, pitch), row_scale=rs, rows_per_sample=100, ones_col=181, slope=0.2)
    _run_both("split_terms", kw, ["out_f32"], 2e-6)
    # the components are exact functions of the fp32 value: compare the reconstructed sums, and the leading segment bitwise
    from kair_b200 import ops
    xg = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in kw.items()}
    xg["out"] = torch.zeros(T, terms * n + 8, device="cuda").bfloat16()
    xg["out_f32"] = torch.zeros(T, pitch, device="cuda")
    ops.split_terms(**xg)
    v = xg["out_f32"][:, :n]
    segs = [xg["out"][:, s * n:(s + 1) * n].float() for s in range(terms)]
    assert torch.equal(segs[0], v.bfloat16().float())
    nc = {1: 1, 3: 2, 6: 3}[terms]
    recon = sum(segs[k] for k in range(nc))
    tol = {1: 2.0 ** -8, 3: 2.0 ** -16, 6: 2.0 ** -23}[terms]
    assert ((recon - v).abs() <= tol * v.abs() + 1e-30).all()
    if terms == 3:
        assert torch.equal(segs[2], segs[0])
    if terms == 6:
        assert torch.equal(segs[3], segs[0]) and torch.equal(segs[5], segs[0]) and torch.equal(segs[4], segs[1])


@pytest.mark.parametrize("terms", [3, 6])
def test_layernorm_split_and_ln_bwd_split(terms):
    T, C, Cp = 500, 180, 192
    g = _g(terms)
    x = torch.zeros(T, Cp); x[:, :C] = torch.randn(T, C, generator=g) * 2 + 0.3
    kw = dict(x=x, gamma=None, beta=None, C=C, xhat_bf16=torch.zeros(T, terms * Cp).bfloat16(), rstd=torch.zeros(T),
              ones_col=C, terms=terms)
    _run_both("layernorm_fwd", kw, ["rstd"], 1e-5)
    _run_both("layernorm_fwd", kw, ["xhat_bf16"], 1e-2)
    gamma, beta = 1 + 0.2 * torch.randn(C, generator=g), 0.1 * torch.randn(C, generator=g)
    kwy = dict(x=x, gamma=gamma, beta=beta, C=C, y_bf16=torch.zeros(T, terms * Cp).bfloat16(), mean=torch.zeros(T),
               rstd=torch.zeros(T), ones_col=C, terms=terms)
    _run_both("layernorm_fwd", kwy, ["y_bf16"], 1e-2)
    gg = torch.zeros(T, Cp); gg[:, :C] = torch.randn(T, C, generator=g)
    resid = torch.zeros(T, Cp); resid[:, :C] = torch.randn(T, C, generator=g)
    kw2 = dict(g=gg, xhat=kw["xhat_bf16"], rstd=kw["rstd"], outF=torch.zeros(T, Cp), C=C, terms=terms, resid=resid,
               out=torch.zeros(T, terms * Cp).bfloat16(), row_scale=torch.rand(5, generator=g) + 0.5, rows_per_sample=100)
    _run_both("ln_bwd_split", kw2, ["outF"], 1e-5)
    _run_both("ln_bwd_split", kw2, ["out"], 1e-2)


@pytest.mark.parametrize("terms", [3, 6])
def test_pack_linear_and_conv_terms(terms):
    from kair_b200 import ops
    g = _g(20 + terms)
    L, out_f, in_f, N_pad, K_pad = 2, 180, 180, 192, 192
    W = [torch.randn(out_f, in_f, generator=g) * 0.1 for _ in range(L)]
    b = [torch.randn(out_f, generator=g) * 0.1 for _ in range(L)]
    gamma = [1 + 0.1 * torch.randn(in_f, generator=g) for _ in range(L)]
    beta = [0.1 * torch.randn(in_f, generator=g) for _ in range(L)]
    rm = torch.tensor([r if r < out_f else -1 for r in range(N_pad)], dtype=torch.int32)
    cm = torch.tensor([c if c < in_f else -1 for c in range(K_pad)], dtype=torch.int32)
    kw = dict(W=W, b=b, gamma=gamma, beta=beta, row_map=rm, col_map=cm, row_scale=None, bias_const=None,
              Wp=torch.zeros(L, N_pad, terms * K_pad).bfloat16(), WpT=torch.zeros(L, K_pad, terms * N_pad).bfloat16(),
              bp=torch.zeros(L, N_pad), out_f=out_f, in_f=in_f, terms=terms)
    _run_both("pack_linear", kw, ["bp"], 1e-5)
    kwg = {k: ([t.cuda() for t in v] if isinstance(v, list) else (v.cuda() if isinstance(v, torch.Tensor) else v))
           for k, v in kw.items()}
    emu_ops.build_pack_linear(**kw).run()
    ops.pack_linear(**kwg)
    assert torch.equal(kwg["Wp"].cpu(), kw["Wp"]) and torch.equal(kwg["WpT"].cpu(), kw["WpT"])
    O, I, O_pad, I_pad = 180, 180, 192, 192
    Wc = [torch.randn(O, I, 3, 3, generator=g) * 0.05 for _ in range(L)]
    bc = [torch.randn(O, generator=g) * 0.1 for _ in range(L)]
    kc = dict(W=Wc, b=bc, Wp=torch.zeros(L, O_pad, 9 * terms * I_pad).bfloat16(),
              WpT=torch.zeros(L, I_pad, 9 * terms * O_pad).bfloat16(), bp=torch.zeros(L, O_pad), O=O, I=I, I_pad=I_pad,
              terms=terms)
    kcg = {k: ([t.cuda() for t in v] if isinstance(v, list) else (v.cuda() if isinstance(v, torch.Tensor) else v))
           for k, v in kc.items()}
    emu_ops.build_pack_conv(**kc).run()
    ops.pack_conv(**kcg)
    assert torch.equal(kcg["Wp"].cpu(), kc["Wp"]) and torch.equal(kcg["WpT"].cpu(), kc["WpT"])
    assert torch.equal(kcg["bp"].cpu(), kc["bp"])


def test_split_gemm_reaches_fp32_accuracy():
    """one kair_gemm_tn launch over split operands vs an fp64 product: ~2^-17 (terms 3) / fp32-accumulate level (6)"""
    from kair_b200 import ops
    g = _g(5)
    M, K, N = 512, 192, 192
    X = torch.randn(M, K, generator=g)
    Wm = torch.randn(N, K, generator=g) * 0.1
    ref = (X.double() @ Wm.double().t()).float()
    errs = {}
    for terms in (1, 3, 6):
        A = emu_ops._split_act(X, terms).cuda() if terms > 1 else X.bfloat16().cuda()
        c = emu_ops._components(Wm)
        B = torch.cat([c[w] for w in emu_ops._seg_w(terms)], 1).bfloat16().cuda()
        out = torch.zeros(M, N, device="cuda")
        ops.gemm_tn(A, B, N=N, K=terms * K, bn=192, outF=out)
        errs[terms] = ((out.cpu() - ref).abs().max() / ref.abs().max()).item()
    assert errs[1] > 1e-3 and errs[3] < 3e-5 and errs[6] < 6e-6, errs   # terms 6: the fp32 accumulation limit


@pytest.mark.parametrize("ws,shift,B,H,W,nH", [(8, 0, 2, 16, 24, 6), (8, 4, 1, 24, 24, 6), (7, 3, 1, 14, 21, 6),
                                               (8, 4, 2, 32, 16, 8)])
def test_window_attention_fp32(ws, shift, B, H, W, nH):
    T = B * H * W
    g = _g(ws * 10 + shift + H)
    qkv = torch.randn(T, 3 * nH * 32, generator=g) * 0.7
    table = torch.randn((2 * ws - 1) ** 2, nH, generator=g) * 0.5
    kw = dict(qkv=qkv, table=table, out=torch.zeros(T, nH * 32), batch=B, H=H, W=W, num_heads=nH, window=ws, shift=shift)
    _run_both("window_attn_fwd", kw, ["out"], 2e-6)
    dout = torch.randn(T, nH * 32, generator=g)
    kb = dict(qkv=qkv, table=table, dout=dout, dqkv=torch.zeros(T, 3 * nH * 32), dtable=torch.zeros_like(table),
              batch=B, H=H, W=W, num_heads=nH, window=ws, shift=shift)
    _run_both("window_attn_bwd", kb, ["dqkv", "dtable"], 2e-5)


def test_thin_and_shuffle_kernels_on_fp32_features():
    g = _g(77)
    B, H, W, C, Cp, co = 2, 12, 20, 180, 192, 3
    T = B * H * W
    feat = torch.zeros(T, Cp); feat[:, :C] = torch.randn(T, C, generator=g)
    Wt = torch.randn(co, C, 3, 3, generator=g) * 0.05
    b = torch.randn(co, generator=g) * 0.1
    mean = torch.rand(co, generator=g)
    skip = torch.rand(B, co, H, W, generator=g)
    kw = dict(feat=feat, W=Wt, b=b, out=torch.zeros(B, co, H, W), out_mean=mean, C=C, pitch=Cp, out_scale=0.5,
              skip=skip, skip_mean=mean, skip_scale=2.0)
    _run_both("conv_out_fwd", kw, ["out"], 2e-6)
    dout = torch.randn(B, co, H, W, generator=g)
    kw2 = dict(dout=dout, feat=feat, W=Wt, dW=torch.zeros(co, C, 3, 3), db=torch.zeros(co), C=C, pitch=Cp, out_scale=0.5)
    _run_both("conv_out_wgrad", kw2, ["dW", "db"], 2e-5)
    for r, nf in ((2, 64), (3, 64)):
        x = torch.randn(T, r * r * nf, generator=g)
        kw3 = dict(x=x, out=torch.zeros(T * r * r, nf), batch=B, H=H, W=W, C=nf, r=r)
        _run_both("pixel_shuffle", kw3, ["out"], 0.0 + 1e-12)
    x = torch.randn(T, 64, generator=g)
    _run_both("upsample_nearest2", dict(x=x, out=torch.zeros(4 * T, 64), batch=B, H=H, W=W, C=64), ["out"], 1e-12)
    x = torch.randn(T, 64, generator=g)
    _run_both("pixel_shuffle_out", dict(x=x, out=torch.zeros(B, 3, 2 * H, 2 * W), mean=mean, r=2, scale=0.7), ["out"], 1e-6)


def _rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return ((a - b).norm() / (b.norm() + 1e-30)).item(), ((a - b).abs().max() / (b.abs().max() + 1e-30)).item()


@pytest.mark.parametrize("terms", [3, 6])
def test_module_fp32_forward_backward_vs_oracle_and_golden(terms):
    """SwinIR-M geometry at depth 2+2 (the golden fixture of the real reference module): 1e-4 / 1e-3"""
    import os
    import numpy as np
    from kair_b200.network_swinir import SwinIR
    case = gc.CASES["swinir_m_gray32"]
    sd, x, tgt = gc.state_and_io(case)
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    y_ref = gc.oracle_forward(case, sdg, x)
    gc.loss_fn(case)(y_ref, tgt).backward()
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().train().set_precision("fp32", terms)
    y = model(x.cuda())
    gc.loss_fn(case)(y, tgt.cuda()).backward()
    torch.cuda.synchronize()
    assert (y.detach().cpu() - y_ref.detach()).abs().max().item() < 1e-4
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "swinir_m_gray32.npz"))
    assert np.abs(y.detach().cpu().numpy() - z["y"]).max() < 1e-4
    bad = []
    for k, p in model.named_parameters():
        r, m = _rel(p.grad.cpu(), sdg[k].grad)
        if not (r < 1e-3 and m < 1e-3):
            bad.append((k, r, m))
    assert not bad, bad[:8]


@pytest.mark.parametrize("name", ["swinir_sr_ps4", "swinir_sr_nearest4", "swinir_s_psd2", "swinir_m_rgb_pad",
                                  "swinir_jpeg_ws7"])
def test_module_fp32_inference_heads(name):
    import os
    import numpy as np
    from kair_b200.network_swinir import SwinIR
    case = gc.CASES[name]
    sd, x, tgt = gc.state_and_io(case)
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().eval().set_precision("fp32")
    with torch.no_grad():
        y = model(x.cuda()).cpu()
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    yr = torch.from_numpy(z["y"])
    assert y.shape == yr.shape
    assert (y - yr).abs().max().item() < 1e-4
    assert abs(gc.psnr_uint8(y, tgt) - gc.psnr_uint8(yr, tgt)) < 1e-3


def test_module_fp32_drop_path_and_bf16_switch():
    """the two precisions share one module: switching back and forth keeps both plan sets valid"""
    from kair_b200.network_swinir import SwinIR
    case = gc.CASES["swinir_m_gray32"]
    sd, x, tgt = gc.state_and_io(case)
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().eval()
    with torch.no_grad():
        yb = model(x.cuda())
        yf = model.set_precision("fp32")(x.cuda())
        yb2 = model.set_precision("bf16")(x.cuda())
        y_ref = gc.oracle_forward(case, sd, x)
    assert torch.equal(yb, yb2)
    eb, ef = (yb.cpu() - y_ref).abs().max().item(), (yf.cpu() - y_ref).abs().max().item()
    assert ef < 1e-4 and ef < eb / 20, (eb, ef)
