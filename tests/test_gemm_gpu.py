"""tcgen05 GEMM family vs plain PyTorch fp32 on the same (bf16-rounded) inputs."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda:0")


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12)).item()


def _rand(shape, g, scale=1.0):
    return (torch.randn(shape, generator=g, device="cpu") * scale).to(_dev())


@pytest.mark.parametrize("M", [128, 1024, 64 * 37])
@pytest.mark.parametrize("N,K,bn", [(192, 192, 192), (576, 192, 192), (192, 384, 192), (192, 576, 192), (64, 64, 64),
                                    (256, 128, 256), (128, 64, 128)])
def test_tn_linear(M, N, K, bn):
    from kair_b200 import ops
    g = torch.Generator().manual_seed(M * 7 + N + K)
    A = _rand((M, K), g).bfloat16()
    B = _rand((N, K), g, 0.1).bfloat16()
    bias = _rand((N,), g)
    outB = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
    outF = torch.empty(M, N, dtype=torch.float32, device=_dev())
    ops.gemm_tn(A, B, N=N, K=K, bn=bn, bias=bias, outB=outB, outF=outF)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().t() + bias
    assert _rel(outF, ref) < 2e-5
    assert _rel(outB, ref) < 6e-3


def test_tn_linear_resid_act_ones():
    from kair_b200 import ops
    M, N, K = 640, 192, 192
    g = torch.Generator().manual_seed(5)
    A = _rand((M, K), g).bfloat16()
    B = _rand((N, K), g, 0.1).bfloat16()
    bias = _rand((N,), g)
    resid = _rand((M, N), g)
    outB = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
    outF = torch.empty(M, N, dtype=torch.float32, device=_dev())
    ops.gemm_tn(A, B, N=N, K=K, bn=192, bias=bias, resid=resid, act=ops.ACT_LRELU, act_slope=0.2, outB=outB,
                outF=outF, ones_col=180)
    torch.cuda.synchronize()
    ref = torch.nn.functional.leaky_relu(A.float() @ B.float().t() + bias + resid, 0.2)
    assert _rel(outF, ref) < 2e-5
    refb = ref.clone()
    refb[:, 180] = 1.0
    assert _rel(outB, refb) < 6e-3
    assert (outB[:, 180].float() == 1.0).all()


def test_tn_gelu():
    from kair_b200 import ops
    M, N, K = 1152, 384, 192
    g = torch.Generator().manual_seed(6)
    A = _rand((M, K), g).bfloat16()
    B = _rand((N, K), g, 0.1).bfloat16()
    bias = _rand((N,), g)
    u = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
    gl = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
    ops.gemm_tn(A, B, N=N, K=K, bn=192, mode=ops.EPI_GELU, bias=bias, outB=gl, outB2=u, ones_col=360)
    torch.cuda.synchronize()
    ref_u = A.float() @ B.float().t() + bias
    ref_g = torch.nn.functional.gelu(ref_u)
    ref_g[:, 360] = 1.0
    assert _rel(u, ref_u) < 6e-3
    assert _rel(gl, ref_g) < 6e-3


@pytest.mark.parametrize("M", [256, 64 * 21])
def test_tn_resid_ln(M):
    from kair_b200 import ops
    N, K, C = 192, 384, 180
    g = torch.Generator().manual_seed(7)
    A = _rand((M, K), g).bfloat16()
    B = _rand((N, K), g, 0.1).bfloat16()
    B[C:] = 0
    bias = _rand((N,), g)
    bias[C:] = 0
    resid = _rand((M, N), g) + 0.5
    resid[:, C:] = 0
    outF = torch.empty(M, N, dtype=torch.float32, device=_dev())
    xhat = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
    rstd = torch.empty(M, dtype=torch.float32, device=_dev())
    ops.gemm_tn(A, B, N=N, K=K, bn=192, mode=ops.EPI_RESID_LN, bias=bias, resid=resid, outF=outF, outB=xhat,
                rowstat=rstd, n_valid=C, ones_col=C, ln_eps=1e-5)
    torch.cuda.synchronize()
    v = A.float() @ B.float().t() + bias + resid
    assert _rel(outF, v) < 2e-5
    vc = v[:, :C]
    mu = vc.mean(1, keepdim=True)
    var = vc.var(1, unbiased=False, keepdim=True)
    ref_rstd = (var + 1e-5).rsqrt()
    ref_xhat = torch.zeros_like(v)
    ref_xhat[:, :C] = (vc - mu) * ref_rstd
    ref_xhat[:, C] = 1.0
    assert _rel(rstd, ref_rstd.squeeze(1)) < 1e-4
    assert (xhat.float() - ref_xhat).abs().max().item() < 3e-2
    assert (xhat[:, C + 1:].float() == 0).all() and (xhat[:, C].float() == 1).all()


def test_tn_dgelu():
    from kair_b200 import ops
    M, N, K = 768, 384, 192
    g = torch.Generator().manual_seed(8)
    A = _rand((M, K), g).bfloat16()
    B = _rand((N, K), g, 0.1).bfloat16()
    u = _rand((M, N), g).bfloat16()
    out = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
    ops.gemm_tn(A, B, N=N, K=K, bn=192, mode=ops.EPI_DGELU, aux=u, outB=out)
    torch.cuda.synchronize()
    uu = u.float().requires_grad_(True)
    torch.nn.functional.gelu(uu).sum().backward()
    ref = (A.float() @ B.float().t()) * uu.grad
    assert _rel(out, ref) < 6e-3


def test_tn_lnbwd():
    from kair_b200 import ops
    M, N, K, C = 896, 192, 384, 180
    g = torch.Generator().manual_seed(9)
    A = _rand((M, K), g).bfloat16()
    B = _rand((N, K), g, 0.1).bfloat16()
    B[C:] = 0
    x = _rand((M, C), g) * 2 + 0.3
    mu = x.mean(1, keepdim=True)
    rstd = (x.var(1, unbiased=False, keepdim=True) + 1e-5).rsqrt()
    xhat = torch.zeros(M, N, device=_dev())
    xhat[:, :C] = (x - mu) * rstd
    xhat[:, C] = 1.0
    xhat_b = xhat.bfloat16()
    dres = _rand((M, N), g)
    dres[:, C:] = 0
    outF = torch.empty(M, N, dtype=torch.float32, device=_dev())
    outB = torch.empty(M, N, dtype=torch.bfloat16, device=_dev())
    ops.gemm_tn(A, B, N=N, K=K, bn=192, mode=ops.EPI_LNBWD, aux=xhat_b, resid=dres, rowstat=rstd.squeeze(1).contiguous(),
                outF=outF, outB=outB, n_valid=C)
    torch.cuda.synchronize()
    dxh = (A.float() @ B.float().t())[:, :C]
    xh = xhat_b.float()[:, :C]
    ref = torch.zeros(M, N, device=_dev())
    ref[:, :C] = dres[:, :C] + rstd * (dxh - dxh.mean(1, keepdim=True) - xh * (dxh * xh).mean(1, keepdim=True))
    assert _rel(outF, ref) < 5e-5
    assert _rel(outB, ref) < 6e-3
    assert (outF[:, C:] == 0).all()


@pytest.mark.parametrize("Bt,H,W,Cin,Cout,bn", [(2, 16, 16, 64, 64, 64), (1, 128, 128, 192, 192, 192),
                                                (2, 24, 40, 192, 192, 192), (1, 32, 64, 128, 256, 256),
                                                (3, 8, 8, 64, 128, 128)])
def test_tn_conv3x3(Bt, H, W, Cin, Cout, bn):
    from kair_b200 import ops
    g = torch.Generator().manual_seed(H * W + Cin)
    x = _rand((Bt, H, W, Cin), g).bfloat16()
    w = _rand((Cout, Cin, 3, 3), g, 0.05).bfloat16()
    bias = _rand((Cout,), g)
    wp = w.permute(0, 2, 3, 1).reshape(Cout, 9 * Cin).contiguous()  # [o][tap][ci]
    outB = torch.empty(Bt, H, W, Cout, dtype=torch.bfloat16, device=_dev())
    outF = torch.empty(Bt, H, W, Cout, dtype=torch.float32, device=_dev())
    ops.gemm_tn(x, wp, N=Cout, K=Cin, bn=bn, bias=bias, act=ops.ACT_LRELU, act_slope=0.0, outB=outB, outF=outF,
                conv=(Bt, H, W))
    torch.cuda.synchronize()
    ref = torch.relu(torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float(), bias, padding=1))
    ref = ref.permute(0, 2, 3, 1)
    assert _rel(outF, ref) < 3e-5
    assert _rel(outB, ref) < 6e-3


def test_tn_conv3x3_resid_ln():
    from kair_b200 import ops
    Bt, H, W, C, Cp = 2, 16, 24, 180, 192
    g = torch.Generator().manual_seed(11)
    x = torch.zeros(Bt, H, W, Cp, device=_dev())
    x[..., :C] = _rand((Bt, H, W, C), g)
    xb = x.bfloat16()
    w = torch.zeros(Cp, Cp, 3, 3, device=_dev())
    w[:C, :C] = _rand((C, C, 3, 3), g, 0.05)
    wb = w.bfloat16()
    bias = torch.zeros(Cp, device=_dev())
    bias[:C] = _rand((C,), g)
    resid = torch.zeros(Bt, H, W, Cp, device=_dev())
    resid[..., :C] = _rand((Bt, H, W, C), g)
    wp = wb.permute(0, 2, 3, 1).reshape(Cp, 9 * Cp).contiguous()
    outF = torch.empty(Bt, H, W, Cp, dtype=torch.float32, device=_dev())
    xhat = torch.empty(Bt, H, W, Cp, dtype=torch.bfloat16, device=_dev())
    rstd = torch.empty(Bt * H * W, dtype=torch.float32, device=_dev())
    ops.gemm_tn(xb, wp, N=Cp, K=Cp, bn=192, mode=ops.EPI_RESID_LN, bias=bias, resid=resid, outF=outF, outB=xhat,
                rowstat=rstd, n_valid=C, ones_col=C, conv=(Bt, H, W))
    torch.cuda.synchronize()
    v = torch.nn.functional.conv2d(xb.float().permute(0, 3, 1, 2), wb.float(), bias, padding=1).permute(0, 2, 3, 1)
    v = v + resid
    assert _rel(outF, v) < 3e-5
    ref_xhat = torch.nn.functional.layer_norm(v[..., :C], (C,), eps=1e-5)
    assert (xhat[..., :C].float() - ref_xhat).abs().max().item() < 3e-2


@pytest.mark.parametrize("T,n1,n2,bn", [(64, 192, 192, 192), (4096, 576, 192, 192), (64 * 33, 192, 384, 192),
                                        (8192, 384, 192, 192), (1024, 64, 64, 64), (2048, 256, 256, 256)])
def test_nt_dense(T, n1, n2, bn):
    from kair_b200 import ops
    g = torch.Generator().manual_seed(T + n1)
    A = _rand((T, n1), g).bfloat16()
    B = _rand((T, n2), g).bfloat16()
    D = torch.zeros(n1, n2, dtype=torch.float32, device=_dev())
    ops.gemm_nt(A, B, D, n1=n1, n2=n2, bn=bn)
    torch.cuda.synchronize()
    ref = A.float().t() @ B.float()
    assert _rel(D, ref) < 1e-4


@pytest.mark.parametrize("Bt,H,W,Cin,Cout,bn", [(2, 16, 16, 64, 64, 64), (1, 64, 64, 192, 192, 192),
                                                (2, 24, 40, 192, 192, 192)])
def test_nt_conv(Bt, H, W, Cin, Cout, bn):
    from kair_b200 import ops
    g = torch.Generator().manual_seed(H + W + Cin)
    x = _rand((Bt, H, W, Cin), g).bfloat16()
    dy = _rand((Bt, H, W, Cout), g).bfloat16()
    D = torch.zeros(Cout, 9 * Cin, dtype=torch.float32, device=_dev())
    ops.gemm_nt(dy, x, D, n1=Cout, n2=Cin, bn=bn, conv=(Bt, H, W))
    torch.cuda.synchronize()
    xx = x.float().permute(0, 3, 1, 2).requires_grad_(False)
    w = torch.zeros(Cout, Cin, 3, 3, device=_dev(), requires_grad=True)
    y = torch.nn.functional.conv2d(xx, w, padding=1)
    y.backward(dy.float().permute(0, 3, 1, 2))
    ref = w.grad.permute(0, 2, 3, 1).reshape(Cout, 9 * Cin)
    assert _rel(D, ref) < 1e-4
