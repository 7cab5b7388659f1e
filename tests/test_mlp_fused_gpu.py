"""kair_mlp_fused (fc1 + GELU + fc2 + DropPath + residual + LayerNorm in one launch, hidden activations on chip) against
the unfused pair of kair_gemm_tn launches it replaces -- bit-identical by construction (same rounding points, same
k-order) -- and against the fp32 PyTorch statement of Mlp.forward + residual + LayerNorm (network_swinir.py:24-30,
:277, :245) at the bf16 tolerance."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(M, N, C, hidden, seed):
    g = torch.Generator().manual_seed(seed)
    F = 2 * N
    X = torch.zeros(M, N); X[:, :C] = torch.randn(M, C, generator=g); X[:, C] = 1.0
    W1 = torch.zeros(F, N); W1[:hidden, :C] = torch.randn(hidden, C, generator=g) * 0.08
    W2 = torch.zeros(N, F); W2[:C, :hidden] = torch.randn(C, hidden, generator=g) * 0.08
    b1 = torch.zeros(F); b1[:hidden] = torch.randn(hidden, generator=g) * 0.1
    b2 = torch.zeros(N); b2[:C] = torch.randn(C, generator=g) * 0.1
    resid = torch.zeros(M, N); resid[:, :C] = torch.randn(M, C, generator=g)
    return X, W1, W2, b1, b2, resid


@pytest.mark.parametrize("M,N,C,hidden", [(128, 192, 180, 360), (1000, 192, 180, 360), (37 * 128 + 5, 192, 180, 360),
                                          (148 * 128 * 2 + 64, 192, 180, 360), (640, 64, 60, 120), (900, 256, 240, 480)])
@pytest.mark.parametrize("ln,train", [(True, False), (True, True), (False, False), (False, True)])
def test_mlp_fused_equals_unfused_pair(M, N, C, hidden, ln, train):
    from kair_b200 import ops
    F = 2 * N
    X, W1, W2, b1, b2, resid = _mk(M, N, C, hidden, M + N + ln + 2 * train)
    tps = 100
    scale = (torch.rand((M + tps - 1) // tps) < 0.8).float() / 0.8
    dev = "cuda"
    Xb, W1b, W2b = X.bfloat16().to(dev), W1.bfloat16().to(dev), W2.bfloat16().to(dev)
    b1, b2, resid, scale = b1.to(dev), b2.to(dev), resid.to(dev), scale.to(dev)
    bf = lambda *s: torch.zeros(*s, dtype=torch.bfloat16, device=dev)
    f32 = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
    # unfused pair
    u0, g0 = bf(M, F), bf(M, F)
    oF0, oB0, rs0 = (f32(M, N) if ln else None), bf(M, N), f32(M)
    ops.gemm_tn(Xb, W1b, N=F, K=N, bn=N if N != 256 else 256, mode=ops.EPI_GELU, bias=b1, outB=g0, outB2=u0, ones_col=hidden)
    if ln:
        ops.gemm_tn(g0, W2b, N=N, K=F, bn=N, mode=ops.EPI_RESID_LN, bias=b2, resid=resid, outF=oF0, outB=oB0, rowstat=rs0,
                    n_valid=C, ones_col=C, acc_scale=scale, tokens_per_sample=tps)
    else:
        ops.gemm_tn(g0, W2b, N=N, K=F, bn=N, mode=ops.EPI_LINEAR, bias=b2, resid=resid, outB=oB0, ones_col=C,
                    acc_scale=scale, tokens_per_sample=tps)
    # fused
    u1, g1 = (bf(M, F), bf(M, F)) if train else (None, None)
    oF1, oB1, rs1 = (f32(M, N) if ln else None), bf(M, N), f32(M)
    ops.mlp_fused(Xb, W1b, W2b, N=N, F=F, bias1=b1, bias2=b2, resid=resid, outF=oF1, outB=oB1, outU=u1, outG=g1,
                  rowstat=rs1 if ln else None, acc_scale=scale, tokens_per_sample=tps, ln=ln, n_valid=C, ones_col=C,
                  ones_col_g=hidden)
    torch.cuda.synchronize()
    assert torch.equal(oB0, oB1)
    if ln:
        assert torch.equal(oF0, oF1) and torch.equal(rs0, rs1)
    if train:
        assert torch.equal(u0, u1) and torch.equal(g0, g1)
    # fp32 statement of the reference arithmetic on the same (bf16-rounded) operands
    x_, w1_, w2_ = Xb.float()[:, :C], W1b.float()[:hidden, :C], W2b.float()[:C, :hidden]
    h = torch.nn.functional.gelu(x_ @ w1_.t() + b1[:hidden])
    v = (h @ w2_.t() + b2[:C]) * scale.repeat_interleave(tps)[:M, None] + resid[:, :C]
    if ln:
        assert (oF1[:, :C] - v).abs().max().item() < 2e-2 * max(1.0, v.abs().max().item())
        ref = torch.nn.functional.layer_norm(v, (C,))
        assert (oB1.float()[:, :C] - ref).abs().max().item() < 5e-2
    else:
        assert (oB1.float()[:, :C] - v).abs().max().item() < 3e-2 * max(1.0, v.abs().max().item())


def test_mlp_fused_plan_equals_unfused_plan():
    """whole-network check: the plans with and without the fused MLP half give identical outputs and gradients"""
    import golden_cases as gc
    from kair_b200 import swinir_engine as se
    from kair_b200.network_swinir import SwinIR
    case = gc.CASES["swinir_m_gray32"]
    sd, x, tgt = gc.state_and_io(case)
    res = {}
    for fuse in (True, False):
        model = SwinIR(drop_path_rate=0.0, **case["kw"])
        model.load_state_dict(sd, strict=True)
        model = model.cuda().train()
        model.engine(torch.device("cuda:0")).fuse_mlp = fuse
        y = model(x.cuda())
        gc.loss_fn(case)(y, tgt.cuda()).backward()
        plan = model.engine(torch.device("cuda:0")).plan(x.shape[0], 32, 32, True)
        assert plan.fuse_mlp == (fuse and se.FUSE_MLP)
        res[fuse] = (y.detach().clone(), {k: p.grad.clone() for k, p in model.named_parameters()})
    assert torch.equal(res[True][0], res[False][0])
    # the saved u / g are bit-identical, so the backward passes see the same operands; the split-K weight-gradient
    # GEMMs reduce with floating-point atomics in a run-dependent order, hence a 1e-5 band instead of equality
    for k in res[True][1]:
        a, b = res[True][1][k], res[False][1][k]
        assert (a - b).abs().max().item() <= 1e-5 * b.abs().max().item() + 1e-12, k
