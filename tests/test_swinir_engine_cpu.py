"""Host-side logic of the SwinIR engine (launch lists, packing maps, gradient routing), checked on CPU by replaying
the launch lists through tests/emu_ops.py and comparing with the unmodified reference module / the oracle."""
import pytest
import torch

import emu_ops
import ref_loader
from kair_b200.network_swinir import SwinIR
from kair_b200.swinir_engine import SwinIREngine, SwinIRFunction

CFG = dict(upscale=1, in_chans=1, img_size=16, window_size=8, img_range=1.0, depths=[2, 2], embed_dim=180,
           num_heads=[6, 6], mlp_ratio=2, upsampler="", resi_connection="1conv")


def _rel_l2(a, b):
    """bf16 tensor-core path vs fp32 reference: relative L2 error of a gradient tensor"""
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


def _loss(out, target):
    # smooth loss: with Charbonnier(eps=1e-9) dL/dy = sign(y - t)/n, whose sign flips under bf16 rounding
    return torch.mean((out - target) ** 2)


def _randomize(model, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for k, p in model.named_parameters():
            if k.endswith("norm1.weight") or k.endswith("norm2.weight") or k.endswith("norm.weight"):
                p.copy_(1.0 + 0.2 * torch.randn(p.shape, generator=g))
            elif k.endswith("bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif k.endswith("relative_position_bias_table"):
                p.copy_(0.3 * torch.randn(p.shape, generator=g))
            else:
                p.mul_(1.5)


def _engine_step(model, x, target, drop=None):
    params = dict(model.named_parameters())
    eng = SwinIREngine(model._spec, params, "cpu", ops=emu_ops)
    B, _, H, W = x.shape
    plan = eng.plan(B, H, W, True)
    if drop is not None:
        plan.dp.copy_(drop)
    out = SwinIRFunction.apply(x, plan, *params.values())
    loss = _loss(out, target)
    loss.backward()
    return out.detach(), {k: p.grad.clone() for k, p in params.items()}


@pytest.mark.skipif(not ref_loader.have_reference(), reason="reference tree not mounted")
@pytest.mark.parametrize("in_chans,H,W,C", [(1, 16, 16, 180), (3, 16, 24, 180), (3, 16, 16, 60)])
def test_engine_matches_reference_module(in_chans, H, W, C):
    ref_mod = ref_loader.load("network_swinir")
    cfg = dict(CFG, in_chans=in_chans, embed_dim=C)   # C=60: SwinIR-S (lightweight SR) geometry, head_dim 10
    torch.manual_seed(0)
    ref = ref_mod.SwinIR(drop_path_rate=0.0, **cfg)
    mine = SwinIR(drop_path_rate=0.0, **cfg)
    assert list(ref.state_dict().keys()) == list(mine.state_dict().keys())
    for (k1, v1), (k2, v2) in zip(ref.state_dict().items(), mine.state_dict().items()):
        assert v1.shape == v2.shape and v1.dtype == v2.dtype, k1
        if k1.endswith("attn_mask") or k1.endswith("relative_position_index"):
            assert torch.equal(v1, v2), k1
    _randomize(ref, 1)
    mine.load_state_dict(ref.state_dict(), strict=True)
    g = torch.Generator().manual_seed(2)
    x = torch.rand(2, in_chans, H, W, generator=g)
    tgt = torch.rand(2, in_chans, H, W, generator=g)
    ref.train()
    y_ref = ref(x)
    _loss(y_ref, tgt).backward()
    y, grads = _engine_step(mine, x, tgt)
    assert (y - y_ref.detach()).abs().max().item() < 2e-2
    for k, p in ref.named_parameters():
        gr = p.grad
        assert _rel_l2(grads[k], gr) < 5e-2, (k, _rel_l2(grads[k], gr))


def test_engine_matches_oracle_with_drop_path():
    from oracle import swinir_oracle as so
    mine = SwinIR(drop_path_rate=0.1, **CFG)
    _randomize(mine, 3)
    L = 4
    g = torch.Generator().manual_seed(4)
    x = torch.rand(3, 1, 16, 16, generator=g)
    tgt = torch.rand(3, 1, 16, 16, generator=g)
    drop = ((torch.rand(2 * L, 3, generator=g) < 0.8).float() / 0.8)
    y, grads = _engine_step(mine, x, tgt, drop)
    sd = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in mine.state_dict().items()}
    cfg = so.SwinIRConfig(img_size=16, in_chans=1, embed_dim=180, depths=(2, 2), num_heads=(6, 6), window_size=8,
                          mlp_ratio=2, upscale=1, img_range=1.0, upsampler="")
    y_ref = so.swinir_forward(sd, x, cfg, drop_masks=[(drop[l], drop[L + l]) for l in range(L)])
    _loss(y_ref, tgt).backward()
    assert (y - y_ref.detach()).abs().max().item() < 2e-2
    for k, p in mine.named_parameters():
        gr = sd[k].grad
        assert _rel_l2(grads[k], gr) < 5e-2, (k, _rel_l2(grads[k], gr))


@pytest.mark.parametrize("name", ["swinir_sr_ps4", "swinir_sr_nearest4", "swinir_s_psd2"])
def test_sr_heads_forward_vs_oracle(name):
    """Reconstruction heads (network_swinir.py:740-760, 816-831): pixelshuffle x4, nearest+conv x4 and the lightweight
    pixelshuffledirect x2 (embed_dim 60), inference plan."""
    import golden_cases as gc
    case = gc.CASES[name]
    sd, x, _ = gc.state_and_io(case)
    mine = SwinIR(drop_path_rate=0.0, **case["kw"])
    mine.load_state_dict(sd, strict=True)
    y_ref = gc.oracle_forward(case, sd, x)
    params = dict(mine.named_parameters())
    eng = SwinIREngine(mine._spec, params, "cpu", ops=emu_ops)
    B, _, H, W = x.shape
    y = eng.plan(B, H, W, False).run_forward(x)
    assert y.shape == y_ref.shape
    assert (y - y_ref).abs().max().item() < 2e-2


class _Fp32Buffers:
    """Allocate fp32 wherever the engine / the emulation ask for bf16: what is left is pure host logic."""

    def __enter__(self):
        self.real = torch.bfloat16
        torch.bfloat16 = torch.float32

    def __exit__(self, *exc):
        torch.bfloat16 = self.real


@pytest.mark.parametrize("name", ["swinir_sr_ps4", "swinir_sr_nearest4", "swinir_s_psd2", "swinir_3conv", "swinir_l_small", "swinir_jpeg_ws7"])
def test_sr_heads_backward_vs_oracle(name):
    """Training through the reconstruction heads and the '3conv' residual connection (options/swinir/train_swinir_sr_*.json,
    SwinIR-L): every parameter gradient of the launch lists (emulated) against autograd through the oracle.  Checked
    twice: with fp32 buffers (exact host logic, 1e-4) and with the real bf16 operands (rounding only)."""
    import golden_cases as gc
    case = gc.CASES[name]
    sd, x, tgt = gc.state_and_io(case)
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    y_ref = gc.oracle_forward(case, sdg, x)
    _loss(y_ref, tgt).backward()

    mine = SwinIR(drop_path_rate=0.0, **case["kw"])
    mine.load_state_dict(sd, strict=True)
    with _Fp32Buffers():
        y, grads = _engine_step(mine, x, tgt)
    assert (y - y_ref.detach()).abs().max().item() < 1e-4
    bad = [(k, _rel_l2(grads[k], sdg[k].grad)) for k in grads if not _rel_l2(grads[k], sdg[k].grad) < 1e-4]
    assert not bad, bad[:8]

    mine = SwinIR(drop_path_rate=0.0, **case["kw"])
    mine.load_state_dict(sd, strict=True)
    y, grads = _engine_step(mine, x, tgt)
    assert (y - y_ref.detach()).abs().max().item() < 2e-2
    # bf16 operands: the bias-table gradient is a sum over few windows with heavy cancellation (2x the bound); the
    # 12-channel LeakyReLU bottleneck of the C=48 '3conv' case amplifies rounding for everything upstream of it
    wide = 2.5e-1 if name == "swinir_3conv" else 5e-2
    tol = lambda k: max(wide, 1e-1) if k.endswith("relative_position_bias_table") else wide
    bad = [(k, _rel_l2(grads[k], sdg[k].grad)) for k in grads if not _rel_l2(grads[k], sdg[k].grad) < tol(k)]
    assert not bad, bad[:8]
