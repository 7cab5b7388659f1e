"""Host logic of the fp32-accurate mode (kair_b200/swinir_engine_fp32.py): the launch lists replayed on the CPU through
tests/emu_ops.py -- which states the split-operand semantics of include/kair_b200.h ("SPLIT PRECISION") with real bf16
rounding of every component -- against the oracle at north_star's fp32-mode bounds: 1e-4 max-abs forward and 1e-3
relative per-tensor gradients.  The same comparison on the CUDA kernels is tests/test_swinir_fp32_gpu.py."""
import pytest
import torch

import emu_ops
import golden_cases as gc
from kair_b200.network_swinir import SwinIR
from kair_b200.swinir_engine import SwinIREngine, SwinIRFunction
from oracle import swinir_oracle as so


def _rel_l2(a, b):
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def _run(case, terms, train=True, drop=None):
    sd, x, tgt = gc.state_and_io(case)
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    params = dict(model.named_parameters())
    eng = SwinIREngine(model._spec, params, "cpu", ops=emu_ops)
    B, _, H, W = x.shape
    plan = eng.plan(B, H, W, train, "fp32", terms)
    if not train:
        with torch.no_grad():
            return plan.run_forward(x), None, sd, x, tgt
    if drop is not None:
        plan.dp.copy_(drop)
    out = SwinIRFunction.apply(x, plan, *params.values())
    gc.loss_fn(case)(out, tgt).backward()
    return out.detach(), {k: p.grad.clone() for k, p in params.items()}, sd, x, tgt


@pytest.mark.parametrize("terms", [3, 6])
def test_fp32_plan_forward_backward_vs_oracle(terms):
    case = gc.CASES["swinir_m_gray32"]
    y, grads, sd, x, tgt = _run(case, terms)
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    y_ref = gc.oracle_forward(case, sdg, x)
    gc.loss_fn(case)(y_ref, tgt).backward()
    assert (y - y_ref.detach()).abs().max().item() < 1e-4
    bad = [(k, _rel_l2(g, sdg[k].grad)) for k, g in grads.items() if not _rel_l2(g, sdg[k].grad) < 1e-3]
    assert not bad, bad[:8]


def test_fp32_plan_drop_path_replayed():
    case = gc.CASES["swinir_m_gray32"]
    g = torch.Generator().manual_seed(7)
    L, B = 4, case["x"][0]
    drop = (torch.rand(2 * L, B, generator=g) < 0.7).float() / 0.7
    y, grads, sd, x, tgt = _run(case, 3, drop=drop)
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    y_ref = so.swinir_forward(sdg, x, gc.swinir_cfg(case["kw"]), drop_masks=[(drop[l], drop[L + l]) for l in range(L)])
    gc.loss_fn(case)(y_ref, tgt).backward()
    assert (y - y_ref.detach()).abs().max().item() < 1e-4
    bad = [(k, _rel_l2(g_, sdg[k].grad)) for k, g_ in grads.items() if not _rel_l2(g_, sdg[k].grad) < 1e-3]
    assert not bad, bad[:8]


@pytest.mark.parametrize("name", ["swinir_sr_ps4", "swinir_sr_nearest4", "swinir_s_psd2", "swinir_m_rgb_pad",
                                  "swinir_jpeg_ws7"])
def test_fp32_plan_inference_heads(name):
    """forward-only plans of the SR heads / other window sizes in fp32 mode vs the oracle at 1e-4"""
    case = gc.CASES[name]
    sd, x, tgt = gc.state_and_io(case)
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model._ops = emu_ops
    model.set_precision("fp32")
    with torch.no_grad():
        y = model(x)
        y_ref = gc.oracle_forward(case, sd, x)
    scale = max(1.0, case["kw"]["img_range"] / 255.0 * 0 + 1.0)
    assert y.shape == y_ref.shape
    assert (y - y_ref).abs().max().item() < 1e-4 * scale


def test_fp32_mode_rejects_what_it_does_not_cover():
    case = gc.CASES["swinir_3conv"]
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model._ops = emu_ops
    model.set_precision("fp32")
    with pytest.raises(NotImplementedError):
        model(torch.rand(1, 3, 16, 16))
    with pytest.raises(ValueError):
        model.set_precision("fp16")
