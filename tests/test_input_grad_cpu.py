"""d(loss)/d(input) through the drop-in modules (the reference gets it from autograd for free: unrolled / plug-and-play
schemes, input-space regularisers) and BatchNorm buffer versioning, on CPU through tests/emu_ops.py vs the oracle."""
import pytest
import torch

import emu_ops
import golden_cases as gc
from oracle import convnets_oracle as co
from oracle import swinir_oracle as so


def _rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


# bf16 mode: the image gradient is a heavily cancelling sum (180 channels x 9 taps of conv_first^T), so the bf16 rounding
# noise of the backward pass shows up amplified: 4-10 % rel-L2 measured (direction intact: cosine > 0.995); the fp32
# mode meets north_star's 1e-3
@pytest.mark.parametrize("name,precision,tol", [("swinir_m_gray32", "bf16", 1.5e-1), ("swinir_m_gray32", "fp32", 1e-3),
                                                ("swinir_sr_ps4", "bf16", 1.5e-1), ("swinir_m_rgb_pad", "bf16", 1.5e-1)])
def test_swinir_input_gradient(name, precision, tol):
    from kair_b200.network_swinir import SwinIR
    case = gc.CASES[name]
    sd, x, tgt = gc.state_and_io(case)
    xr = x.clone().requires_grad_(True)
    gc.loss_fn(case)(gc.oracle_forward(case, sd, xr), tgt).backward()
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model._ops = emu_ops
    model.set_precision(precision)
    for p in model.parameters():          # frozen network, gradient w.r.t. the image only
        p.requires_grad_(False)
    xm = x.clone().requires_grad_(True)
    y = model(xm)
    assert y.requires_grad
    gc.loss_fn(case)(y, tgt).backward()
    assert xm.grad is not None and xm.grad.shape == x.shape
    assert _rel(xm.grad, xr.grad) < tol, _rel(xm.grad, xr.grad)
    assert torch.nn.functional.cosine_similarity(xm.grad.flatten(), xr.grad.flatten(), dim=0).item() > 0.995


@pytest.mark.parametrize("name", ["dncnn_r17", "fdncnn_r8", "unetres_nb2"])
def test_convnet_input_gradient(name):
    from kair_b200.network_dncnn import DnCNN, FDnCNN
    from kair_b200.network_unet import UNetRes
    case = gc.CASES[name]
    sd, x, tgt = gc.state_and_io(case)
    xr = x.clone().requires_grad_(True)
    gc.loss_fn(case)(gc.oracle_forward(case, sd, xr), tgt).backward()
    cls = {"dncnn": DnCNN, "fdncnn": FDnCNN, "unetres": UNetRes}[case["net"]]
    model = cls(**case["kw"])
    model.load_state_dict(sd, strict=True)
    model._ops = emu_ops
    xm = x.clone().requires_grad_(True)
    gc.loss_fn(case)(model(xm), tgt).backward()
    assert _rel(xm.grad, xr.grad) < 1.5e-1, _rel(xm.grad, xr.grad)
    assert torch.nn.functional.cosine_similarity(xm.grad.flatten(), xr.grad.flatten(), dim=0).item() > 0.995
    assert all(p.grad is not None for p in model.parameters())


def test_bn_buffers_versioned_after_train_forward():
    """train-mode forward with NO optimizer step (BN recalibration, gradient accumulation before validation): the
    running statistics changed, so the eval-mode forward must re-fold them -- compared with the oracle."""
    from kair_b200.network_dncnn import DnCNN
    case = gc.CASES["dncnn_br_train"]
    sd, x, tgt = gc.state_and_io(case)
    model = DnCNN(**case["kw"])
    model.load_state_dict(sd, strict=True)
    model._ops = emu_ops
    model.eval()
    with torch.no_grad():
        y0 = model(x)                               # folds the initial running statistics
    v0 = model.model[3].running_mean._version
    model.train()
    with torch.no_grad():
        model(x)                                    # updates running statistics, no optimizer step
    assert model.model[3].running_mean._version > v0
    new_stats = {}
    co.dncnn_forward(sd, x, nb=case["kw"]["nb"], act_mode="BR", training=True, new_stats=new_stats)
    sd2 = dict(sd)
    sd2.update(new_stats)
    model.eval()
    with torch.no_grad():
        y1 = model(x)
        y_ref = co.dncnn_forward(sd2, x, nb=case["kw"]["nb"], act_mode="BR", training=False)
    assert (y1 - y_ref).abs().max().item() < 2e-2
    assert (y1 - y0).abs().max().item() > 1e-3      # the stale fold would have reproduced y0


@pytest.mark.parametrize("net", ["swinir", "dncnn"])
def test_two_forwards_before_one_backward(net):
    """y1 = m(x1); y2 = m(x2); (l(y1) + l(y2)).backward() -- siamese / two-pass losses of the reference work because
    autograd keeps both graphs; the engines keep two activation sets per shape (kair_b200.swinir_engine.pick_plan)."""
    from kair_b200.network_dncnn import DnCNN
    from kair_b200.network_swinir import SwinIR
    case = gc.CASES["swinir_m_gray32" if net == "swinir" else "dncnn_r17"]
    sd, x, tgt = gc.state_and_io(case)
    model = (SwinIR(drop_path_rate=0.0, **case["kw"]) if net == "swinir" else DnCNN(**case["kw"]))
    model.load_state_dict(sd, strict=True)
    model._ops = emu_ops
    x2 = x.flip(3).contiguous()
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    (gc.oracle_forward(case, sdg, x).square().mean() + gc.oracle_forward(case, sdg, x2).abs().mean()).backward()
    y1 = model(x)
    y2 = model(x2)
    (y1.square().mean() + y2.abs().mean()).backward()
    worst = max(_rel(p.grad, sdg[k].grad) for k, p in model.named_parameters())
    assert worst < 6e-2, worst
    # a third pending forward takes over the older activation set: its own backward must then refuse, loudly
    y1 = model(x); y2 = model(x2); y3 = model(x)
    y3.mean().backward(); y2.mean().backward()
    with pytest.raises(RuntimeError, match="two sets are kept"):
        y1.mean().backward()
