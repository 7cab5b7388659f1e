"""SURVEY 8(f1) on CPU: the oracle of the step tail (oracle/train_step_oracle.py) pinned against torch.optim.Adam and the
reference's CharbonnierLoss; the host logic of kair_b200.optim.FusedAdamEMA / kair_b200.loss (state_dict layout, launch
caching, lr schedules, ragged groups) through the emulated launches of tests/emu_ops.py."""
import copy

import pytest
import torch

import emu_ops
import ref_loader
from oracle import train_step_oracle as to
from kair_b200.loss import CharbonnierLoss, L1Loss
from kair_b200.optim import FusedAdamEMA, update_ema

SHAPES = [(7,), (3, 5), (1,), (33, 17, 3), (8192,), (8193,), (2, 3, 3, 3)]


def _params(seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(s, generator=g)) for s in SHAPES]


def _set_grads(ps, seed):
    g = torch.Generator().manual_seed(seed)
    for p in ps:
        p.grad = torch.randn(p.shape, generator=g) * 0.1


@pytest.mark.parametrize("wd", [0.0, 1e-2])
def test_oracle_matches_torch_adam_and_update_e(wd):
    ps = _params(1)
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    opt = torch.optim.Adam(ref, lr=2e-4, betas=(0.9, 0.999), weight_decay=wd, foreach=False)
    ema_ref = [p.detach().clone() for p in ref]
    mine = [p.detach().clone() for p in ps]
    m = [torch.zeros_like(p) for p in mine]
    v = [torch.zeros_like(p) for p in mine]
    ema = [p.clone() for p in mine]
    for step in range(1, 6):
        _set_grads(ref, 10 + step)
        opt.step()
        for e, p in zip(ema_ref, ref):     # models/model_base.py:247-252
            e.mul_(0.999).add_(p.data, alpha=1 - 0.999)
        to.adam_ema_step(mine, [p.grad for p in ref], m, v, ema, step, lr=2e-4, weight_decay=wd, ema_decay=0.999)
    for a, b in zip(mine, ref):
        assert torch.allclose(a, b.data, rtol=1e-6, atol=1e-7)
    for a, b in zip(ema, ema_ref):
        assert torch.allclose(a, b, rtol=1e-6, atol=1e-7)


@pytest.mark.skipif(not ref_loader.have_reference(), reason="reference tree not mounted")
def test_oracle_charbonnier_matches_reference_module():
    ref = ref_loader.load("loss").CharbonnierLoss()
    g = torch.Generator().manual_seed(3)
    x, y = torch.rand(2, 1, 16, 16, generator=g, requires_grad=True), torch.rand(2, 1, 16, 16, generator=g)
    l = ref(x, y)
    l.backward()
    assert abs(to.charbonnier(x.detach(), y).item() - l.item()) < 1e-7
    assert torch.allclose(to.charbonnier_grad(x.detach(), y), x.grad, rtol=1e-5, atol=1e-9)


def test_fused_adam_ema_host_logic_and_state_dict():
    ps = _params(2)
    ref = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    ema_mine = [p.detach().clone() for p in ps]
    ema_ref = [p.detach().clone() for p in ps]
    opt_ref = torch.optim.Adam(ref, lr=1e-3, betas=(0.9, 0.99), foreach=False)
    opt = FusedAdamEMA(ps, lr=1e-3, betas=(0.9, 0.99), ema_params=ema_mine, ema_decay=0.999, ops=emu_ops)
    sch_ref = torch.optim.lr_scheduler.MultiStepLR(opt_ref, milestones=[2], gamma=0.5)
    sch = torch.optim.lr_scheduler.MultiStepLR(opt, milestones=[2], gamma=0.5)
    for step in range(4):
        _set_grads(ps, 20 + step)
        for a, b in zip(ref, ps):
            a.grad = b.grad.clone()
        if step == 2:
            ps[3].grad = None   # a parameter without a gradient is skipped (its step count then lags: separate launch)
            ref[3].grad = None
        opt.step(); opt_ref.step()
        update_ema(ema_ref, ref, 0.999)
        sch.step(); sch_ref.step()
    for a, b in zip(ps, ref):
        assert torch.allclose(a.data, b.data, rtol=1e-6, atol=1e-7)
    assert all(p._version >= 3 for p in ps)   # the engines watch parameter versions to re-pack their operands
    for i, (a, b) in enumerate(zip(ema_mine, ema_ref)):
        if i != 3:   # update_E also averages parameters that got no gradient; the fused step leaves them alone
            assert torch.allclose(a, b, rtol=1e-6, atol=1e-7)
    sd, sd_ref = opt.state_dict(), opt_ref.state_dict()
    assert sd["state"].keys() == sd_ref["state"].keys()
    for k in sd["state"]:
        assert sd["state"][k].keys() == sd_ref["state"][k].keys()
        assert float(sd["state"][k]["step"]) == float(sd_ref["state"][k]["step"])
        assert torch.allclose(sd["state"][k]["exp_avg_sq"], sd_ref["state"][k]["exp_avg_sq"], rtol=1e-6, atol=1e-12)
    assert set(sd_ref["param_groups"][0]) - {"decoupled_weight_decay"} <= set(sd["param_groups"][0]) | {"initial_lr"}
    # a torch.optim.Adam checkpoint resumes in the fused optimizer (model_base.py:232-238 load_optimizer)
    opt2 = FusedAdamEMA([torch.nn.Parameter(p.detach().clone()) for p in ref], lr=1e-3, betas=(0.9, 0.99), ops=emu_ops)
    opt2.load_state_dict(copy.deepcopy(sd_ref))
    assert float(opt2.state[opt2.param_groups[0]["params"][0]]["step"]) == 4.0
    with pytest.raises(NotImplementedError):
        FusedAdamEMA(_params(3), amsgrad=True, ops=emu_ops)


@pytest.mark.parametrize("cls,ref", [(CharbonnierLoss, lambda a, b: to.charbonnier(a, b)),
                                     (L1Loss, lambda a, b: torch.mean(torch.abs(a - b)))])
def test_loss_modules_host_logic(cls, ref):
    g = torch.Generator().manual_seed(5)
    x = torch.rand(2, 3, 8, 8, generator=g, requires_grad=True)
    y = torch.rand(2, 3, 8, 8, generator=g)
    mod = cls()
    mod._ops = emu_ops
    l = mod(x, y) * 2.0
    l.backward()
    xr = x.detach().clone().requires_grad_(True)
    lr = ref(xr, y) * 2.0
    lr.backward()
    assert l.shape == () and abs(l.item() - lr.item()) < 1e-6
    assert torch.allclose(x.grad, xr.grad, rtol=1e-5, atol=1e-9)
    with torch.no_grad():
        assert abs(mod(x, y).item() * 2.0 - lr.item()) < 1e-6
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cls()(x, y)


def test_engines_repack_after_an_optimizer_step_that_hides_from_version_counters():
    """torch.optim.Adam(fused=True) (and apex) update parameters without bumping Tensor._version: the engines must still
    see the new weights on their next forward (kair_b200/_dirty.py: global optimizer post-step hook)."""
    from kair_b200.network_dncnn import DnCNN
    from kair_b200 import _dirty
    torch.manual_seed(0)
    model = DnCNN(in_nc=1, out_nc=1, nc=64, nb=4, act_mode="R")
    model._ops = emu_ops
    x = torch.rand(1, 1, 16, 16)
    with torch.no_grad():
        y0 = model(x)

    class Sneaky(torch.optim.Optimizer):      # writes through .data views: no version bump on the parameters
        def __init__(self, params):
            super().__init__(params, {})

        def step(self, closure=None):
            for g in self.param_groups:
                for p in g["params"]:
                    torch.Tensor.add_(p.data, 0.05)

    opt = Sneaky(model.parameters())
    v = [p._version for p in model.parameters()]
    e0 = _dirty.optimizer_epoch()
    opt.step()
    assert [p._version for p in model.parameters()] == v and _dirty.optimizer_epoch() == e0 + 1
    with torch.no_grad():
        y1 = model(x)
    assert (y1 - y0).abs().max().item() > 1e-3


@pytest.mark.parametrize("method", ["linear", "log", "anscombe"])
def test_poisson_ll_oracle_and_module(method):
    """PoissonLLLoss (the SPECT fork's loss): oracle vs the reference module where mounted, drop-in module (emulated
    kernel) vs the oracle incl. the gradient and the clamp at epsilon (negative predictions)."""
    from kair_b200.loss import PoissonLLLoss
    g = torch.Generator().manual_seed(11)
    x = (torch.rand(2, 1, 16, 16, generator=g) * 0.6 - 0.05)     # some negative predictions -> clamped, zero gradient
    y = torch.poisson(torch.rand(2, 1, 16, 16, generator=g) * 20.0) / 150.0
    if method == "log":
        y = torch.log(y * 150.0 + 1.0) / torch.log(torch.tensor(151.0))
    xr = x.clone().requires_grad_(True)
    lo = to.poisson_ll(xr, y, method)
    lo.backward()
    if ref_loader.have_reference():
        ref = ref_loader.load("loss").PoissonLLLoss(method, 150.0, 1e-9)
        xq = x.clone().requires_grad_(True)
        lq = ref(xq, y)
        lq.backward()
        assert abs(lq.item() - lo.item()) < 1e-6 * max(1.0, abs(lq.item()))
        assert torch.allclose(xq.grad, xr.grad, rtol=1e-5, atol=1e-8)
    mod = PoissonLLLoss(method, 150.0, 1e-9)
    mod._ops = emu_ops
    xm = x.clone().requires_grad_(True)
    lm = mod(xm, y)
    lm.backward()
    assert abs(lm.item() - lo.item()) < 1e-5 * max(1.0, abs(lo.item()))
    assert torch.allclose(xm.grad, xr.grad, rtol=1e-4, atol=1e-7)
    with pytest.raises(ValueError):
        PoissonLLLoss("sqrt")


def test_psnr_metric_host_logic():
    """kair_b200.metrics.psnr_uint8 == golden_cases.psnr_uint8 (the oracle's restatement of tensor2uint +
    calculate_psnr), incl. border cropping and identical images."""
    import golden_cases as gc
    from kair_b200 import metrics
    g = torch.Generator().manual_seed(2)
    a = torch.rand(1, 1, 20, 24, generator=g) * 1.2 - 0.1
    b = a + torch.randn(a.shape, generator=g) * 0.05
    assert abs(metrics.psnr_uint8(a, b, ops=emu_ops) - gc.psnr_uint8(a, b)) < 1e-4
    assert abs(metrics.psnr_uint8(a, b, border=3, ops=emu_ops) - gc.psnr_uint8(a[..., 3:-3, 3:-3], b[..., 3:-3, 3:-3])) < 1e-4
    assert metrics.psnr_uint8(a, a, ops=emu_ops) == float("inf")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        metrics.psnr_uint8(a, b)
