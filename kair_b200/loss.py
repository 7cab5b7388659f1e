"""Fused pixel losses: drop-ins for `CharbonnierLoss` (models/loss.py:208-218 of the reference) and `nn.L1Loss`
(select_network / model_plain.py:113-137 `define_loss`), computed together with their gradient by one sm_100a kernel
(`kair_pixel_loss`, kair_b200/csrc/train_step.cu).  SURVEY.md section 8(f1).  No CPU fallback."""
import torch
import torch.nn as nn

from . import ops as _real_ops


class _PixelLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y, owner):
        o = owner._ops or _real_ops
        if not x.is_cuda and owner._ops is None:
            raise RuntimeError("kair_b200 losses run on sm_100a GPUs only; got a %s tensor (there is no CPU fallback)"
                               % x.device.type)
        xc, yc = x.detach().float().contiguous(), y.detach().float().contiguous()
        need = x.requires_grad or y.requires_grad
        dx = torch.empty_like(xc) if need else None
        loss = torch.empty(1, dtype=torch.float32, device=x.device)
        o.build_pixel_loss(xc, yc, dx, loss, owner._scratch_for(x.device), eps=owner.eps, mode=owner._mode,
                           max_pixel=getattr(owner, "max_pixel", 0.0)).run(o.stream_handle())
        ctx.dx = dx
        ctx.needs = (x.requires_grad, y.requires_grad)
        return loss.view(())

    @staticmethod
    def backward(ctx, gout):
        g = ctx.dx * gout
        return (g if ctx.needs[0] else None), (-g if ctx.needs[1] else None), None


class _PixelLoss(nn.Module):
    _mode = 0

    def __init__(self):
        super().__init__()
        self._scratch = {}
        self._ops = None   # test hook: tests/emu_ops.py

    def _scratch_for(self, device):
        s = self._scratch.get(device)
        if s is None:
            s = self._scratch[device] = torch.zeros(2, dtype=torch.float64, device=device)
        return s

    def forward(self, x, y):
        return _PixelLossFn.apply(x, y, self)


class CharbonnierLoss(_PixelLoss):
    """mean(sqrt((x - y)^2 + eps)), eps = 1e-9 by default (models/loss.py:208-218)."""
    _mode = 0

    def __init__(self, eps=1e-9):
        super().__init__()
        self.eps = eps


class L1Loss(_PixelLoss):
    """nn.L1Loss() with the default 'mean' reduction (the DRUNet / DnCNN training loss, model_plain.py:117)."""
    _mode = 1
    eps = 0.0


class PoissonLLLoss(_PixelLoss):
    """Poisson negative log-likelihood on de-normalised SPECT counts (models/loss.py:294-345 of the reference; the fork's
    `G_lossfn_type: poisson`, model_plain.py:193-202): lambda = clamp(denorm(pred), min=epsilon), k = clamp(denorm(target),
    min=0), loss = -mean(k*log(lambda) - lambda); denorm per `normalization_method` as utils/utils_spect.py:21-35."""
    _MODES = {"linear": 2, "log": 3, "anscombe": 4}

    def __init__(self, normalization_method="linear", max_pixel=150.0, epsilon=1e-9):
        super().__init__()
        if normalization_method not in self._MODES:
            raise ValueError(f"Unknown normalization method: {normalization_method}")
        self.normalization_method = normalization_method
        self.max_pixel = float(max_pixel)
        self.epsilon = self.eps = epsilon
        self._mode = self._MODES[normalization_method]
