"""TEST INFRASTRUCTURE ONLY (imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs; never by the product).

CPU restatement of the serial tail of the reference's training step:
  * CharbonnierLoss            models/loss.py:208-218      loss = mean(sqrt((x - y)^2 + eps)), eps = 1e-9
  * PoissonLLLoss              models/loss.py:294-345      -mean(k log(lambda) - lambda) on de-normalised counts
  * Adam                       torch.optim.Adam as built by models/model_plain.py:139-152 (betas, weight decay = L2 added
                               to the gradient, no amsgrad); the published update rule, in torch's operation order
  * update_E                   models/model_base.py:247-252 e = e*decay + p*(1 - decay), after the optimizer step
Parity is pinned in tests/test_train_step_cpu.py against torch.optim.Adam itself (the implementation the reference calls)
and against the reference's CharbonnierLoss module where /root/reference is mounted."""
import math

import torch


def charbonnier(x, y, eps=1e-9):
    d = x - y
    return torch.mean(torch.sqrt(d * d + eps))


def charbonnier_grad(x, y, eps=1e-9):
    d = x - y
    return d / torch.sqrt(d * d + eps) / x.numel()


def poisson_ll(pred, target, method="linear", max_pixel=150.0, eps=1e-9):
    """PoissonLLLoss.forward (models/loss.py:320-345) with denormalize_spect (:308-318 == utils/utils_spect.py:21-35)."""
    def den(v):
        if method == "log":
            return torch.exp(v * math.log(max_pixel + 1.0)) - 1.0
        if method == "anscombe":
            return (math.sqrt(max_pixel + 3 / 8) * v) ** 2 - 3 / 8
        return v * max_pixel
    lam = torch.clamp(den(pred), min=eps)
    k = torch.clamp(den(target), min=0.0)
    return -torch.mean(k * torch.log(lam) - lam)


def adam_ema_step(params, grads, exp_avg, exp_avg_sq, ema, step, *, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0,
                  ema_decay=0.0):
    """One optimizer step (step = 1-based count AFTER this update) + EMA, in place on lists of fp32 tensors."""
    b1, b2 = betas
    bc1 = 1.0 - b1 ** step
    bc2_sqrt = math.sqrt(1.0 - b2 ** step)
    for i, p in enumerate(params):
        g = grads[i] + weight_decay * p if weight_decay else grads[i]
        exp_avg[i].copy_(exp_avg[i] + (1.0 - b1) * (g - exp_avg[i]))
        exp_avg_sq[i].copy_(b2 * exp_avg_sq[i] + (1.0 - b2) * g * g)
        denom = exp_avg_sq[i].sqrt() / bc2_sqrt + eps
        p.copy_(p - (lr / bc1) * (exp_avg[i] / denom))
        if ema is not None and ema_decay > 0:
            ema[i].copy_(ema[i] * ema_decay + p * (1.0 - ema_decay))
