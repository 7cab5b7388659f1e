"""GPU-resident validation metric: PSNR of the uint8-quantised images, i.e. `calculate_psnr(tensor2uint(E), tensor2uint(H),
border)` of the reference (utils/utils_image.py:296-300, 629-645) without copying the images to the host
(the fork validates on full 2x1024x256 SPECT frames, models/model_base.py:308-348).  One launch of `kair_pixel_loss`
(mode 5: mean squared difference of round(clamp(v,0,1)*255)); the logarithm is taken on the host from one scalar."""
import math

import torch

from . import ops as _real_ops

_SCRATCH = {}


def mse_uint8(x, y, border=0, ops=None):
    """Mean squared difference of the uint8 images of x and y ((..., H, W) tensors in [0, 1]); a device scalar."""
    o = ops or _real_ops
    if not x.is_cuda and ops is None:
        raise RuntimeError("kair_b200 metrics run on sm_100a GPUs only; got a %s tensor (there is no CPU fallback)"
                           % x.device.type)
    if x.shape != y.shape:
        raise ValueError("Input images must have the same dimensions.")
    if border:
        x, y = x[..., border:x.shape[-2] - border, border:x.shape[-1] - border], \
            y[..., border:y.shape[-2] - border, border:y.shape[-1] - border]
    xc, yc = x.detach().float().contiguous(), y.detach().float().contiguous()
    scratch = _SCRATCH.get(x.device)
    if scratch is None:
        scratch = _SCRATCH[x.device] = torch.zeros(2, dtype=torch.float64, device=x.device)
    out = torch.empty(1, dtype=torch.float32, device=x.device)
    o.build_pixel_loss(xc, yc, None, out, scratch, eps=0.0, mode=5).run(o.stream_handle())
    return out.view(())


def psnr_uint8(x, y, border=0, ops=None):
    """20*log10(255/sqrt(mse)) as a Python float (inf for identical images), like utils_image.calculate_psnr."""
    mse = float(mse_uint8(x, y, border, ops))
    return float("inf") if mse == 0 else 20.0 * math.log10(255.0 / math.sqrt(mse))
