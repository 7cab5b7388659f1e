"""SPECT data path on the device: patch sampling for training and full-frame validation with the fork's metrics
(SURVEY.md section 8(f3)).

The reference reads every 2 x 1024 x 256 .dat frame pair from disk in a DataLoader worker, crops / flips / rotates /
normalises it in numpy and hands 64..128-pixel patches to the GPU (data/dataset_spect.py:57-118); validation runs the
network on full frames, de-normalises on the CPU and computes PSNR / SSIM / Poisson log-likelihood in numpy + OpenCV
(models/model_base.py:283-560).  Here the frame stacks live in HBM (2 MB per frame: a 10 000-frame study is 20 GB of
the 180 GB), a training batch is ONE kernel launch (`kair_spect_patches`) and the validation metrics are one launch
per image and scale (`kair_spect_metrics`), so the input pipeline can never starve a 240 patches/s training step.

Same arithmetic, same random-number protocol: `SpectPatchSampler` draws (top, left, mode) with Python's `random` in the
order `DatasetSPECT.__getitem__` does (randint(0, H - P), randint(0, W - P), randint(0, 7)), so a seeded run picks the
same patches.  LPIPS (a pretrained AlexNet / VGG, weights not in the reference tree) is not computed.
"""
import math
import random

import torch

from . import ops as _ops


def _method(normalization):
    t = (normalization or {}).get("type", "log") if isinstance(normalization, dict) else (normalization or "log")
    mp = (normalization or {}).get("max_pixel", 150) if isinstance(normalization, dict) else 150
    if t not in ("log", "anscombe", "linear"):
        raise ValueError("Unknown normalization method: %s" % t)
    return t, float(mp)


def normalize_spect(x, method="log", max_pixel=150, ops=None):
    """utils_spect.normalize_spect on a device tensor (fp32)."""
    o = ops or _ops
    x = x.float().contiguous()
    y = torch.empty_like(x)
    o.build_spect_normalize(x, y, method=method, max_pixel=max_pixel, inverse=False).run(o.stream_handle(x.device))
    return y


def denormalize_spect(x, method="log", max_pixel=150, ops=None):
    """utils_spect.denormalize_spect on a device tensor (fp32)."""
    o = ops or _ops
    x = x.float().contiguous()
    y = torch.empty_like(x)
    o.build_spect_normalize(x, y, method=method, max_pixel=max_pixel, inverse=True).run(o.stream_handle(x.device))
    return y


class SpectPatchSampler:
    """Training batches from device-resident frame stacks.

    frames_L, frames_H: fp32 [n_files, 2, 1024, 256] raw counts exactly as stored in the .dat files (view 0 anterior,
    view 1 posterior, un-mirrored).  `sample(indices)` reproduces DatasetSPECT.__getitem__ (phase 'train') for those
    file indices as one batch {'L': (B, 2, P, P), 'H': (B, 2, P, P)}."""

    def __init__(self, frames_L, frames_H, patch_size=64, normalization=None, rng=None, ops=None):
        self.ops = ops or _ops
        if frames_L.shape != frames_H.shape or frames_L.dim() != 4 or frames_L.shape[1] != 2:
            raise ValueError("frames must be two [n_files, 2, H, W] stacks of one shape")
        self.L, self.H = frames_L.float().contiguous(), frames_H.float().contiguous()
        self.P = int(patch_size)
        self.method, self.max_pixel = _method(normalization)
        self.rng = rng or random
        self._cache = {}

    def __len__(self):
        return self.L.shape[0]

    def draw(self, indices):
        """(file, top, left, mode) rows in the reference's draw order"""
        _, _, FH, FW = self.L.shape
        rows = []
        for f in indices:
            top = self.rng.randint(0, max(0, FH - self.P))
            left = self.rng.randint(0, max(0, FW - self.P))
            mode = self.rng.randint(0, 7)
            rows.append((int(f), top, left, mode))
        return rows

    def sample(self, indices, rows=None):
        rows = rows if rows is not None else self.draw(indices)
        B, dev = len(rows), self.L.device
        st = self._cache.get(B)
        if st is None:
            samp = torch.zeros(B, 4, dtype=torch.int32, device=dev)
            oL = torch.empty(B, 2, self.P, self.P, dtype=torch.float32, device=dev)
            oH = torch.empty_like(oL)
            launch = self.ops.build_spect_patches(self.L, self.H, samp, oL, oH, patch=self.P, method=self.method,
                                                  max_pixel=self.max_pixel)
            st = self._cache[B] = (samp, oL, oH, launch)
        samp, oL, oH, launch = st
        samp.copy_(torch.tensor(rows, dtype=torch.int32), non_blocking=True)
        launch.run(self.ops.stream_handle(dev))
        return {"L": oL.clone(), "H": oH.clone()}


def test_frames(frames, normalization=None, ops=None):
    """DatasetSPECT.__getitem__ (phase 'test') for a whole stack: [n, 2, H, W] raw -> normalised, posterior mirrored."""
    method, max_pixel = _method(normalization)
    x = torch.stack([frames[:, 0], frames[:, 1].flip(-1)], dim=1)
    return normalize_spect(x, method, max_pixel, ops=ops)


def image_metrics(E, H, max_val_global, ops=None):
    """The per-image metric sums of models/model_base.py:379-405 (without LPIPS) for one de-normalised image pair
    (C, h, w): PSNR / SSIM summed over channels at the global and at the local (max of this H) scale, and pll."""
    o = ops or _ops
    E, H = E.float().contiguous(), H.float().contiguous()
    C, h, w = E.shape
    max_local = float(H.max().item())
    out = {}
    for tag, mv in (("global", float(max_val_global)), ("local", max_local)):
        acc = torch.zeros(C, 4, dtype=torch.float64, device=E.device)
        o.build_spect_metrics(E, H, acc, max_val=mv).run(o.stream_handle(E.device))
        acc = acc.cpu()
        psnr = 0.0
        for c in range(C):
            mse = acc[c, 0].item() / (h * w)
            psnr += float("inf") if mse == 0 else 20 * math.log10(255.0 / math.sqrt(mse))
        out["psnr_" + tag] = psnr
        out["ssim_" + tag] = (acc[:, 1].sum() / ((h - 10) * (w - 10))).item()
        out["poisson_ll"] = (acc[:, 2].sum() / (C * h * w)).item()
    return out


@torch.no_grad()
def validate(model, frames_L, frames_H, normalization=None, loss_fn=None, ops=None):
    """`ModelBase.test_spect` without files and figures (models/model_base.py:283-305, 308-348, 491-560): every frame
    pair through the network at full size (2 x 1024 x 256), de-normalised, global max over all H, per-image metrics,
    averages over images x channels.  Returns the reference's metric dict (LPIPS keys absent)."""
    method, max_pixel = _method(normalization)
    L = test_frames(frames_L, normalization, ops=ops)
    H = test_frames(frames_H, normalization, ops=ops)
    was_training = model.training
    model.eval()
    E_all, H_all, loss_sum = [], [], 0.0
    for i in range(L.shape[0]):
        E = model(L[i:i + 1])
        if loss_fn is not None:
            loss_sum += float(loss_fn(E, H[i:i + 1]).item())
        E_all.append(denormalize_spect(E[0], method, max_pixel, ops=ops))
        H_all.append(denormalize_spect(H[i], method, max_pixel, ops=ops))
    model.train(was_training)
    max_global = max(float(h.max().item()) for h in H_all)
    sums = {"psnr_global": 0.0, "ssim_global": 0.0, "psnr_local": 0.0, "ssim_local": 0.0, "poisson_ll": 0.0}
    for E, Hh in zip(E_all, H_all):
        m = image_metrics(E, Hh, max_global, ops=ops)
        for k in sums:
            sums[k] += m[k]
    n, C = len(E_all), L.shape[1]
    out = {k: sums[k] / (n * C) for k in ("psnr_global", "ssim_global", "psnr_local", "ssim_local")}
    out["poisson_ll"] = sums["poisson_ll"] / n
    out["loss"] = loss_sum / n if loss_fn is not None else None
    return out
