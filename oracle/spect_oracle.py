"""ORACLE (test infrastructure, not product code): CPU restatement (numpy, float64 where the reference uses it) of the
SPECT data path and validation metrics of the reference fork.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.  Pinned by
tests/test_spect_cpu.py against the reference's own functions (data/dataset_spect.py, utils/utils_spect.py,
utils/utils_image.py -- imported from /root/reference where it is mounted) and against the committed golden vectors
tests/golden/kat/spect_golden.npz generated from them by tests/golden/make_spect_golden.py.
"""
import math

import numpy as np


def normalize_spect(img, method="log", max_pixel=150):
    """utils/utils_spect.py:3-18."""
    if method == "log":
        return np.log(img + 1.0) / np.log(max_pixel + 1.0)
    if method == "anscombe":
        return np.sqrt(img + 3 / 8) / np.sqrt(max_pixel + 3 / 8)
    if method == "linear":
        return img / max_pixel
    raise ValueError("Unknown normalization method: %s" % method)


def denormalize_spect(img, method="log", max_pixel=150):
    """utils/utils_spect.py:20-35."""
    if method == "log":
        return np.exp(img * np.log(max_pixel + 1.0)) - 1.0
    if method == "anscombe":
        return (np.sqrt(max_pixel + 3 / 8) * img) ** 2 - 3 / 8
    if method == "linear":
        return img * max_pixel
    raise ValueError("Unknown normalization method: %s" % method)


def augment_img(img, mode=0):
    """utils/utils_image.py:387-405 (HWC arrays)."""
    if mode == 0:
        return img
    if mode == 1:
        return np.flipud(np.rot90(img))
    if mode == 2:
        return np.flipud(img)
    if mode == 3:
        return np.rot90(img, k=3)
    if mode == 4:
        return np.flipud(np.rot90(img, k=2))
    if mode == 5:
        return np.rot90(img)
    if mode == 6:
        return np.rot90(img, k=2)
    return np.flipud(np.rot90(img, k=3))


def stack_views(frame):
    """data/dataset_spect.py:71-79: (2, 1024, 256) .dat frame -> (1024, 256, 2) with the posterior view mirrored."""
    return np.stack([frame[0], np.fliplr(frame[1])], axis=2)


def train_patch(frame_L, frame_H, top, left, mode, patch, method, max_pixel):
    """data/dataset_spect.py:82-118 for given random draws: crop, augment, normalise, HWC -> CHW float32."""
    out = []
    for f in (frame_L, frame_H):
        img = stack_views(f)[top:top + patch, left:left + patch, :]
        img = augment_img(img, mode)
        img = normalize_spect(img, method, max_pixel)
        out.append(np.ascontiguousarray(img.transpose((2, 0, 1))).astype(np.float32))
    return out


def test_frame(frame, method, max_pixel):
    """data/dataset_spect.py:119-131: full frame, normalised, CHW float32."""
    img = normalize_spect(stack_views(frame), method, max_pixel)
    return np.ascontiguousarray(img.transpose((2, 0, 1))).astype(np.float32)


def quantize(ch, max_val):
    """models/model_base.py:358-366: clip to [0, max], / max * 255, astype(uint8) (truncation)."""
    if max_val > 0:
        return (np.clip(ch, 0, max_val) / max_val * 255).astype(np.uint8)
    return np.zeros_like(ch, dtype=np.uint8)


def psnr_uint8(a, b):
    """utils/utils_image.py:629-644."""
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return float("inf") if mse == 0 else 20 * math.log10(255.0 / math.sqrt(mse))


def gaussian_kernel_11():
    """cv2.getGaussianKernel(11, 1.5): exp(-(i - 5)^2 / (2 sigma^2)), normalised to sum 1."""
    g = np.exp(-((np.arange(11) - 5.0) ** 2) / (2 * 1.5 ** 2))
    return g / g.sum()


def _filter_valid(img, g):
    """cv2.filter2D(img, -1, outer(g, g))[5:-5, 5:-5] for a symmetric separable window, float64."""
    h, w = img.shape
    rows = np.zeros((h, w - 10))
    for k in range(11):
        rows += g[k] * img[:, k:k + w - 10]
    out = np.zeros((h - 10, w - 10))
    for k in range(11):
        out += g[k] * rows[k:k + h - 10, :]
    return out


def ssim_uint8(a, b):
    """utils/utils_image.py:677-700 on one channel."""
    C1, C2 = (0.01 * 255) ** 2, (0.03 * 255) ** 2
    a, b = a.astype(np.float64), b.astype(np.float64)
    g = gaussian_kernel_11()
    mu1, mu2 = _filter_valid(a, g), _filter_valid(b, g)
    s1 = _filter_valid(a * a, g) - mu1 ** 2
    s2 = _filter_valid(b * b, g) - mu2 ** 2
    s12 = _filter_valid(a * b, g) - mu1 * mu2
    m = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 ** 2 + mu2 ** 2 + C1) * (s1 + s2 + C2))
    return m.mean()


def pll(mean_image, observed_image):
    """utils/utils_image.py:833-865 with border 0: mean Poisson log-pmf of round(observed) under max(mean, 1e-9)."""
    k = np.round(observed_image).astype(np.int32).astype(np.float64)
    mu = np.maximum(mean_image, 1e-9).astype(np.float64)   # the clamp happens in the image's own dtype (float32)
    lg = np.vectorize(math.lgamma)(k + 1.0)
    return float(np.mean(k * np.log(mu) - mu - lg))


def image_metrics(E_img, H_img, max_val_global):
    """models/model_base.py:379-405 without LPIPS: per-image sums over channels of PSNR / SSIM at the global and the
    local (this image's max of H) scale, and pll.  E_img, H_img: (h, w, C) de-normalised."""
    max_local = np.max(H_img)
    out = {"psnr_global": 0.0, "ssim_global": 0.0, "psnr_local": 0.0, "ssim_local": 0.0}
    for ch in range(E_img.shape[2]):
        for tag, mv in (("global", max_val_global), ("local", max_local)):
            e8, h8 = quantize(E_img[:, :, ch], mv), quantize(H_img[:, :, ch], mv)
            out["psnr_" + tag] += psnr_uint8(e8, h8)
            out["ssim_" + tag] += ssim_uint8(e8, h8)
    out["poisson_ll"] = pll(E_img, H_img)
    return out
