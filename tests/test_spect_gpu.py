"""SPECT data path on the CUDA kernels (through the C ABI) vs the oracle / the golden vectors of the reference: patch
sampling (crop, posterior mirror, 8 augment modes, 3 normalisations), normalise / de-normalise, PSNR / SSIM / Poisson
log-likelihood sums, and full-frame (2 x 1024 x 256) validation of a SwinIR through kair_b200.spect.validate."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import spect_oracle as spo
from spect_cases import synth_frames

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kat", "spect_golden.npz"))


@pytest.mark.parametrize("method", ["log", "anscombe", "linear"])
def test_patch_sampler_matches_reference_items(method):
    from kair_b200.spect import SpectPatchSampler, test_frames
    L, H = synth_frames(2, 7, 1024, 256)
    s = SpectPatchSampler(torch.from_numpy(L).cuda(), torch.from_numpy(H).cuda(), 64, {"type": method, "max_pixel": 150})
    random.seed(1234)
    idx = [0, 1, 1, 0, 1, 0, 0, 1]
    rows = s.draw(idx)
    assert rows == [tuple(r) for r in G["draws_" + method].tolist()]
    batch = s.sample(idx, rows=rows)
    assert np.allclose(batch["L"].cpu().numpy(), G["patch_L_" + method], rtol=0, atol=2e-6)
    assert np.allclose(batch["H"].cpu().numpy(), G["patch_H_" + method], rtol=0, atol=2e-6)
    full = test_frames(torch.from_numpy(L).cuda(), {"type": method, "max_pixel": 150})
    assert np.allclose(full[0].cpu().numpy()[:, :160, :96], G["test_L_" + method], rtol=0, atol=2e-6)
    # every augment mode at the patch corners (exact index arithmetic), larger patch, non-square frame offsets
    s2 = SpectPatchSampler(torch.from_numpy(L).cuda(), torch.from_numpy(H).cuda(), 128, {"type": method, "max_pixel": 150})
    rows = [(m % 2, 896 - 7 * m, 128 - 3 * m, m) for m in range(8)]
    got = s2.sample(None, rows=rows)
    for b, (f, top, left, mode) in enumerate(rows):
        pl, ph = spo.train_patch(L[f], H[f], top, left, mode, 128, method, 150)
        assert np.allclose(got["L"][b].cpu().numpy(), pl, rtol=0, atol=2e-6)
        assert np.allclose(got["H"][b].cpu().numpy(), ph, rtol=0, atol=2e-6)


def test_normalize_roundtrip_and_golden():
    from kair_b200 import spect
    x = torch.from_numpy(G["norm_x"]).cuda()
    for method in ("log", "anscombe", "linear"):
        n = spect.normalize_spect(x, method, 150)
        assert np.allclose(n.cpu().numpy(), G["norm_" + method], rtol=2e-6, atol=1e-7)
        d = spect.denormalize_spect(torch.from_numpy(G["norm_" + method].astype(np.float32)).cuda(), method, 150)
        assert np.allclose(d.cpu().numpy(), G["denorm_" + method], rtol=2e-5, atol=2e-5)


def test_metric_kernel_matches_reference_values():
    from kair_b200 import spect
    E, H = G["metric_E"], G["metric_H"]                     # (h, w, 2) counts
    Et = torch.from_numpy(np.ascontiguousarray(E.transpose(2, 0, 1))).cuda()
    Ht = torch.from_numpy(np.ascontiguousarray(H.transpose(2, 0, 1))).cuda()
    m = spect.image_metrics(Et, Ht, float(G["metric_maxvals"][0]))
    ref = G["metric_psnr_ssim"]
    # max_val_local of image_metrics is max(H) == metric_maxvals[1]
    assert abs(m["psnr_global"] - ref[:2, 0].sum()) < 1e-9 and abs(m["psnr_local"] - ref[2:, 0].sum()) < 1e-9
    assert abs(m["ssim_global"] - ref[:2, 1].sum()) < 1e-9 and abs(m["ssim_local"] - ref[2:, 1].sum()) < 1e-9
    assert abs(m["poisson_ll"] - float(G["metric_pll"][0])) < 1e-9


def test_full_frame_validation_of_swinir():
    """2 x 1024 x 256 frames through a (small) SwinIR at full frame size, metrics on the device vs the oracle's numpy
    on the same network outputs"""
    from kair_b200 import spect
    from kair_b200.network_swinir import SwinIR
    L, H = synth_frames(2, 5, 1024, 256)
    torch.manual_seed(0)
    model = SwinIR(upscale=1, in_chans=2, img_size=64, window_size=8, img_range=1.0, depths=[2], embed_dim=60,
                   num_heads=[6], mlp_ratio=2, upsampler="", resi_connection="1conv").cuda()
    norm = {"type": "log", "max_pixel": 150}
    got = spect.validate(model, torch.from_numpy(L).cuda(), torch.from_numpy(H).cuda(), norm, loss_fn=torch.nn.L1Loss())
    model.eval()
    E_imgs, H_imgs = [], []
    with torch.no_grad():
        for i in range(2):
            l = torch.from_numpy(spo.test_frame(L[i], "log", 150))[None].cuda()
            e = model(l)[0].cpu().numpy().transpose(1, 2, 0)
            E_imgs.append(spo.denormalize_spect(e, "log", 150).astype(np.float32))
            H_imgs.append(spo.denormalize_spect(spo.test_frame(H[i], "log", 150).transpose(1, 2, 0), "log", 150).astype(np.float32))
    mx = max(x.max() for x in H_imgs)
    sums = {}
    for e, h in zip(E_imgs, H_imgs):
        for k, v in spo.image_metrics(e, h, mx).items():
            sums[k] = sums.get(k, 0.0) + v
    # float32 exp / log on the device vs numpy: a few uint8 truncations may differ
    for k in ("psnr_global", "psnr_local"):
        assert abs(got[k] - sums[k] / 4) < 2e-2, (k, got[k], sums[k] / 4)
    for k in ("ssim_global", "ssim_local"):
        assert abs(got[k] - sums[k] / 4) < 1e-3, (k, got[k], sums[k] / 4)
    assert abs(got["poisson_ll"] - sums["poisson_ll"] / 2) < 1e-3 * abs(sums["poisson_ll"] / 2) + 1e-3
    assert got["loss"] is not None and np.isfinite(got["loss"])
