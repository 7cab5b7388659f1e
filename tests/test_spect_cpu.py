"""SPECT data path: the oracle restatement against the committed golden vectors generated from the
unmodified reference (tests/golden/make_spect_golden.py) and -- where /root/reference is mounted -- against the live
reference functions; then the host logic of kair_b200.spect through tests/emu_ops.py."""
import os
import random

import numpy as np
import pytest
import torch

import emu_ops
import ref_loader
from oracle import spect_oracle as spo
from spect_cases import synth_frames

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kat", "spect_golden.npz"))


@pytest.mark.parametrize("method", ["log", "anscombe", "linear"])
def test_oracle_matches_golden_dataset_items(method):
    L, H = synth_frames(2, 7, 1024, 256)
    for b, (f, top, left, mode) in enumerate(G["draws_" + method].tolist()):
        pl, ph = spo.train_patch(L[f], H[f], top, left, mode, 64, method, 150)
        assert np.array_equal(pl, G["patch_L_" + method][b]) and np.array_equal(ph, G["patch_H_" + method][b])
    assert np.array_equal(spo.test_frame(L[0], method, 150)[:, :160, :96], G["test_L_" + method])
    x = G["norm_x"]
    assert np.allclose(spo.normalize_spect(x, method, 150), G["norm_" + method], rtol=0, atol=0)
    n32 = np.asarray(spo.normalize_spect(x, method, 150), dtype=np.float32)
    assert np.allclose(spo.denormalize_spect(n32, method, 150), G["denorm_" + method], rtol=0, atol=0)


def test_oracle_matches_golden_metrics():
    E, H = G["metric_E"], G["metric_H"]
    rows = []
    for mv in G["metric_maxvals"]:
        for ch in range(2):
            e8, h8 = spo.quantize(E[:, :, ch], mv), spo.quantize(H[:, :, ch], mv)
            rows.append((spo.psnr_uint8(e8, h8), spo.ssim_uint8(e8, h8)))
    assert np.allclose(np.array(rows), G["metric_psnr_ssim"], rtol=1e-12, atol=0)
    assert abs(spo.pll(E, H) - float(G["metric_pll"][0])) < 1e-9
    m = spo.image_metrics(E, H, G["metric_maxvals"][0])
    assert abs(m["psnr_global"] - G["metric_psnr_ssim"][:2, 0].sum()) < 1e-9


@pytest.mark.skipif(not ref_loader.have_reference(), reason="reference tree not mounted")
def test_oracle_matches_live_reference_functions():
    util, uspect, _ = ref_loader.load_utils()
    rng = np.random.RandomState(0)
    img = (rng.rand(48, 40, 2) * 180).astype(np.float32)
    for mode in range(8):
        assert np.array_equal(spo.augment_img(img, mode), util.augment_img(img, mode))
    for method in ("log", "anscombe", "linear"):
        assert np.array_equal(spo.normalize_spect(img, method, 150), uspect.normalize_spect(img, method, 150))
        assert np.array_equal(spo.denormalize_spect(img / 200, method, 150), uspect.denormalize_spect(img / 200, method, 150))
    a = (rng.rand(64, 72) * 255).astype(np.uint8)
    b = np.clip(a.astype(np.int32) + rng.randint(-9, 9, a.shape), 0, 255).astype(np.uint8)
    assert abs(spo.psnr_uint8(a, b) - util.calculate_psnr(a, b)) < 1e-12
    assert abs(spo.ssim_uint8(a, b) - util.calculate_ssim(a, b)) < 1e-12
    mean, obs = rng.rand(30, 20, 2) * 50, rng.poisson(20, (30, 20, 2)).astype(np.float32)
    assert abs(spo.pll(mean, obs) - util.pll(mean, obs)) < 1e-9


def test_sampler_reproduces_the_reference_draws_and_patches():
    """kair_b200.spect.SpectPatchSampler through emu_ops == the golden DatasetSPECT items, same `random` protocol"""
    from kair_b200.spect import SpectPatchSampler
    L, H = synth_frames(2, 7, 1024, 256)
    for method in ("log", "linear"):
        s = SpectPatchSampler(torch.from_numpy(L), torch.from_numpy(H), 64, {"type": method, "max_pixel": 150}, ops=emu_ops)
        random.seed(1234)
        idx = [0, 1, 1, 0, 1, 0, 0, 1]
        rows = s.draw(idx)
        assert rows == [tuple(r) for r in G["draws_" + method].tolist()]
        batch = s.sample(idx, rows=rows)
        assert np.array_equal(batch["L"].numpy(), G["patch_L_" + method])
        assert np.array_equal(batch["H"].numpy(), G["patch_H_" + method])


def test_validate_host_logic_matches_the_oracle_loop():
    """kair_b200.spect.validate (emu ops, a stand-in network) vs the reference's loop restated on the oracle"""
    from kair_b200 import spect
    L, H = synth_frames(3, 11, 64, 48)
    net = torch.nn.Conv2d(2, 2, 3, padding=1)
    torch.manual_seed(0)
    with torch.no_grad():
        net.weight.mul_(0.1)
        net.weight[0, 0, 1, 1] += 1.0
        net.weight[1, 1, 1, 1] += 1.0
    norm = {"type": "log", "max_pixel": 150}
    got = spect.validate(net, torch.from_numpy(L), torch.from_numpy(H), norm, loss_fn=torch.nn.L1Loss(), ops=emu_ops)
    E_imgs, H_imgs, loss = [], [], 0.0
    for i in range(3):
        l, h = spo.test_frame(L[i], "log", 150), spo.test_frame(H[i], "log", 150)
        with torch.no_grad():
            e = net(torch.from_numpy(l)[None])
        loss += torch.nn.functional.l1_loss(e, torch.from_numpy(h)[None]).item()
        # the device path keeps the de-normalised counts in float32 (the reference's numpy promotes them to float64
        # under NumPy >= 2 and keeps float32 under NumPy 1.x: a handful of uint8 truncations differ, < 1e-3 dB)
        E_imgs.append(spo.denormalize_spect(e[0].numpy().transpose(1, 2, 0), "log", 150).astype(np.float32))
        H_imgs.append(spo.denormalize_spect(h.transpose(1, 2, 0), "log", 150).astype(np.float32))
    mx = max(x.max() for x in H_imgs)
    sums = {}
    for e, h in zip(E_imgs, H_imgs):
        for k, v in spo.image_metrics(e, h, mx).items():
            sums[k] = sums.get(k, 0.0) + v
    for k in ("psnr_global", "ssim_global", "psnr_local", "ssim_local"):
        assert abs(got[k] - sums[k] / 6) < 1e-6, k
    assert abs(got["poisson_ll"] - sums["poisson_ll"] / 3) < 1e-6
    assert abs(got["loss"] - loss / 3) < 1e-6
