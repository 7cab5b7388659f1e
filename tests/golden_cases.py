"""TEST INFRASTRUCTURE: the seeded cases shared by tests/golden/make_golden.py (reference side) and the parity tests
(oracle side / CUDA side).  Weights come from oracle.swinir_oracle.deterministic_state (one CPU generator per key),
inputs and targets from a seeded CPU generator, so both sides rebuild identical tensors without storing them."""
import math

import torch

from oracle import convnets_oracle as co
from oracle import swinir_oracle as so

_SW = dict(net="swinir", patch_norm=True, qkv_bias=True)
CASES = {
    # SwinIR-M shaped (C=180, 6 heads, ws 8, mlp 2) gray denoiser, the cfg2 architecture at reduced depth / size
    "swinir_m_gray32": dict(_SW, kw=dict(img_size=32, in_chans=1, embed_dim=180, depths=[2, 2], num_heads=[6, 6],
                                         window_size=8, mlp_ratio=2, upscale=1, img_range=1.0, upsampler="",
                                         resi_connection="1conv"), x=(2, 1, 32, 32), seed=3, loss="charbonnier"),
    # colour, input not a multiple of the window and != img_size: reflect pad + dynamic mask (network_swinir.py:262)
    "swinir_m_rgb_pad": dict(_SW, kw=dict(img_size=32, in_chans=3, embed_dim=180, depths=[2], num_heads=[6],
                                          window_size=8, mlp_ratio=2, upscale=1, img_range=1.0, upsampler="",
                                          resi_connection="1conv"), x=(1, 3, 27, 35), seed=4, loss="l1"),
    # classical SR x4 head (cfg3 architecture, reduced): pixelshuffle
    "swinir_sr_ps4": dict(_SW, kw=dict(img_size=16, in_chans=3, embed_dim=180, depths=[2], num_heads=[6],
                                       window_size=8, mlp_ratio=2, upscale=4, img_range=1.0,
                                       upsampler="pixelshuffle", resi_connection="1conv"), x=(2, 3, 16, 16), seed=5,
                          loss="l1"),
    # real-world SR x4 head (cfg5 architecture, reduced): nearest+conv
    "swinir_sr_nearest4": dict(_SW, kw=dict(img_size=16, in_chans=3, embed_dim=180, depths=[2], num_heads=[6],
                                            window_size=8, mlp_ratio=2, upscale=4, img_range=1.0,
                                            upsampler="nearest+conv", resi_connection="1conv"), x=(1, 3, 16, 24),
                               seed=6, loss="l1"),
    # lightweight SR: pixelshuffledirect, C=60 (head_dim 10)
    "swinir_s_psd2": dict(_SW, kw=dict(img_size=16, in_chans=3, embed_dim=60, depths=[2], num_heads=[6],
                                       window_size=8, mlp_ratio=2, upscale=2, img_range=1.0,
                                       upsampler="pixelshuffledirect", resi_connection="1conv"), x=(1, 3, 16, 16),
                          seed=7, loss="l1"),
    # '3conv' residual connection (SwinIR-L style), C=48
    "swinir_3conv": dict(_SW, kw=dict(img_size=16, in_chans=3, embed_dim=48, depths=[2], num_heads=[6],
                                      window_size=8, mlp_ratio=2, upscale=1, img_range=1.0, upsampler="",
                                      resi_connection="3conv"), x=(1, 3, 16, 16), seed=8, loss="l1"),
    # SwinIR-L geometry at reduced depth (real-world SR large model, main_test_swinir.py:163-168): C=240, 8 heads,
    # '3conv', nearest+conv x4
    "swinir_l_small": dict(_SW, kw=dict(img_size=16, in_chans=3, embed_dim=240, depths=[2], num_heads=[8],
                                        window_size=8, mlp_ratio=2, upscale=4, img_range=1.0, upsampler="nearest+conv",
                                        resi_connection="3conv"), x=(1, 3, 16, 16), seed=14, loss="l1"),
    # JPEG-artifact model geometry: window 7, img_range 255
    "swinir_jpeg_ws7": dict(_SW, kw=dict(img_size=28, in_chans=1, embed_dim=180, depths=[2], num_heads=[6],
                                         window_size=7, mlp_ratio=2, upscale=1, img_range=255.0, upsampler="",
                                         resi_connection="1conv"), x=(1, 1, 28, 28), seed=9, loss="charbonnier"),
    # DnCNN-17 'R' (the inference checkpoint's architecture, cfg1)
    "dncnn_r17": dict(net="dncnn", kw=dict(in_nc=1, out_nc=1, nc=64, nb=17, act_mode="R"), x=(1, 1, 40, 40), seed=10,
                      loss="l1"),
    # DnCNN 'BR' in training mode: batch-statistics BN (momentum 0.9, eps 1e-4) + running-stat update
    "dncnn_br_train": dict(net="dncnn", kw=dict(in_nc=1, out_nc=1, nc=64, nb=5, act_mode="BR"), x=(4, 1, 24, 24),
                           seed=11, loss="l1", train_mode=True),
    "dncnn_br_eval": dict(net="dncnn", kw=dict(in_nc=3, out_nc=3, nc=64, nb=5, act_mode="BR"), x=(2, 3, 24, 24),
                          seed=12, loss="l1"),
    # FDnCNN (network_dncnn.py:126-148): noise-level map as second input channel, no residual learning
    "fdncnn_r8": dict(net="fdncnn", kw=dict(in_nc=2, out_nc=1, nc=64, nb=8, act_mode="R"), x=(2, 2, 24, 32), seed=15,
                      loss="l1"),
    # IRCNN (network_dncnn.py:77-116): dilated 3x3 stack 1-2-3-4-3-2-1, the plug-and-play denoiser prior
    "ircnn": dict(net="ircnn", kw=dict(in_nc=3, out_nc=3, nc=64), x=(2, 3, 24, 40), seed=17, loss="l1"),
    # FFDNet colour geometry (network_ffdnet.py:27-72: nc=96 -> 128 padded channels), odd image size (replicate pad)
    "ffdnet_c96": dict(net="ffdnet", kw=dict(in_nc=3, out_nc=3, nc=96, nb=5, act_mode="R"), x=(2, 3, 23, 30), seed=18,
                       loss="l1"),
    # RRDBNet_noUp (the fork's `rrdbnet_no_up`, SPECToptions/network_configs.json:16-31: 2 SPECT views in / out), nb reduced
    "rrdbnet_nb2": dict(net="rrdbnet", kw=dict(in_nc=2, out_nc=2, nf=64, nb=2, gc=32), x=(1, 2, 24, 32), seed=16, loss="l1"),
    # DRUNet (cfg4 architecture, nb reduced to 2)
    "unetres_nb2": dict(net="unetres", kw=dict(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=2, act_mode="R",
                                               downsample_mode="strideconv", upsample_mode="convtranspose",
                                               bias=False), x=(1, 4, 32, 32), seed=13, loss="l1"),
}


def sigma_of(case):
    """noise levels (B, 1, 1, 1) of an FFDNet case: 15/255 .. 50/255, one per sample"""
    B = case["x"][0]
    return (torch.linspace(15.0, 50.0, B) / 255.0).view(B, 1, 1, 1)


def loss_fn(case):
    if case["loss"] == "charbonnier":
        return so.charbonnier
    return lambda a, b: torch.mean(torch.abs(a - b))


def swinir_cfg(kw):
    return so.SwinIRConfig(img_size=kw["img_size"], in_chans=kw["in_chans"], embed_dim=kw["embed_dim"],
                           depths=tuple(kw["depths"]), num_heads=tuple(kw["num_heads"]),
                           window_size=kw["window_size"], mlp_ratio=kw["mlp_ratio"], upscale=kw["upscale"],
                           img_range=kw["img_range"], upsampler=kw["upsampler"],
                           resi_connection=kw["resi_connection"])


def state_and_io(case):
    """(state_dict, x, target) of a case -- identical on every machine."""
    kw, net = case["kw"], case["net"]
    if net == "swinir":
        cfg = swinir_cfg(kw)
        sd = so.deterministic_state(so.state_spec(cfg), case["seed"], cfg)
        up = kw["upscale"]
    elif net in ("dncnn", "fdncnn"):
        sd = so.deterministic_state(co.dncnn_state_spec(**kw), case["seed"])
        up = 1
    elif net == "rrdbnet":
        sd = so.deterministic_state(co.rrdbnet_state_spec(**kw), case["seed"])
        up = 1
    elif net == "ircnn":
        sd = so.deterministic_state(co.ircnn_state_spec(**kw), case["seed"])
        up = 1
    elif net == "ffdnet":
        sd = so.deterministic_state(co.ffdnet_state_spec(**kw), case["seed"])
        up = 1
    else:
        sd = so.deterministic_state(co.unetres_state_spec(kw["in_nc"], kw["out_nc"], tuple(kw["nc"]), kw["nb"],
                                                          kw["bias"]), case["seed"])
        up = 1
    g = torch.Generator().manual_seed(1000 + case["seed"])
    B, c, H, W = case["x"]
    x = torch.rand(B, c, H, W, generator=g)
    co_ = kw.get("out_nc", c)
    tgt = torch.rand(B, co_, H * up, W * up, generator=g)
    return sd, x, tgt


def oracle_forward(case, sd, x, new_stats=None):
    kw, net = case["kw"], case["net"]
    if net == "swinir":
        return so.swinir_forward(sd, x, swinir_cfg(kw))
    if net in ("dncnn", "fdncnn"):
        return co.dncnn_forward(sd, x, nb=kw["nb"], act_mode=kw["act_mode"], training=case.get("train_mode", False),
                                new_stats=new_stats, residual=net == "dncnn")
    if net == "rrdbnet":
        return co.rrdbnet_noup_forward(sd, x, nb=kw["nb"])
    if net == "ircnn":
        return co.ircnn_forward(sd, x)
    if net == "ffdnet":
        return co.ffdnet_forward(sd, x, sigma_of(case), nb=kw["nb"])
    return co.unetres_forward(sd, x, nb=kw["nb"])


def build_reference(case, ref_loader):
    """The real reference module with the case's weights loaded (needs /root/reference)."""
    kw, net = case["kw"], case["net"]
    sd, x, tgt = state_and_io(case)
    if net == "swinir":
        model = ref_loader.load("network_swinir").SwinIR(drop_path_rate=0.0, **kw)
    elif net == "dncnn":
        model = ref_loader.load("network_dncnn").DnCNN(**kw)
    elif net == "fdncnn":
        model = ref_loader.load("network_dncnn").FDnCNN(**kw)
    elif net == "rrdbnet":
        model = ref_loader.load("network_rrdbnet").RRDBNet_noUp(**kw)
    elif net == "ircnn":
        model = ref_loader.load("network_dncnn").IRCNN(**kw)
    elif net == "ffdnet":
        ffd = ref_loader.load("network_ffdnet").FFDNet(**kw)
        sig = sigma_of(case)

        class _WithSigma(torch.nn.Module):      # the golden harness calls model(x)
            def __init__(self):
                super().__init__()
                self.net = ffd

            def forward(self, x):
                return self.net(x, sig)

            def load_state_dict(self, sd, strict=True):
                return self.net.load_state_dict(sd, strict=strict)

            def named_parameters(self, *a, **k):
                return self.net.named_parameters(*a, **k)

            def state_dict(self, *a, **k):
                return self.net.state_dict(*a, **k)
        model = _WithSigma()
    else:
        model = ref_loader.load("network_unet").UNetRes(**kw)
    missing = model.load_state_dict(sd, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return model, sd, x, tgt


def dncnn_cfg1_images(size=256, sigma=25.0):
    """cfg1 protocol (main_test_dncnn.py:151-152 of the reference): a clean synthetic image in [0,1] (sum of sinusoids
    and an edge), plus np.random.seed(0) Gaussian noise of level sigma/255.  Returns (clean, noisy) as (1,1,H,W) fp32."""
    import numpy as np
    yy, xx = np.meshgrid(np.arange(size, dtype=np.float64), np.arange(size, dtype=np.float64), indexing="ij")
    clean = 0.5 + 0.2 * np.sin(2 * np.pi * xx / 64.0) * np.cos(2 * np.pi * yy / 48.0) + 0.15 * np.sin(2 * np.pi * (xx + yy) / 23.0)
    clean += 0.1 * ((xx > size * 0.6) & (yy > size * 0.3))
    clean = np.clip(clean, 0.0, 1.0)
    np.random.seed(seed=0)
    noisy = clean + np.random.normal(0, sigma / 255.0, clean.shape)
    to_t = lambda a: torch.from_numpy(a.astype(np.float32))[None, None]
    return to_t(clean), to_t(noisy)


def psnr_uint8(a, b):
    """utils/utils_image.py:296-300, 629-645 of the reference: PSNR on round(clip(x,0,1)*255) uint8 images."""
    qa = (a.clamp(0, 1) * 255.0).round()
    qb = (b.clamp(0, 1) * 255.0).round()
    mse = torch.mean((qa - qb) ** 2).item()
    return float("inf") if mse == 0 else 20.0 * math.log10(255.0 / math.sqrt(mse))
