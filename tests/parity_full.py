"""TEST INFRASTRUCTURE: full-architecture parity measurements of the CUDA path against the CPU oracle at the
BASELINE.json configurations themselves (cfg2 .. cfg5), shared by tests/test_parity_full_gpu.py (asserts) and
tools/parity_table.py (writes profiles/parity_r2.json).  Nothing here is imported by the product package.

Each `measure_*` builds the drop-in module at its full depth with deterministic weights, runs the oracle
(oracle/swinir_oracle.py, oracle/convnets_oracle.py: restatements of models/network_swinir.py:805-839 and
models/network_unet.py:59-76 of the reference) on the host cores and the CUDA path on the same inputs, and returns
a dict of error figures: forward max-abs, uint8 PSNR difference, and -- for the training configs -- per-tensor
gradient errors (relative L2, max-abs error over max-abs reference, cosine) with the worst tensor named."""
import time

import torch

from oracle import convnets_oracle as co
from oracle import swinir_oracle as so

SWINIR_M = dict(window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180, num_heads=[6] * 6, mlp_ratio=2,
                resi_connection="1conv")


def _ocfg(kw):
    return so.SwinIRConfig(img_size=kw["img_size"], in_chans=kw["in_chans"], embed_dim=kw["embed_dim"],
                           depths=tuple(kw["depths"]), num_heads=tuple(kw["num_heads"]), window_size=kw["window_size"],
                           mlp_ratio=kw["mlp_ratio"], upscale=kw["upscale"], img_range=kw["img_range"],
                           upsampler=kw["upsampler"], resi_connection=kw["resi_connection"])


def grad_errors(named_grads, ref_grads):
    """per-tensor error figures; returns (table, summary)"""
    rows = {}
    for k, g in named_grads.items():
        a, b = g.detach().double().cpu().flatten(), ref_grads[k].detach().double().flatten()
        nb = b.norm().item()
        rel = ((a - b).norm().item() / nb) if nb > 0 else (a.norm().item())
        mx = ((a - b).abs().max().item() / b.abs().max().item()) if nb > 0 else a.abs().max().item()
        cos = (torch.dot(a, b).item() / (a.norm().item() * nb)) if nb > 0 and a.norm().item() > 0 else 1.0
        rows[k] = (rel, mx, cos)
    worst = max(rows, key=lambda k: rows[k][0])
    worst_mx = max(rows, key=lambda k: rows[k][1])
    allg = torch.cat([g.detach().double().cpu().flatten() for g in named_grads.values()])
    allr = torch.cat([ref_grads[k].detach().double().flatten() for k in named_grads])
    summary = {
        "grad_rel_l2_all": ((allg - allr).norm() / allr.norm()).item(),
        "grad_rel_l2_worst": rows[worst][0], "grad_rel_l2_worst_tensor": worst,
        "grad_maxrel_worst": rows[worst_mx][1], "grad_maxrel_worst_tensor": worst_mx,
        "grad_cos_min": min(r[2] for r in rows.values()),
        "grad_rel_l2_median": sorted(r[0] for r in rows.values())[len(rows) // 2],
        "n_tensors": len(rows),
    }
    return rows, summary


def _fwd_errors(y, y_ref, tgt):
    """tgt is replaced by a 'ground truth' 30 dB away from the oracle output (sigma 0.03 noise), so that the uint8
    PSNR difference is measured where restoration PSNRs live (random-init nets are ~8 dB from any random target)"""
    y, y_ref = y.detach().float().cpu(), y_ref.detach().float()
    g = torch.Generator().manual_seed(977)
    tgt = (y_ref + 0.03 * torch.randn(y_ref.shape, generator=g)).clamp(0, 1)
    return {"fwd_max_abs": (y - y_ref).abs().max().item(),
            "fwd_rel_l2": ((y - y_ref).norm() / y_ref.norm()).item(),
            "psnr_ref": so.psnr_uint8(y_ref, tgt), "psnr_cuda": so.psnr_uint8(y, tgt),
            "dpsnr_uint8": abs(so.psnr_uint8(y, tgt) - so.psnr_uint8(y_ref, tgt))}


def _swinir(kw, seed, precision):
    from kair_b200.network_swinir import SwinIR
    cfg = _ocfg(kw)
    sd = so.deterministic_state(so.state_spec(cfg), seed, cfg)
    model = SwinIR(drop_path_rate=0.0, **kw)
    model.load_state_dict(sd, strict=True)
    if precision != "bf16":
        model.set_precision(precision)
    return model, sd, cfg


def measure_cfg2(precision="bf16", B=2, size=128):
    """SwinIR-M gray denoising training (6x6 RSTB, 1 x 128 x 128 patches, Charbonnier): forward + every gradient."""
    kw = dict(SWINIR_M, upscale=1, in_chans=1, img_size=size, upsampler="")
    model, sd, cfg = _swinir(kw, 31, precision)
    # north_star states the forward bounds on [0,1] images: scale conv_last so that the net is a denoiser-like
    # x + residual with a residual of the sigma=25/255 noise level (std ~0.1) instead of ~0.3 at the seeded init
    with torch.no_grad():
        for k in ("conv_last.weight", "conv_last.bias"):
            sd[k] = sd[k] * 0.3
        model.load_state_dict(sd, strict=True)
    g = torch.Generator().manual_seed(131)
    x, tgt = torch.rand(B, 1, size, size, generator=g), torch.rand(B, 1, size, size, generator=g)
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    t0 = time.time()
    y_ref = so.swinir_forward(sdg, x, cfg)
    so.charbonnier(y_ref, tgt).backward()
    t_cpu = time.time() - t0
    model = model.cuda().train()
    y = model(x.cuda())
    so.charbonnier(y, tgt.cuda()).backward()
    torch.cuda.synchronize()
    out = {"config": "cfg2 SwinIR-M gray-DN training, B=%d 1x%dx%d, depths 6x6, Charbonnier" % (B, size, size),
           "precision": precision, "oracle_seconds": round(t_cpu, 1)}
    out.update(_fwd_errors(y, y_ref, tgt))
    rows, summ = grad_errors({k: p.grad for k, p in model.named_parameters()},
                             {k: sdg[k].grad for k, _ in model.named_parameters()})
    out.update(summ)
    return out, rows


def measure_cfg3(precision="bf16", B=2, size=64):
    """SwinIR-M classical SR x4 (pixelshuffle), 3 x 64 x 64 LR patches, inference."""
    kw = dict(SWINIR_M, upscale=4, in_chans=3, img_size=size, upsampler="pixelshuffle")
    model, sd, cfg = _swinir(kw, 32, precision)
    g = torch.Generator().manual_seed(132)
    x, tgt = torch.rand(B, 3, size, size, generator=g), torch.rand(B, 3, 4 * size, 4 * size, generator=g)
    t0 = time.time()
    with torch.no_grad():
        y_ref = so.swinir_forward(sd, x, cfg)
    t_cpu = time.time() - t0
    model = model.cuda().eval()
    with torch.no_grad():
        y = model(x.cuda())
    out = {"config": "cfg3 SwinIR-M SR x4 pixelshuffle inference, B=%d 3x%dx%d, depths 6x6" % (B, size, size),
           "precision": precision, "oracle_seconds": round(t_cpu, 1)}
    out.update(_fwd_errors(y, y_ref, tgt))
    return out, {}


def measure_cfg4(precision="bf16", B=1, size=128):
    """DRUNet (UNetRes nb=4, strideconv / convtranspose, bias=False) training on 4 x 128 x 128 (image + noise map)."""
    from kair_b200.network_unet import UNetRes
    kw = dict(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", downsample_mode="strideconv",
              upsample_mode="convtranspose", bias=False)
    sd = so.deterministic_state(co.unetres_state_spec(4, 3, (64, 128, 256, 512), 4, False), 33)
    model = UNetRes(**kw)
    model.load_state_dict(sd, strict=True)
    if precision != "bf16":
        model.set_precision(precision)
    g = torch.Generator().manual_seed(133)
    x, tgt = torch.rand(B, 4, size, size, generator=g), torch.rand(B, 3, size, size, generator=g)
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    t0 = time.time()
    y_ref = co.unetres_forward(sdg, x, nb=4)
    torch.mean(torch.abs(y_ref - tgt)).backward()
    t_cpu = time.time() - t0
    model = model.cuda().train()
    y = model(x.cuda())
    torch.mean(torch.abs(y - tgt.cuda())).backward()
    torch.cuda.synchronize()
    out = {"config": "cfg4 DRUNet nb=4 training, B=%d 4x%dx%d, L1" % (B, size, size), "precision": precision,
           "oracle_seconds": round(t_cpu, 1)}
    out.update(_fwd_errors(y, y_ref, tgt))
    rows, summ = grad_errors({k: p.grad for k, p in model.named_parameters()},
                             {k: sdg[k].grad for k, _ in model.named_parameters()})
    out.update(summ)
    return out, rows


def reference_tiled_loop(img_lq, model, tile, tile_overlap, sf):
    """Literal restatement of main_test_swinir.py:256-284 (`test` with args.tile set): one tile per forward."""
    b, c, h, w = img_lq.size()
    tile = min(tile, h, w)
    stride = tile - tile_overlap
    h_idx_list = list(range(0, h - tile, stride)) + [h - tile]
    w_idx_list = list(range(0, w - tile, stride)) + [w - tile]
    E = torch.zeros(b, c, h * sf, w * sf).type_as(img_lq)
    W = torch.zeros_like(E)
    for h_idx in h_idx_list:
        for w_idx in w_idx_list:
            out_patch = model(img_lq[..., h_idx:h_idx + tile, w_idx:w_idx + tile])
            E[..., h_idx * sf:(h_idx + tile) * sf, w_idx * sf:(w_idx + tile) * sf].add_(out_patch)
            W[..., h_idx * sf:(h_idx + tile) * sf, w_idx * sf:(w_idx + tile) * sf].add_(torch.ones_like(out_patch))
    return E.div_(W)


def measure_cfg5(precision="bf16", size=280, tile=256, overlap=32):
    """SwinIR-M real-world SR x4 (nearest+conv) tiled inference: flip padding to the window, 256^2 tiles with 32 px
    overlap, overlap-add -- kair_b200.tiling.tiled_forward on the CUDA path vs the literal loop on the oracle."""
    from kair_b200 import tiling
    kw = dict(SWINIR_M, upscale=4, in_chans=3, img_size=64, upsampler="nearest+conv")
    model, sd, cfg = _swinir(kw, 34, precision)
    g = torch.Generator().manual_seed(134)
    img = torch.rand(1, 3, size, size, generator=g)
    tgt = torch.rand(1, 3, 4 * size, 4 * size, generator=g)
    padded = tiling.pad_to_window(img, 8)
    t0 = time.time()
    with torch.no_grad():
        y_ref = reference_tiled_loop(padded, lambda p: so.swinir_forward(sd, p, cfg), tile, overlap, 4)
    t_cpu = time.time() - t0
    model = model.cuda().eval()
    with torch.no_grad():
        y = tiling.tiled_forward(model, padded.cuda(), tile, overlap, 4, 8, tile_batch=4)
    y, y_ref = y[..., :4 * size, :4 * size], y_ref[..., :4 * size, :4 * size]
    n_tiles = len(tiling.tile_origins(padded.shape[2], padded.shape[3], tile, overlap))
    out = {"config": "cfg5 SwinIR-M real-SR x4 nearest+conv tiled inference, %dx%d image -> %d tiles of %d^2, "
                     "overlap %d" % (size, size, n_tiles, tile, overlap),
           "precision": precision, "oracle_seconds": round(t_cpu, 1)}
    out.update(_fwd_errors(y, y_ref, tgt))
    return out, {}


MEASURES = {"cfg2": measure_cfg2, "cfg3": measure_cfg3, "cfg4": measure_cfg4, "cfg5": measure_cfg5}
