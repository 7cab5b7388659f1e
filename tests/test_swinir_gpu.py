"""Parity of the drop-in SwinIR module (CUDA path, through the C ABI) against the CPU oracle on the same seeded
inputs and weights.  Tolerances: bf16 tensor-core mode, |out - ref|_max <= 2e-2 on [0,1] images (north_star);
gradients are compared in relative L2 (<= 3e-2) because the fp32-mode 1e-3 bound does not apply to bf16 operands."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

CFG = dict(upscale=1, in_chans=1, img_size=32, window_size=8, img_range=1.0, depths=[2, 2], embed_dim=180,
           num_heads=[6, 6], mlp_ratio=2, upsampler="", resi_connection="1conv")


def _oracle_cfg(cfg):
    from oracle import swinir_oracle as so
    return so.SwinIRConfig(img_size=cfg["img_size"], in_chans=cfg["in_chans"], embed_dim=cfg["embed_dim"],
                           depths=tuple(cfg["depths"]), num_heads=tuple(cfg["num_heads"]),
                           window_size=cfg["window_size"], mlp_ratio=cfg["mlp_ratio"], upscale=cfg["upscale"],
                           img_range=cfg["img_range"], upsampler=cfg["upsampler"])


def _randomize(model, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for k, p in model.named_parameters():
            if "norm" in k and k.endswith("weight"):
                p.copy_(1.0 + 0.2 * torch.randn(p.shape, generator=g))
            elif k.endswith("bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif k.endswith("relative_position_bias_table"):
                p.copy_(0.3 * torch.randn(p.shape, generator=g))
            else:
                p.mul_(1.5)


def _rel_l2(a, b):
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


@pytest.mark.parametrize("in_chans,B,H,W,C", [(1, 2, 32, 32, 180), (3, 1, 24, 40, 180), (2, 3, 16, 16, 180),
                                              (3, 2, 32, 32, 60), (1, 1, 16, 24, 60)])
def test_forward_backward_vs_oracle(in_chans, B, H, W, C):
    from kair_b200.network_swinir import SwinIR
    from oracle import swinir_oracle as so
    cfg = dict(CFG, in_chans=in_chans, embed_dim=C)   # C=60: SwinIR-S geometry (64 padded channels, 64-wide tiles)
    torch.manual_seed(0)
    model = SwinIR(drop_path_rate=0.0, **cfg)
    _randomize(model, 11)
    sd = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in model.state_dict().items()}
    g = torch.Generator().manual_seed(5)
    x = torch.rand(B, in_chans, H, W, generator=g)
    tgt = torch.rand(B, in_chans, H, W, generator=g)
    y_ref = so.swinir_forward(sd, x, _oracle_cfg(cfg))
    torch.mean((y_ref - tgt) ** 2).backward()

    model = model.cuda().train()
    y = model(x.cuda())
    torch.mean((y - tgt.cuda()) ** 2).backward()
    torch.cuda.synchronize()
    err = (y.detach().cpu() - y_ref.detach()).abs().max().item()
    assert err < 2e-2, err
    bad = []
    for k, p in model.named_parameters():
        r = _rel_l2(p.grad.cpu(), sd[k].grad)
        if not r < 3e-2:
            bad.append((k, r))
    assert not bad, bad[:8]


def test_eval_matches_train_forward_and_pads():
    """no_grad/eval plan (shared activation buffers) == training plan; odd sizes are reflect-padded and cropped."""
    from kair_b200.network_swinir import SwinIR
    from oracle import swinir_oracle as so
    torch.manual_seed(0)
    model = SwinIR(drop_path_rate=0.1, **CFG)
    _randomize(model, 12)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    x = torch.rand(2, 1, 27, 35, generator=torch.Generator().manual_seed(6))
    y_ref = so.swinir_forward(sd, x, _oracle_cfg(CFG))
    model = model.cuda().eval()
    with torch.no_grad():
        y_eval = model(x.cuda())
    y_train = model(x.cuda())          # grad enabled, eval mode: DropPath off
    assert y_eval.shape == (2, 1, 27, 35)
    assert (y_eval.cpu() - y_ref).abs().max().item() < 2e-2
    assert torch.equal(y_eval, y_train.detach())


def test_drop_path_scales_replayed():
    """DropPath: inject the per-sample scales, compare with the oracle replaying the same masks."""
    from kair_b200.network_swinir import SwinIR
    from oracle import swinir_oracle as so
    torch.manual_seed(0)
    model = SwinIR(drop_path_rate=0.1, **CFG)
    _randomize(model, 13)
    sd = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in model.state_dict().items()}
    g = torch.Generator().manual_seed(7)
    B, L = 4, 4
    x = torch.rand(B, 1, 32, 32, generator=g)
    drop = (torch.rand(2 * L, B, generator=g) < 0.8).float() / 0.8
    y_ref = so.swinir_forward(sd, x, _oracle_cfg(CFG), drop_masks=[(drop[l], drop[L + l]) for l in range(L)])
    y_ref.square().mean().backward()
    model = model.cuda().train()
    eng = model.engine(torch.device("cuda:0"))
    plan = eng.plan(B, 32, 32, True)
    from kair_b200.swinir_engine import SwinIRFunction
    plan.dp.copy_(drop.cuda())
    y = SwinIRFunction.apply(x.cuda(), plan, *eng.P.values())
    y.square().mean().backward()
    assert (y.detach().cpu() - y_ref.detach()).abs().max().item() < 2e-2
    bad = [(k, _rel_l2(p.grad.cpu(), sd[k].grad)) for k, p in model.named_parameters()
           if not _rel_l2(p.grad.cpu(), sd[k].grad) < 5e-2]
    assert not bad, bad[:8]


def test_cpu_tensor_is_rejected():
    from kair_b200.network_swinir import SwinIR
    model = SwinIR(drop_path_rate=0.0, **CFG)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        model(torch.rand(1, 1, 32, 32))


def test_full_size_properties():
    """cfg2 shape (B=8, 1x128x128, SwinIR-M): finite outputs, identity-at-zero-weights property, determinism."""
    from kair_b200.network_swinir import SwinIR
    torch.manual_seed(0)
    model = SwinIR(upscale=1, in_chans=1, img_size=128, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
                   num_heads=[6] * 6, mlp_ratio=2, upsampler="", resi_connection="1conv").cuda().train()
    x = torch.rand(8, 1, 128, 128, device="cuda")
    y1 = model(x)
    y1.square().mean().backward()
    g1 = model.conv_first.weight.grad.clone()
    assert torch.isfinite(y1).all() and all(torch.isfinite(p.grad).all() for p in model.parameters())
    model.zero_grad()
    model.eval()
    with torch.no_grad():
        y2 = model(x)
    # conv_last zeroed => network output == input (x + conv_last(res) with zero weights), exactly
    with torch.no_grad():
        model.conv_last.weight.zero_(); model.conv_last.bias.zero_()
        y3 = model(x)
    assert torch.equal(y3, x)
    assert torch.isfinite(y2).all() and g1.abs().sum().item() > 0


def test_golden_fixture_swinir_m_gray32():
    """CUDA path vs the committed golden vector produced by the real reference module (tests/golden/)."""
    import json
    import os
    import numpy as np
    import golden_cases as gc
    from kair_b200.network_swinir import SwinIR
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "swinir_m_gray32.npz"))
    case = gc.CASES["swinir_m_gray32"]
    sd, x, tgt = gc.state_and_io(case)
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().train()
    y = model(x.cuda())
    loss = gc.loss_fn(case)(y, tgt.cuda())
    loss.backward()
    assert np.abs(y.detach().cpu().numpy() - z["y"]).max() < 2e-2
    assert abs(loss.item() - float(z["loss"])) < 2e-3
    grads = json.loads(str(z["grads"]))
    bad = []
    for k, p in model.named_parameters():
        n = p.grad.double().norm().item()
        if abs(n - grads[k][0]) > 3e-2 * grads[k][0] + 1e-6:
            bad.append((k, n, grads[k][0]))
    assert not bad, bad[:6]


@pytest.mark.parametrize("name", ["swinir_sr_ps4", "swinir_sr_nearest4", "swinir_s_psd2", "swinir_3conv", "swinir_l_small", "swinir_jpeg_ws7"])
def test_golden_fixture_sr_heads(name):
    """cfg3 / cfg5 architectures (reduced depth): x4 pixelshuffle and nearest+conv heads under no_grad vs the golden
    output of the real reference module; bf16 result vs fp32 reference: 2e-2 max-abs and 0.02 dB PSNR (north_star)."""
    import os
    import numpy as np
    import golden_cases as gc
    from kair_b200.network_swinir import SwinIR
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    case = gc.CASES[name]
    sd, x, tgt = gc.state_and_io(case)
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().eval()
    with torch.no_grad():
        y = model(x.cuda()).cpu()
    yr = torch.from_numpy(z["y"])
    assert y.shape == yr.shape
    assert (y - yr).abs().max().item() < 2e-2
    assert abs(gc.psnr_uint8(y, tgt) - gc.psnr_uint8(yr, tgt)) < 0.02


def test_sr_x4_full_size_cfg3():
    """cfg3 shape: SwinIR-M x4 pixelshuffle, batch 32 of 3x64x64 LR patches, no_grad; batch-split invariance."""
    from kair_b200.network_swinir import SwinIR
    torch.manual_seed(0)
    model = SwinIR(upscale=4, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
                   num_heads=[6] * 6, mlp_ratio=2, upsampler="pixelshuffle", resi_connection="1conv").cuda().eval()
    x = torch.rand(32, 3, 64, 64, device="cuda")
    with torch.no_grad():
        y = model(x)
        y4 = model(x[4:8])
    assert y.shape == (32, 3, 256, 256) and torch.isfinite(y).all()
    assert torch.equal(y[4:8], y4)


@pytest.mark.parametrize("name", ["swinir_sr_ps4", "swinir_sr_nearest4", "swinir_s_psd2", "swinir_3conv", "swinir_l_small", "swinir_jpeg_ws7"])
def test_sr_heads_training_vs_oracle(name):
    """Autograd through the reconstruction heads (classical / lightweight / real-world SR training,
    options/swinir/train_swinir_sr_*.json): output and every parameter gradient vs autograd through the oracle."""
    import golden_cases as gc
    from kair_b200.network_swinir import SwinIR
    case = gc.CASES[name]
    sd, x, tgt = gc.state_and_io(case)
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    y_ref = gc.oracle_forward(case, sdg, x)
    torch.mean((y_ref - tgt) ** 2).backward()
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().train()
    y = model(x.cuda())
    torch.mean((y - tgt.cuda()) ** 2).backward()
    torch.cuda.synchronize()
    assert (y.detach().cpu() - y_ref.detach()).abs().max().item() < 2e-2
    # the 12-channel LeakyReLU bottleneck of the C=48 '3conv' case amplifies bf16 rounding upstream of it; the host
    # logic of every case is pinned exactly (fp32 buffers) in tests/test_swinir_engine_cpu.py
    wide = 2.5e-1 if name == "swinir_3conv" else 5e-2
    tol = lambda k: max(wide, 1e-1) if k.endswith("relative_position_bias_table") else wide
    bad = []
    for k, p in model.named_parameters():
        r = _rel_l2(p.grad.cpu(), sdg[k].grad)
        if not r < tol(k):
            bad.append((k, r))
    assert not bad, bad[:8]


def test_sr_x2_training_step_cfg_classical():
    """train_swinir_sr_classical.json geometry (SwinIR-M x2, 48x48 LR patches, batch 4): two Adam steps reduce the
    loss on a fixed batch and every gradient is finite."""
    from kair_b200.network_swinir import SwinIR
    torch.manual_seed(0)
    model = SwinIR(upscale=2, in_chans=3, img_size=48, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
                   num_heads=[6] * 6, mlp_ratio=2, upsampler="pixelshuffle", resi_connection="1conv").cuda().train()
    opt = torch.optim.Adam(model.parameters(), lr=2e-4)
    g = torch.Generator().manual_seed(3)
    x = torch.rand(4, 3, 48, 48, generator=g).cuda()
    t = torch.rand(4, 3, 96, 96, generator=g).cuda()
    losses = []
    for _ in range(3):
        opt.zero_grad()
        loss = torch.mean(torch.abs(model(x) - t))
        loss.backward()
        assert all(torch.isfinite(p.grad).all().item() for p in model.parameters())
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0], losses


# every model main_test_swinir.py:124-175 of the reference can construct (define_model), at its full architecture
TASKS = {
    "classical_sr_x2": dict(upscale=2, in_chans=3, img_size=48, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
                            num_heads=[6] * 6, mlp_ratio=2, upsampler="pixelshuffle", resi_connection="1conv"),
    "classical_sr_x3": dict(upscale=3, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
                            num_heads=[6] * 6, mlp_ratio=2, upsampler="pixelshuffle", resi_connection="1conv"),
    "classical_sr_x8": dict(upscale=8, in_chans=3, img_size=48, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
                            num_heads=[6] * 6, mlp_ratio=2, upsampler="pixelshuffle", resi_connection="1conv"),
    "lightweight_sr_x3": dict(upscale=3, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6] * 4, embed_dim=60,
                              num_heads=[6] * 4, mlp_ratio=2, upsampler="pixelshuffledirect", resi_connection="1conv"),
    "real_sr_x4_large": dict(upscale=4, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6] * 9, embed_dim=240,
                             num_heads=[8] * 9, mlp_ratio=2, upsampler="nearest+conv", resi_connection="3conv"),
    "color_dn": dict(upscale=1, in_chans=3, img_size=128, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180,
                     num_heads=[6] * 6, mlp_ratio=2, upsampler="", resi_connection="1conv"),
    "jpeg_car": dict(upscale=1, in_chans=1, img_size=126, window_size=7, img_range=255.0, depths=[6] * 6, embed_dim=180,
                     num_heads=[6] * 6, mlp_ratio=2, upsampler="", resi_connection="1conv"),
}


@pytest.mark.parametrize("task", sorted(TASKS))
def test_every_published_task_runs(task):
    """Full-depth models of every task: inference on an input that is not a window multiple, batch-split invariance
    (bit-exact: no cross-sample arithmetic), and one training step with finite gradients for all parameters."""
    from kair_b200.network_swinir import SwinIR
    kw = TASKS[task]
    torch.manual_seed(0)
    model = SwinIR(**kw).cuda().eval()
    g = torch.Generator().manual_seed(2)
    x = torch.rand(2, kw["in_chans"], 37, 45, generator=g).cuda()
    with torch.no_grad():
        y = model(x)
        y0 = model(x[:1])
    assert y.shape == (2, kw["in_chans"], 37 * kw["upscale"], 45 * kw["upscale"])
    assert torch.isfinite(y).all().item()
    assert torch.equal(y[:1], y0)
    model.train()
    ws = kw["window_size"]
    xt = torch.rand(2, kw["in_chans"], 4 * ws, 4 * ws, generator=g).cuda()
    out = model(xt)
    out.abs().mean().backward()
    for k, p in model.named_parameters():
        assert p.grad is not None and torch.isfinite(p.grad).all().item(), k
    assert sum(p.grad.abs().sum().item() for p in model.parameters()) > 0


def test_amp_gradscaler_clip_and_frozen_params():
    """The training loop of models/model_plain.py:273-300 of the reference around the drop-in module: autocast +
    GradScaler (scaled loss -> unscale_ -> clip_grad_norm_ -> step), use_checkpoint=True accepted, and parameters with
    requires_grad=False (main_test_dncnn.py:123-124) get no gradient while the rest match the unscaled run."""
    from kair_b200.network_swinir import SwinIR
    cfg = dict(CFG, depths=[2], num_heads=[6], use_checkpoint=True)
    torch.manual_seed(0)
    model = SwinIR(drop_path_rate=0.0, **cfg).cuda().train()
    _randomize(model, 21)
    g = torch.Generator().manual_seed(8)
    x = torch.rand(2, 1, 32, 32, generator=g).cuda()
    t = torch.rand(2, 1, 32, 32, generator=g).cuda()
    charb = lambda a, b: torch.mean(torch.sqrt((a - b) ** 2 + 1e-9))

    charb(model(x), t).backward()
    plain = {k: p.grad.clone() for k, p in model.named_parameters()}
    model.zero_grad(set_to_none=True)

    frozen = [k for k in plain if "blocks.0.mlp" in k]
    for k, p in model.named_parameters():
        p.requires_grad_(k not in frozen)
    opt = torch.optim.Adam([p for p in model.parameters() if p.requires_grad], lr=2e-4)
    scaler = torch.amp.GradScaler("cuda", init_scale=1024.0)
    before = {k: p.detach().clone() for k, p in model.named_parameters()}
    with torch.autocast("cuda", dtype=torch.float16):
        out = model(x)
        loss = charb(out, t)
    assert out.dtype == torch.float32
    scaler.scale(loss).backward()
    scaler.unscale_(opt)
    total = torch.nn.utils.clip_grad_norm_(model.parameters(), max_norm=1e9)
    assert torch.isfinite(total).item()
    for k, p in model.named_parameters():
        if k in frozen:
            assert p.grad is None, k
        else:
            assert _rel_l2(p.grad, plain[k]) < 2e-2, (k, _rel_l2(p.grad, plain[k]))
    scaler.step(opt)
    scaler.update()
    for k, p in model.named_parameters():
        assert torch.equal(p.detach(), before[k]) == (k in frozen or plain[k].abs().max().item() == 0), k


@pytest.mark.parametrize("precision,tol", [("bf16", 1.5e-1), ("fp32", 1e-3)])
def test_input_gradient_and_two_forwards(precision, tol):
    """d(loss)/d(x) through the module (frozen net, plug-and-play use) vs autograd through the oracle, and two forwards
    of one shape before a single backward (two activation sets per shape)."""
    import golden_cases as gc
    from kair_b200.network_swinir import SwinIR
    case = gc.CASES["swinir_m_gray32"]
    sd, x, tgt = gc.state_and_io(case)
    xr = x.clone().requires_grad_(True)
    gc.loss_fn(case)(gc.oracle_forward(case, sd, xr), tgt).backward()
    model = SwinIR(drop_path_rate=0.0, **case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().set_precision(precision)
    for p in model.parameters():
        p.requires_grad_(False)
    xm = x.clone().cuda().requires_grad_(True)
    gc.loss_fn(case)(model(xm), tgt.cuda()).backward()
    assert _rel_l2(xm.grad.cpu(), xr.grad) < tol, _rel_l2(xm.grad.cpu(), xr.grad)
    for p in model.parameters():
        p.requires_grad_(True)
    x2 = x.flip(3).contiguous()
    sdg = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point) for k, v in sd.items()}
    (gc.oracle_forward(case, sdg, x).square().mean() + gc.oracle_forward(case, sdg, x2).abs().mean()).backward()
    y1, y2 = model(x.cuda()), model(x2.cuda())
    (y1.square().mean() + y2.abs().mean()).backward()
    worst = max(_rel_l2(p.grad.cpu(), sdg[k].grad) for k, p in model.named_parameters())
    assert worst < (6e-2 if precision == "bf16" else 1e-3), worst


def test_launch_list_is_one_ffi_call_and_reports_failures():
    """kair_run_list replays a plan's pass in one call; a failing entry is reported with its index and name"""
    import ctypes
    from kair_b200 import _lib, ops
    lib = _lib.load()
    assert lib.kair_op_id(b"kair_gemm_tn") >= 0 and lib.kair_op_id(b"kair_last_error") == -1
    y, xx = torch.ones(1024, device="cuda"), torch.full((1024,), 2.0, device="cuda")
    lst = ops.make_list([ops.build_add_inplace(y, xx), ops.build_add_inplace(y, xx)])
    lst.run(ops.stream_handle())
    torch.cuda.synchronize()
    assert torch.equal(y, torch.full((1024,), 5.0, device="cuda"))
    bad = ops.build_add_inplace(y, xx)
    bad.args.n = 3          # not a multiple of 4: the launcher refuses
    with pytest.raises(RuntimeError, match="launch 1"):
        ops.make_list([ops.build_add_inplace(y, xx), bad]).run(ops.stream_handle())
