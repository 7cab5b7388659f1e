"""GPU parity of the conv-net path (through the C ABI): the new helper kernels vs their PyTorch statements in
tests/emu_ops.py, and the drop-in DnCNN / UNetRes modules vs the CPU oracle, the committed golden fixtures (made from
the real reference modules) and the model_zoo/dncnn_25.pth known-answer fixture.
Tolerances: bf16 tensor-core mode, 2e-2 max-abs on [0,1]-range outputs (north_star); gradients in relative L2."""
import json
import os

import numpy as np
import pytest
import torch

import emu_ops
import golden_cases as gc
from test_elementwise_gpu import _g, _run_both

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ------------------------------------------------------------------------------------------------ kernels

@pytest.mark.parametrize("B,H,W,C", [(2, 8, 12, 64), (1, 16, 16, 256), (3, 4, 4, 512)])
def test_regroup(B, H, W, C):
    g = _g(C + H)
    fine = torch.randn(B * H * W, C, generator=g).bfloat16()
    kw = dict(fine=fine, coarse=torch.zeros(B * H * W // 4, 4 * C).bfloat16(), batch=B, H=H, W=W, C=C)
    _run_both("space_to_depth", kw, ["coarse"], 1e-6)
    for src in (torch.randn(B * H * W // 4, 4 * C, generator=g), kw["coarse"]):
        kw2 = dict(coarse=src, batch=B, H=H, W=W, C=C, out_f32=torch.zeros(B * H * W, C),
                   out_bf16=torch.zeros(B * H * W, C).bfloat16())
        _run_both("depth_to_space", kw2, ["out_f32"], 1e-6)
        _run_both("depth_to_space", kw2, ["out_bf16"], 1e-2)
    # round trip
    from kair_b200 import ops
    f = fine.cuda(); c = torch.zeros(B * H * W // 4, 4 * C, device="cuda", dtype=torch.bfloat16)
    back = torch.zeros_like(f)
    ops.space_to_depth(f, c, batch=B, H=H, W=W, C=C)
    ops.depth_to_space(c, batch=B, H=H, W=W, C=C, out_bf16=back)
    assert torch.equal(back, f)


@pytest.mark.parametrize("T,C,dt", [(1000, 64, torch.bfloat16), (4096, 512, torch.bfloat16), (777, 128, torch.float32)])
def test_colsum(T, C, dt):
    x = torch.randn(T, C, generator=_g(T)).to(dt)
    _run_both("colsum", dict(x=x, out=torch.zeros(C), C=C, scale=0.5), ["out"], 1e-4)


@pytest.mark.parametrize("T,C,slope", [(2 * 24 * 24, 64, 0.0), (1000, 128, 0.2)])
def test_batchnorm_kernels(T, C, slope):
    g = _g(T + C)
    z = torch.randn(T, C, generator=g) * 1.7 + 0.3 * torch.randn(C, generator=g)
    gamma, beta = 1 + 0.2 * torch.randn(C, generator=g), 0.1 * torch.randn(C, generator=g)
    rm, rv = 0.1 * torch.randn(C, generator=g), 0.5 + torch.rand(C, generator=g)
    acc = torch.zeros(2, C, dtype=torch.float64)
    _run_both("bn_stats", dict(z=z, acc=acc, C=C), ["acc"], 1e-5)
    mean, rstd = torch.zeros(C), torch.zeros(C)
    nbt = torch.zeros((), dtype=torch.int64)
    kw = dict(acc=acc, mean=mean, rstd=rstd, running_mean=rm, running_var=rv, num_batches_tracked=nbt, C=C, T=T,
              eps=1e-4, momentum=0.9)
    _run_both("bn_finalize", kw, ["mean", "rstd", "running_mean", "running_var", "num_batches_tracked", "acc"], 1e-5)
    assert int(nbt) == 1
    y = torch.zeros(T, C).bfloat16()
    kwa = dict(z=z, mean=mean, rstd=rstd, gamma=gamma, beta=beta, C=C, act=emu_ops.ACT_LRELU, act_slope=slope,
               out_bf16=y, out_f32=torch.zeros(T, C))
    _run_both("bn_apply", kwa, ["out_f32"], 1e-5)
    _run_both("bn_apply", kwa, ["out_bf16"], 1e-2)
    for dy in (torch.randn(T, C, generator=g), torch.randn(T, C, generator=g).bfloat16()):
        dgamma, dbeta = torch.zeros(C), torch.zeros(C)
        kwr = dict(z=z, y=y, dy=dy, mean=mean, rstd=rstd, dgamma=dgamma, dbeta=dbeta, C=C, act=emu_ops.ACT_LRELU,
                   act_slope=slope)
        _run_both("bn_bwd_reduce", kwr, ["dgamma", "dbeta"], 1e-4)
        kwb = dict(z=z, y=y, dy=dy, mean=mean, rstd=rstd, gamma=gamma, dgamma=dgamma, dbeta=dbeta,
                   dz=torch.zeros(T, C).bfloat16(), C=C, act=emu_ops.ACT_LRELU, act_slope=slope, dbias=torch.zeros(C))
        _run_both("bn_bwd_apply", kwb, ["dz"], 1e-2)
    sc, sh = torch.zeros(C), torch.zeros(C)
    _run_both("bn_fold", dict(gamma=gamma, beta=beta, running_mean=rm, running_var=rv, scale=sc, shift=sh, C=C,
                              eps=1e-4), ["scale", "shift"], 1e-5)


@pytest.mark.parametrize("r,C", [(2, 64), (2, 6), (3, 64), (4, 4)])
def test_pixel_shuffle_and_nearest(r, C):
    B, H, W = 2, 6, 10
    x = torch.randn(B * H * W, C * r * r, generator=_g(r)).bfloat16()
    _run_both("pixel_shuffle", dict(x=x, out=torch.zeros(B * H * r * W * r, C).bfloat16(), batch=B, H=H, W=W, C=C, r=r),
              ["out"], 1e-6)
    # the emulation follows torch.nn.PixelShuffle's channel order
    ref = torch.nn.functional.pixel_shuffle(x.float().view(B, H, W, -1).permute(0, 3, 1, 2), r)
    out = torch.zeros(B * H * r * W * r, C).bfloat16()
    emu_ops.build_pixel_shuffle(x, out, batch=B, H=H, W=W, C=C, r=r).run()
    assert torch.equal(out.float().view(B, H * r, W * r, C).permute(0, 3, 1, 2), ref)
    dy = torch.randn(B * H * r * W * r, C, generator=_g(r + 10)).bfloat16()
    _run_both("pixel_shuffle_bwd", dict(dy=dy, dx=torch.zeros(B * H * W, C * r * r).bfloat16(), batch=B, H=H, W=W, C=C, r=r),
              ["dx"], 1e-6)
    if r == 2:
        x8 = torch.randn(B * H * W, 64, generator=_g(5)).bfloat16()
        _run_both("upsample_nearest2", dict(x=x8, out=torch.zeros(B * 4 * H * W, 64).bfloat16(), batch=B, H=H, W=W, C=64),
                  ["out"], 1e-6)


@pytest.mark.parametrize("conv,N,K", [(True, 64, 64), (True, 128, 128), (False, 256, 128), (True, 512, 512)])
def test_gemm_tn_activation_backward_mask(conv, N, K):
    """LINEAR epilogue with aux: (acc + bias) * (aux > 0 ? 1 : slope) [+ resid] -- the ReLU backward fused into dgrad."""
    B, H, W = 2, 12, 20
    M = B * H * W
    g = _g(N + K)
    A = torch.randn(M, K, generator=g).bfloat16()
    taps = 9 if conv else 1
    Bw = (torch.randn(N, taps * K, generator=g) * 0.05).bfloat16()
    aux = torch.randn(M, N, generator=g).bfloat16()
    resid = torch.randn(M, N, generator=g)
    bn = 256 if N % 256 == 0 else (128 if N % 128 == 0 else 64)
    for use_resid, slope in ((False, 0.0), (True, 0.2)):
        kw = dict(A=A, B=Bw, N=N, K=K, bn=bn, aux=aux, act_slope=slope, resid=resid if use_resid else None,
                  outB=torch.zeros(M, N).bfloat16(), outF=torch.zeros(M, N), conv=(B, H, W) if conv else None)
        _run_both("gemm_tn", kw, ["outF"], 2e-3)
        _run_both("gemm_tn", kw, ["outB"], 1e-2)


def test_conv_out_dgrad_mask_and_pack_conv_fold():
    B, co, H, W, C = 2, 1, 10, 14, 64
    g = _g(3)
    dout = torch.randn(B, co, H, W, generator=g)
    Wt = torch.randn(co, C, 3, 3, generator=g) * 0.1
    mask = torch.randn(B * H * W, C, generator=g).bfloat16()
    kw = dict(dout=dout, W=Wt, C=C, pitch=C, out_scale=-1.0, dfeat_f32=torch.zeros(B * H * W, C), mask=mask,
              mask_slope=0.1)
    _run_both("conv_out_dgrad", kw, ["dfeat_f32"], 1e-5)
    L = 3
    Ws = [torch.randn(C, C, 3, 3, generator=g) * 0.05 for _ in range(L)]
    bs = [torch.randn(C, generator=g) for _ in range(L)]
    osc = [1 + 0.3 * torch.randn(C, generator=g) for _ in range(L)]
    osh = [torch.randn(C, generator=g) for _ in range(L)]
    kwp = dict(W=Ws, b=bs, Wp=torch.zeros(L, C, 9 * C).bfloat16(), WpT=torch.zeros(L, C, 9 * C).bfloat16(),
               bp=torch.zeros(L, C), O=C, I=C, I_pad=C, oscale=osc, oshift=osh)
    _run_both("pack_conv", kwp, ["bp"], 1e-6)
    _run_both("pack_conv", kwp, ["Wp", "WpT"], 1e-2)


# ------------------------------------------------------------------------------------------------ networks

def _rel_l2(a, b):
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


def _build(case):
    from kair_b200.network_dncnn import DnCNN, FDnCNN
    from kair_b200.network_unet import UNetRes
    kw = case["kw"]
    model = {"dncnn": DnCNN, "fdncnn": FDnCNN, "unetres": UNetRes}[case["net"]](**kw)
    sd, x, tgt = gc.state_and_io(case)
    model.load_state_dict(sd, strict=True)
    return model, sd, x, tgt


def _check_against_golden(name, grad_tol, out_tol=2e-2):
    case = gc.CASES[name]
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    model, sd, x, tgt = _build(case)
    model = model.cuda().train(case.get("train_mode", False))
    y = model(x.cuda())
    loss = gc.loss_fn(case)(y, tgt.cuda())
    loss.backward()
    torch.cuda.synchronize()
    yr = z["y"]
    assert np.abs(y.detach().cpu().numpy() - yr).max() < out_tol * max(1.0, np.abs(yr).max())
    grads = json.loads(str(z["grads"]))
    gmax = max(v[0] for v in grads.values())
    bad, e2, r2 = [], 0.0, 0.0
    for k, p in model.named_parameters():
        n = p.grad.double().norm().item()
        e2, r2 = e2 + (n - grads[k][0]) ** 2, r2 + grads[k][0] ** 2
        if abs(n - grads[k][0]) > 3 * grad_tol * grads[k][0] + 2e-3 * gmax:
            bad.append((k, n, grads[k][0]))
    assert not bad, bad[:6]
    assert (e2 / r2) ** 0.5 < grad_tol
    return model, z


def test_golden_dncnn_r17():
    _check_against_golden("dncnn_r17", 4e-2)


def test_golden_fdncnn():
    """FDnCNN (network_dncnn.py:126-148): 2 input channels (image + noise-level map), 1 output, no residual."""
    _check_against_golden("fdncnn_r8", 4e-2)


def test_golden_dncnn_br_train():
    model, z = _check_against_golden("dncnn_br_train", 0.1)
    sd = model.state_dict()
    for k in z.files:
        if k.startswith("stat:"):
            assert np.allclose(sd[k[5:]].cpu().numpy(), z[k], rtol=2e-2, atol=1e-2), k
    assert int(sd["model.3.num_batches_tracked"]) == 1


def test_golden_dncnn_br_eval_forward():
    """eval-mode BN: running statistics folded into the packed conv weights (no_grad path)."""
    case = gc.CASES["dncnn_br_eval"]
    z = np.load(os.path.join(GOLDEN, "dncnn_br_eval.npz"))
    model, sd, x, tgt = _build(case)
    model = model.cuda().eval()
    with torch.no_grad():
        y = model(x.cuda())
    assert np.abs(y.cpu().numpy() - z["y"]).max() < 2e-2 * max(1.0, np.abs(z["y"]).max())
    with pytest.raises(NotImplementedError, match="eval-mode BatchNorm"):
        model(x.cuda())


def test_golden_unetres_nb2():
    _check_against_golden("unetres_nb2", 6e-2)


@pytest.mark.parametrize("net,shape", [("dncnn", (2, 1, 33, 47)), ("dncnn_bn", (3, 3, 24, 40)), ("unet", (2, 4, 40, 24))])
def test_networks_vs_oracle_all_gradients(net, shape):
    """Every parameter gradient (full tensors, not fingerprints) vs the oracle on odd / non-square sizes."""
    from kair_b200.network_dncnn import DnCNN
    from kair_b200.network_unet import UNetRes
    from oracle import convnets_oracle as co
    from test_convnet_engine_cpu import _check_grads, _sd, _seed_params
    torch.manual_seed(0)
    B, c, H, W = shape
    if net == "unet":
        model = UNetRes(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=2, act_mode="R", bias=False)
        _seed_params(model, 61, scale=0.7)
        fwd = lambda sd, x: co.unetres_forward(sd, x, nb=2)
        co_ = 3
    else:
        act = "BR" if net == "dncnn_bn" else "R"
        model = DnCNN(in_nc=c, out_nc=c, nc=64, nb=6, act_mode=act)
        _seed_params(model, 62)
        model.train(True)
        fwd = lambda sd, x: co.dncnn_forward(sd, x, nb=6, act_mode=act, training=True)
        co_ = c
    sd = _sd(model)
    g = torch.Generator().manual_seed(8)
    x, tgt = torch.rand(B, c, H, W, generator=g), torch.rand(B, co_, H, W, generator=g)
    y_ref = fwd(sd, x)
    torch.mean((y_ref - tgt) ** 2).backward()
    model = model.cuda()
    y = model(x.cuda())
    torch.mean((y - tgt.cuda()) ** 2).backward()
    torch.cuda.synchronize()
    assert (y.detach().cpu() - y_ref.detach()).abs().max().item() < 2e-2 * max(1.0, y_ref.detach().abs().max().item())
    _check_grads(model, sd, {k: p.grad.cpu() for k, p in model.named_parameters()}, 0.1 if net == "dncnn_bn" else 6e-2)


def test_dncnn25_known_answer_cfg1():
    """BASELINE configs[0]: DnCNN-17 'R' + model_zoo/dncnn_25.pth on the seeded 256x256 image; y from the real reference."""
    from kair_b200.network_dncnn import DnCNN
    z = np.load(os.path.join(GOLDEN, "kat", "dncnn_25.npz"))
    sd = {str(k): torch.from_numpy(z["w:" + str(k)]) for k in z["keys"]}
    model = DnCNN(in_nc=1, out_nc=1, nc=64, nb=17, act_mode="R")
    model.load_state_dict(sd, strict=True)
    model = model.cuda().eval()
    for p in model.parameters():
        p.requires_grad = False            # main_test_dncnn.py:123-124
    torch.manual_seed(0)
    x = torch.rand(1, 1, 256, 256)
    y = model(x.cuda()).cpu()
    yr = torch.from_numpy(z["y"])
    err = (y - yr).abs().max().item()
    assert err < 2e-2, err
    # cfg1 denoising protocol: sigma=25 noise on a clean synthetic image, PSNR (uint8) within 0.02 dB of the reference
    clean, noisy = gc.dncnn_cfg1_images()
    y_dn = model(noisy.cuda()).cpu()
    assert (y_dn - torch.from_numpy(z["y_denoise"])).abs().max().item() < 2e-2
    assert abs(gc.psnr_uint8(noisy, clean) - float(z["psnr_in"])) < 1e-6
    assert abs(gc.psnr_uint8(y_dn, clean) - float(z["psnr_out"])) < 0.02, (gc.psnr_uint8(y_dn, clean), float(z["psnr_out"]))
    assert abs(y.mean().item() - 0.5027428) < 2e-3
    # batch of 5 odd-sized images: same weights, per-image equality with the batch-1 result of the same image
    xb = torch.rand(5, 1, 63, 97, generator=torch.Generator().manual_seed(3)).cuda()
    yb = model(xb)
    y1 = model(xb[2:3])
    assert torch.equal(yb[2:3], y1)


def test_full_size_properties_drunet_cfg4():
    """cfg4 shape (B=16, 4x128x128, DRUNet nb=4): finite outputs and gradients, linearity of the bias-free net in its
    last layer (tail weights zeroed => exact zeros), determinism of two identical steps."""
    from kair_b200.network_unet import UNetRes
    torch.manual_seed(0)
    model = UNetRes(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", downsample_mode="strideconv",
                    upsample_mode="convtranspose", bias=False).cuda().train()
    x = torch.rand(16, 4, 128, 128, device="cuda")
    y1 = model(x)
    y1.abs().mean().backward()
    g1 = [p.grad.clone() for p in model.parameters()]
    assert y1.shape == (16, 3, 128, 128) and torch.isfinite(y1).all() and all(torch.isfinite(g).all() for g in g1)
    model.zero_grad()
    y2 = model(x)
    y2.abs().mean().backward()
    assert torch.equal(y1, y2) or (y1 - y2).abs().max().item() < 1e-6
    with torch.no_grad():
        model.m_tail.weight.zero_()
        assert torch.count_nonzero(model(x)).item() == 0


def test_cpu_tensor_is_rejected():
    from kair_b200.network_dncnn import DnCNN
    from kair_b200.network_unet import UNetRes
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        DnCNN()(torch.rand(1, 1, 16, 16))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        UNetRes()(torch.rand(1, 3, 16, 16))


@pytest.mark.parametrize("net", ["dncnn", "swinir"])
def test_cuda_graph_replay_matches_eager(net):
    """kair_b200.graphs.GraphedInference: one captured forward replayed on new inputs and after a weight change."""
    from kair_b200.graphs import GraphedInference
    torch.manual_seed(0)
    if net == "dncnn":
        from kair_b200.network_dncnn import DnCNN
        model = DnCNN(in_nc=1, out_nc=1, nc=64, nb=17, act_mode="R").cuda().eval()
        shape = (1, 1, 64, 64)
    else:
        from kair_b200.network_swinir import SwinIR
        model = SwinIR(upscale=2, in_chans=3, img_size=32, window_size=8, img_range=1.0, depths=[2, 2], embed_dim=60,
                       num_heads=[6, 6], mlp_ratio=2, upsampler="pixelshuffledirect").cuda().eval()
        shape = (2, 3, 32, 32)
    g = torch.Generator().manual_seed(1)
    x0, x1 = torch.rand(shape, generator=g).cuda(), torch.rand(shape, generator=g).cuda()
    graphed = GraphedInference(model, x0)
    with torch.no_grad():
        assert torch.equal(graphed(x1).clone(), model(x1))
        assert torch.equal(graphed(x0).clone(), model(x0))
        for p in model.parameters():
            p.mul_(1.01)                       # weights change: the packed operands must follow before the replay
        y_new = model(x1)
        assert torch.equal(graphed(x1).clone(), y_new)
    with pytest.raises(RuntimeError, match="captured for shape"):
        graphed(torch.rand(1, shape[1], 16, 16).cuda())


def test_x8_ensemble_on_the_gpu_path():
    """tiling.test_x8 (utils_model.test_x8, :176-185) around the drop-in DnCNN on the device kernels
    (kair_tiles_gather / kair_tiles_accumulate): (i) against the ORACLE network evaluated one transform at a time with the
    reference's definition (bf16 tolerance), (ii) against the one-at-a-time evaluation of the CUDA path itself (only the
    order of the eight additions differs: 1e-6)."""
    import golden_cases as gc
    from kair_b200 import tiling
    from kair_b200.network_dncnn import DnCNN
    from oracle import convnets_oracle as co
    case = gc.CASES["dncnn_r17"]
    sd, _, _ = gc.state_and_io(case)
    model = DnCNN(**case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().eval()
    for shape in [(1, 1, 48, 48), (2, 1, 32, 56)]:
        L = torch.rand(shape, generator=torch.Generator().manual_seed(3))
        with torch.no_grad():
            got = tiling.test_x8(model, L.cuda())
            seq = torch.stack([tiling.dihedral(model(tiling.dihedral(L.cuda(), i).contiguous()), tiling.dihedral_inverse_mode(i))
                               for i in range(8)], 0).mean(0)
            want = torch.stack([tiling.dihedral(co.dncnn_forward(sd, tiling.dihedral(L, i).contiguous(), nb=17, act_mode="R"),
                                                tiling.dihedral_inverse_mode(i)) for i in range(8)], 0).mean(0)
        assert got.shape == want.shape
        assert (got - seq).abs().max().item() < 1e-6
        assert (got.cpu() - want).abs().max().item() < 2e-2


def test_tile_kernels_match_their_statement():
    """kair_tiles_gather / kair_tiles_accumulate vs tests/emu_ops.py: every dihedral mode, non-square tiles, overlaps"""
    import emu_ops
    from test_elementwise_gpu import _run_both
    g = torch.Generator().manual_seed(9)
    img = torch.rand(3, 40, 52, generator=g)
    rows = [(m % 3, 2 * m, m) for m in range(8)]
    for th, tw in ((16, 24), (20, 20)):
        kw = dict(img=img, origins=torch.tensor(rows, dtype=torch.int32), tiles=torch.zeros(8, 3, th, tw))
        _run_both("tiles_gather", kw, ["tiles"], 1e-12)
        tiles = torch.rand(8, 3, th, tw, generator=g)
        kw2 = dict(tiles=tiles, origins=torch.tensor(rows, dtype=torch.int32), E=torch.rand(3, 60, 60, generator=g),
                   Wsum=torch.zeros(60, 60), weight=0.125)
        _run_both("tiles_accumulate", kw2, ["E", "Wsum"], 1e-6)


def test_tiled_forward_and_split_on_the_gpu_path():
    """main_test_swinir.test (tiles + overlap-add) and utils_model.test_split_fn around the drop-in DnCNN vs the literal
    reference loops on the oracle network"""
    import golden_cases as gc
    import parity_full as pf
    from kair_b200 import tiling
    from kair_b200.network_dncnn import DnCNN
    from oracle import convnets_oracle as co
    case = gc.CASES["dncnn_r17"]
    sd, _, _ = gc.state_and_io(case)
    model = DnCNN(**case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().eval()
    oracle = lambda p: co.dncnn_forward(sd, p, nb=17, act_mode="R")
    img = tiling.pad_to_window(torch.rand(1, 1, 70, 90, generator=torch.Generator().manual_seed(4)), 8)
    with torch.no_grad():
        got = tiling.tiled_forward(model, img.cuda(), 32, 8, 1, 8, tile_batch=3).cpu()
        want = pf.reference_tiled_loop(img, oracle, 32, 8, 1)
        assert (got - want).abs().max().item() < 2e-2
        L = torch.rand(1, 1, 72, 88, generator=torch.Generator().manual_seed(5))
        got2 = tiling.test_mode(model, L.cuda(), mode=2, refield=16, min_size=40, sf=1, modulo=8).cpu()
        want2 = tiling.test_split_fn(oracle, L, refield=16, min_size=40, sf=1, modulo=8)   # host logic pinned on CPU
        assert (got2 - want2).abs().max().item() < 2e-2


def test_rrdbnet_noup_forward_backward_vs_oracle_and_golden():
    """RRDBNet_noUp (the fork's `rrdbnet_no_up`) on the CUDA path: golden output of the real reference module, every
    parameter gradient and the input gradient vs autograd through the oracle; eval plan == training plan forward."""
    import os
    import numpy as np
    import golden_cases as gc
    from kair_b200.network_rrdbnet import RRDBNet_noUp
    case = gc.CASES["rrdbnet_nb2"]
    sd, x, tgt = gc.state_and_io(case)
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rrdbnet_nb2.npz"))
    sdg = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    xr = x.clone().requires_grad_(True)
    gc.loss_fn(case)(gc.oracle_forward(case, sdg, xr), tgt).backward()
    model = RRDBNet_noUp(**case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().train()
    xm = x.clone().cuda().requires_grad_(True)
    y = model(xm)
    gc.loss_fn(case)(y, tgt.cuda()).backward()
    assert np.abs(y.detach().cpu().numpy() - z["y"]).max() < 2e-2
    bad = [(k, _rel_l2(p.grad.cpu(), sdg[k].grad)) for k, p in model.named_parameters()
           if not _rel_l2(p.grad.cpu(), sdg[k].grad) < 6e-2]
    assert not bad, bad[:6]
    assert _rel_l2(xm.grad.cpu(), xr.grad) < 1.5e-1
    model.eval()
    with torch.no_grad():
        y2 = model(x.cuda())
    assert torch.equal(y2, y.detach())
    # the fork's shape: 2 SPECT views, 64 x 64 patches, batch 4, full depth nb=20 -- finite, and one Adam step moves it
    torch.manual_seed(0)
    big = RRDBNet_noUp(in_nc=2, out_nc=2, nf=64, nb=20, gc=32).cuda().train()
    opt = torch.optim.Adam(big.parameters(), lr=1e-4)
    xb, tb = torch.rand(4, 2, 64, 64, device="cuda"), torch.rand(4, 2, 64, 64, device="cuda")
    l0 = None
    for _ in range(3):
        opt.zero_grad()
        loss = torch.mean(torch.abs(big(xb) - tb))
        loss.backward()
        assert all(torch.isfinite(p.grad).all().item() for p in big.parameters())
        opt.step()
        l0 = l0 if l0 is not None else loss.item()
    assert loss.item() < l0


@pytest.mark.parametrize("d", [2, 3, 4])
def test_dilated_conv_kernels(d):
    """kair_gemm_tn / kair_gemm_nt conv mode with `dilation` vs their statement (F.conv2d with dilation = padding = d)"""
    import emu_ops
    from test_elementwise_gpu import _run_both
    g = torch.Generator().manual_seed(40 + d)
    B, H, W, C = 2, 20, 28, 64
    T = B * H * W
    x = (torch.randn(T, C, generator=g)).bfloat16()
    Wm = (torch.randn(C, 9 * C, generator=g) * 0.05).bfloat16()
    bias = torch.randn(C, generator=g) * 0.1
    kw = dict(A=x, B=Wm, N=C, K=C, bn=64, conv=(B, H, W), bias=bias, act=emu_ops.ACT_LRELU, act_slope=0.0,
              outB=torch.zeros(T, C).bfloat16(), outF=torch.zeros(T, C), dilation=d)
    _run_both("gemm_tn", kw, ["outF"], 2e-3)
    gy = (torch.randn(T, C, generator=g)).bfloat16()
    kw2 = dict(A=gy, B=x, D=torch.zeros(C, 9 * C), n1=C, n2=C, bn=64, conv=(B, H, W), dilation=d)
    _run_both("gemm_nt", kw2, ["D"], 2e-3)


def test_ircnn_forward_backward_vs_oracle_and_golden():
    import os
    import numpy as np
    import golden_cases as gc
    from kair_b200.network_dncnn import IRCNN
    case = gc.CASES["ircnn"]
    sd, x, tgt = gc.state_and_io(case)
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ircnn.npz"))
    sdg = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    gc.loss_fn(case)(gc.oracle_forward(case, sdg, x), tgt).backward()
    model = IRCNN(**case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().train()
    y = model(x.cuda())
    gc.loss_fn(case)(y, tgt.cuda()).backward()
    assert np.abs(y.detach().cpu().numpy() - z["y"]).max() < 2e-2
    bad = [(k, _rel_l2(p.grad.cpu(), sdg[k].grad)) for k, p in model.named_parameters()
           if not _rel_l2(p.grad.cpu(), sdg[k].grad) < 6e-2]
    assert not bad, bad[:6]
    model.eval()
    with torch.no_grad():
        big = model(torch.rand(1, 3, 250, 190, device="cuda"))     # not a tile multiple: TMA out-of-bounds fill at dilation 4
    assert big.shape == (1, 3, 250, 190) and torch.isfinite(big).all()


def test_ffdnet_forward_backward_vs_oracle_and_golden():
    import os
    import numpy as np
    import golden_cases as gc
    from kair_b200.network_ffdnet import FFDNet
    case = gc.CASES["ffdnet_c96"]
    sd, x, tgt = gc.state_and_io(case)
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ffdnet_c96.npz"))
    sdg = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    gc.loss_fn(case)(gc.oracle_forward(case, sdg, x), tgt).backward()
    model = FFDNet(**case["kw"])
    model.load_state_dict(sd, strict=True)
    model = model.cuda().train()
    sig = gc.sigma_of(case).cuda()
    y = model(x.cuda(), sig)
    gc.loss_fn(case)(y, tgt.cuda()).backward()
    assert y.shape == tuple(z["y"].shape) and np.abs(y.detach().cpu().numpy() - z["y"]).max() < 2e-2
    bad = [(k, _rel_l2(p.grad.cpu(), sdg[k].grad)) for k, p in model.named_parameters()
           if not _rel_l2(p.grad.cpu(), sdg[k].grad) < 6e-2]
    assert not bad, bad[:6]
    # the gray model at its published depth (nc=64, nb=15) on an even-sized image: inference, finite, right shape
    torch.manual_seed(0)
    gray = FFDNet(in_nc=1, out_nc=1, nc=64, nb=15, act_mode="R").cuda().eval()
    with torch.no_grad():
        out = gray(torch.rand(2, 1, 240, 240, device="cuda"), torch.full((2, 1, 1, 1), 25 / 255.0, device="cuda"))
    assert out.shape == (2, 1, 240, 240) and torch.isfinite(out).all()
