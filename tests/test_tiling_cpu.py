"""Host logic of kair_b200.tiling against a literal restatement of main_test_swinir.py:68-74, 256-284 (flip padding,
tile origins, uniform overlap-add), with a stand-in model so that no GPU is needed; plus rank sharding: the partial
sums of 1, 2 and 3 ranks merge to the same image (no collective on the data path)."""
import pytest
import torch

import emu_ops
import ref_loader
from kair_b200 import tiling


def _model(x):
    """stand-in x2 'network' with a position-dependent response so that tile placement errors show"""
    y = torch.nn.functional.interpolate(x, scale_factor=2, mode="nearest")
    return y * 0.5 + y.mean(dim=(2, 3), keepdim=True)


def _reference_test(img_lq, model, tile, tile_overlap, sf, window_size):
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


def test_pad_and_tiles_match_reference_loop():
    g = torch.Generator().manual_seed(0)
    img = torch.rand(1, 3, 45, 61, generator=g)
    ws = 8
    h_old, w_old = img.shape[2:]
    h_pad, w_pad = (h_old // ws + 1) * ws - h_old, (w_old // ws + 1) * ws - w_old
    ref = torch.cat([img, torch.flip(img, [2])], 2)[:, :, :h_old + h_pad, :]
    ref = torch.cat([ref, torch.flip(ref, [3])], 3)[:, :, :, :w_old + w_pad]
    padded = tiling.pad_to_window(img, ws)
    assert torch.equal(padded, ref) and padded.shape[2] % ws == 0 and padded.shape[3] % ws == 0
    want = _reference_test(padded, _model, 24, 8, 2, ws)
    for tb in (1, 3, 64):
        got = tiling.tiled_forward(_model, padded, 24, 8, 2, ws, tile_batch=tb, ops=emu_ops)
        assert torch.allclose(got, want, atol=1e-6)


def test_cfg5_geometry():
    """2048^2 -> flip-padded 2056^2 -> tile 256 / overlap 32 -> 10 x 10 tiles."""
    o = tiling.tile_origins(2056, 2056, 256, 32)
    assert len(o) == 100 and o[0] == (0, 0) and o[-1] == (1800, 1800)
    counts = [len([i for i in range(100) if i % 8 == r]) for r in range(8)]
    assert max(counts) == 13 and sum(counts) == 100


def test_rank_sharding_merges_to_the_same_image():
    g = torch.Generator().manual_seed(1)
    padded = tiling.pad_to_window(torch.rand(1, 1, 50, 70, generator=g), 8)
    want = tiling.tiled_forward(_model, padded, 16, 8, 2, 8, ops=emu_ops)
    for world in (2, 3):
        parts = [tiling.tiled_partial(_model, padded, 16, 8, 2, 8, tile_batch=2, rank=r, world=world, ops=emu_ops) for r in range(world)]
        assert sum(p[2] for p in parts) == len(tiling.tile_origins(padded.shape[2], padded.shape[3], 16, 8))
        assert torch.allclose(tiling.merge_partials(parts), want, atol=1e-6)


def test_x8_ensemble_matches_the_sequential_definition():
    """kair_b200.tiling.test_x8 vs utils_model.test_x8's definition evaluated one transform at a time (square and
    non-square inputs; an un-equivariant conv so that every inverse transform matters)."""
    import torch
    from kair_b200 import tiling
    torch.manual_seed(0)
    conv = torch.nn.Conv2d(2, 2, 3, padding=1)
    calls = []

    def model(x):
        calls.append(tuple(x.shape))
        return conv(x)

    for shape in [(2, 2, 12, 12), (1, 2, 8, 14)]:
        L = torch.rand(shape)
        for i in range(8):   # every transform is undone by the mode test_x8 pairs it with
            assert torch.equal(tiling.dihedral(tiling.dihedral(L, i), tiling.dihedral_inverse_mode(i)), L)
        with torch.no_grad():
            calls.clear()
            got = tiling.test_x8(model, L, ops=emu_ops)
            n_fwd = len(calls)
            ref = torch.stack([tiling.dihedral(conv(tiling.dihedral(L, i)), tiling.dihedral_inverse_mode(i))
                               for i in range(8)], 0).mean(0)
        assert n_fwd == (1 if shape[2] == shape[3] else 2)
        assert torch.allclose(got, ref, atol=1e-6)
    ident = tiling.test_x8(lambda x: x, torch.arange(24.0).view(1, 1, 4, 6), ops=emu_ops)
    assert torch.allclose(ident, torch.arange(24.0).view(1, 1, 4, 6))


@pytest.mark.skipif(not ref_loader.have_reference(), reason="reference tree not mounted")
@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("shape,sf", [((1, 2, 70, 90), 1), ((2, 1, 64, 64), 2), ((1, 3, 45, 130), 1)])
def test_test_mode_matches_the_reference_utils_model(mode, shape, sf):
    """kair_b200.tiling.test_mode vs the UNMODIFIED utils/utils_model.test_mode (pad, recursive split, x8, split + x8)
    around the same stand-in network (position dependent, not equivariant, so every index matters)"""
    import importlib
    ref_loader.load_utils()
    um = importlib.import_module("utils.utils_model")
    torch.manual_seed(0)
    conv = torch.nn.Conv2d(shape[1], shape[1], 3, padding=1)

    def model(x):
        y = conv(x)
        if sf > 1:
            y = torch.nn.functional.interpolate(y, scale_factor=sf, mode="nearest")
        ramp = torch.linspace(0, 1, y.shape[-1])[None, None, None, :] * torch.linspace(1, 2, y.shape[-2])[None, None, :, None]
        return y * (1 + 0.1 * ramp)

    L = torch.rand(shape, generator=torch.Generator().manual_seed(5))
    with torch.no_grad():
        want = um.test_mode(model, L, mode=mode, refield=16, min_size=40, sf=sf, modulo=8)
        got = tiling.test_mode(model, L, mode=mode, refield=16, min_size=40, sf=sf, modulo=8, ops=emu_ops)
    assert got.shape == want.shape
    assert torch.allclose(got, want, atol=1e-6), (got - want).abs().max()
