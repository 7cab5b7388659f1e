"""Tiled inference for large images: the arithmetic of main_test_swinir.py:68-74 (flip padding to a window multiple)
and :256-284 (`test`: overlapping tiles, uniform overlap-add E / W) of the reference, batched and shardable.

Differences in execution, not in arithmetic: tiles are stacked `tile_batch` at a time into one forward (every tile
has the same shape, so one engine plan serves them all), the overlap-add runs on the device, and a `(rank, world)`
pair selects a contiguous-by-index subset of the tiles (`tile_index % world == rank`) so that N GPUs process N
disjoint subsets with NO collective on the data path: each rank returns its partial (E, W) sums, and
`merge_partials` adds them and divides (tile-sharded inference).
"""
import torch


def pad_to_window(img, window_size):
    """main_test_swinir.py:68-72: mirror-extend H and W to the next multiple of window_size (always pads >= 1 row)."""
    _, _, h, w = img.shape
    hp = (h // window_size + 1) * window_size - h
    wp = (w // window_size + 1) * window_size - w
    img = torch.cat([img, torch.flip(img, [2])], 2)[:, :, :h + hp, :]
    return torch.cat([img, torch.flip(img, [3])], 3)[:, :, :, :w + wp]


def tile_origins(h, w, tile, overlap):
    """main_test_swinir.py:266-268: top-left corners of the tiles, the last one flush with the border."""
    stride = tile - overlap
    hs = list(range(0, h - tile, stride)) + [h - tile]
    ws = list(range(0, w - tile, stride)) + [w - tile]
    return [(y, x) for y in hs for x in ws]


def tiled_partial(model, img, tile, overlap, scale, window_size, tile_batch=4, rank=0, world=1):
    """Partial overlap-add sums (E, W, n_tiles) of this rank's tiles.  img: (1, C, h, w) already window-padded."""
    b, c, h, w = img.shape
    if b != 1:
        raise ValueError("tiled inference takes one image at a time (the tiles form the batch)")
    tile = min(tile, h, w)
    assert tile % window_size == 0, "tile size should be a multiple of window_size"
    origins = [o for i, o in enumerate(tile_origins(h, w, tile, overlap)) if i % world == rank]
    E = torch.zeros(1, c, h * scale, w * scale, dtype=img.dtype, device=img.device)
    W = torch.zeros(1, 1, h * scale, w * scale, dtype=img.dtype, device=img.device)
    ts = tile * scale
    for i in range(0, len(origins), tile_batch):
        chunk = origins[i:i + tile_batch]
        patches = torch.cat([img[..., y:y + tile, x:x + tile] for y, x in chunk], 0)
        out = model(patches)
        for j, (y, x) in enumerate(chunk):
            E[..., y * scale:y * scale + ts, x * scale:x * scale + ts].add_(out[j:j + 1])
            W[..., y * scale:y * scale + ts, x * scale:x * scale + ts].add_(1.0)
    return E, W, len(origins)


def merge_partials(parts):
    """Sum the per-rank (E, W) pairs and divide (main_test_swinir.py:282)."""
    E = sum(p[0] for p in parts)
    W = sum(p[1] for p in parts)
    return E / W


def tiled_forward(model, img, tile, overlap, scale, window_size, tile_batch=4):
    """Single-process equivalent of main_test_swinir.test(img_lq, model, args, window_size) with args.tile set."""
    E, W, _ = tiled_partial(model, img, tile, overlap, scale, window_size, tile_batch)
    return E / W


# ------------------------------------------------------------------------------------------------ x8 self-ensemble

def dihedral(img, mode):
    """The 8 flips / rotations of utils_image.augment_img_tensor4 (utils/utils_image.py:408-426) on (B, C, H, W)."""
    if mode == 0:
        return img
    k, flip = {1: (1, True), 2: (0, True), 3: (3, False), 4: (2, True), 5: (1, False), 6: (2, False), 7: (3, True)}[mode]
    out = img.rot90(k, [2, 3]) if k else img
    return out.flip([2]) if flip else out


def dihedral_inverse_mode(mode):
    """utils_model.test_x8 (:178-182): modes 3 and 5 (pure quarter turns) are each other's inverse, the rest are
    involutions."""
    return 8 - mode if mode in (3, 5) else mode


def test_x8(model, L, sf=1):
    """`utils_model.test_x8(model, L, modulo=1, sf)` (utils/utils_model.py:176-185): mean over the 8 dihedral transforms of
    model(transform(L)) mapped back.  Execution differs, not arithmetic: transforms of equal shape are stacked along the
    batch into one forward (all 8 for square inputs, 2 x 4 otherwise), so one engine plan serves them."""
    outs = [None] * 8
    by_shape = {}
    for i in range(8):
        by_shape.setdefault(tuple(dihedral(L, i).shape), []).append(i)
    b = L.shape[0]
    for modes in by_shape.values():
        y = model(torch.cat([dihedral(L, i) for i in modes], 0).contiguous())
        for j, i in enumerate(modes):
            outs[i] = dihedral(y[j * b:(j + 1) * b], dihedral_inverse_mode(i))
    return torch.stack(outs, dim=0).mean(dim=0, keepdim=False)
