"""Large-image inference drivers: the arithmetic of main_test_swinir.py:68-74 (flip padding to a window multiple) and
:256-284 (`test`: overlapping tiles, uniform overlap-add E / W) and of utils/utils_model.py:50-200 (`test_mode`: pad,
recursive 4-way split, x8 flip / rotation self-ensemble) of the reference, batched, shardable and with the data movement
on the device.

Differences in execution, not in arithmetic:
  * tiles / transformed copies of equal shape are stacked into one forward (one engine plan serves them all);
  * cutting the tiles, the eight dihedral transforms and their inverses, the overlap-add and the ensemble mean are two
    kernels (`kair_tiles_gather`, `kair_tiles_accumulate`: index arithmetic, one thread per pixel, sums in tile order)
    instead of one eager slice / cat / add_ / rot90 / flip per tile;
  * a `(rank, world)` pair selects `tile_index % world == rank`, so N GPUs process N disjoint subsets with NO collective
    on the data path: each rank returns its partial (E, W) sums and `merge_partials` adds them and divides
    (tile-sharded inference).
`ops` is the kernel binding (default: the CUDA library; tests inject tests/emu_ops.py to check the host logic on CPU).
"""
import numpy as np
import torch

from . import ops as _real_ops


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


def _rows(rows, device):
    """[(y, x, mode)] -> int32 device tensor.  Building it is a pageable host-to-device copy, which makes the host wait
    for everything queued on the stream: loops build ONE table up front and pass slices of it."""
    if isinstance(rows, torch.Tensor):
        return rows
    return torch.tensor(rows, dtype=torch.int32, device=device)


def _gather(o, img3, rows, th, tw):
    """(n, C, th, tw) patches of the (C, H, W) image for rows = [(y, x, mode)] (list, or int32 device tensor [n, 3])"""
    org = _rows(rows, img3.device)
    out = torch.empty(org.shape[0], img3.shape[0], th, tw, dtype=torch.float32, device=img3.device)
    o.build_tiles_gather(img3, org, out).run(o.stream_handle(img3.device))
    return out


def _accumulate(o, tiles, rows, E3, W2, weight=1.0):
    org = _rows(rows, E3.device)
    o.build_tiles_accumulate(tiles.float().contiguous(), org, E3, W2, weight=weight).run(o.stream_handle(E3.device))


def tiled_partial(model, img, tile, overlap, scale, window_size, tile_batch=4, rank=0, world=1, ops=None):
    """Partial overlap-add sums (E, W, n_tiles) of this rank's tiles.  img: (1, C, h, w) already window-padded."""
    o = ops or _real_ops
    b, c, h, w = img.shape
    if b != 1:
        raise ValueError("tiled inference takes one image at a time (the tiles form the batch)")
    tile = min(tile, h, w)
    assert tile % window_size == 0, "tile size should be a multiple of window_size"
    origins = [og for i, og in enumerate(tile_origins(h, w, tile, overlap)) if i % world == rank]
    img3 = img[0].float().contiguous()
    E = torch.zeros(c, h * scale, w * scale, dtype=torch.float32, device=img.device)
    W = torch.zeros(h * scale, w * scale, dtype=torch.float32, device=img.device)
    if not origins:
        return E[None], W[None, None], 0
    # both origin tables go to the device once, before the loop: the host then queues gather -> network -> overlap-add
    # of every chunk without ever waiting for the GPU (a per-chunk table upload drained the stream 25 times per image)
    src = _rows([(y, x, 0) for y, x in origins], img.device)
    dst = _rows([(y * scale, x * scale, 0) for y, x in origins], img.device)
    for i in range(0, len(origins), tile_batch):
        patches = _gather(o, img3, src[i:i + tile_batch], tile, tile)
        out = model(patches)
        _accumulate(o, out, dst[i:i + tile_batch], E, W)
    return E[None], W[None, None], len(origins)


def merge_partials(parts):
    """Sum the per-rank (E, W) pairs and divide (main_test_swinir.py:282)."""
    E = sum(p[0] for p in parts)
    W = sum(p[1] for p in parts)
    return E / W


def tiled_forward(model, img, tile, overlap, scale, window_size, tile_batch=4, ops=None):
    """Single-process equivalent of main_test_swinir.test(img_lq, model, args, window_size) with args.tile set."""
    E, W, _ = tiled_partial(model, img, tile, overlap, scale, window_size, tile_batch, ops=ops)
    return E / W


# ------------------------------------------------------------------------------------------------ utils_model.test_mode

def dihedral(img, mode):
    """The 8 flips / rotations of utils_image.augment_img_tensor4 (utils/utils_image.py:408-426) on (B, C, H, W):
    the eager statement of what the kernels' index arithmetic does (kept for callers that want a single transform)."""
    if mode == 0:
        return img
    k, flip = {1: (1, True), 2: (0, True), 3: (3, False), 4: (2, True), 5: (1, False), 6: (2, False), 7: (3, True)}[mode]
    out = img.rot90(k, [2, 3]) if k else img
    return out.flip([2]) if flip else out


def dihedral_inverse_mode(mode):
    """utils_model.test_x8 (:178-182): modes 3 and 5 (pure quarter turns) are each other's inverse, the rest are
    involutions."""
    return 8 - mode if mode in (3, 5) else mode


def test_pad(model, L, modulo=16, sf=1):
    """utils_model.test_pad (:109-116): replicate-pad bottom / right to a multiple of `modulo`, crop the result."""
    h, w = L.shape[-2:]
    pb, pr = int(np.ceil(h / modulo) * modulo - h), int(np.ceil(w / modulo) * modulo - w)
    if pb or pr:
        L = torch.nn.functional.pad(L, (0, pr, 0, pb), mode="replicate")
    return model(L)[..., :h * sf, :w * sf]


def test_x8(model, L, modulo=1, sf=1, ops=None):
    """`utils_model.test_x8(model, L, modulo, sf)` (:176-185): mean over the 8 dihedral transforms of
    test_pad(model, transform(L)) mapped back.  One gather launch cuts the 8 transformed copies (square inputs: one
    forward of batch 8 B; otherwise two forwards of batch 4 B, one per orientation), one accumulate launch per forward
    applies the inverse transforms and the 1/8 weights (x/8 is exact, so the sum equals the reference's
    stack(...).mean up to the order of the eight additions, which is fixed)."""
    o = ops or _real_ops
    B, C, h, w = L.shape
    img3 = L.float().contiguous().view(B * C, h, w)         # the batch rides along as extra channels of one image
    E = None
    groups = [list(range(8))] if h == w else [[0, 2, 4, 6], [1, 3, 5, 7]]
    for modes in groups:
        th, tw = (w, h) if modes[0] % 2 == 1 else (h, w)     # the quarter-turn modes swap the sides
        x = _gather(o, img3, [(0, 0, m) for m in modes], th, tw).view(len(modes) * B, C, th, tw)
        y = test_pad(model, x, modulo=modulo, sf=sf).float().contiguous()
        Co = y.shape[1]
        if E is None:
            E = torch.zeros(B * Co, h * sf, w * sf, dtype=torch.float32, device=L.device)
        _accumulate(o, y.view(len(modes), B * Co, y.shape[2], y.shape[3]), [(0, 0, m) for m in modes], E, None,
                    weight=0.125)
    return E.view(B, -1, h * sf, w * sf)


def test_split_fn(model, L, refield=32, min_size=256, sf=1, modulo=1):
    """utils_model.test_split_fn (:127-164): recursive 4-way split with `refield`-aligned overlapping quadrants; each
    output quadrant is taken from the network output of its own input quadrant.  The four equally shaped quadrants of a
    level go through the network as one batch."""
    h, w = L.shape[-2:]
    if h * w <= min_size ** 2:
        return test_pad(model, L, modulo=modulo, sf=sf)
    top = slice(0, (h // 2 // refield + 1) * refield)
    bottom = slice(h - (h // 2 // refield + 1) * refield, h)
    left = slice(0, (w // 2 // refield + 1) * refield)
    right = slice(w - (w // 2 // refield + 1) * refield, w)
    Ls = [L[..., top, left], L[..., top, right], L[..., bottom, left], L[..., bottom, right]]
    if h * w <= 4 * (min_size ** 2):
        nb = L.shape[0]
        Eb = model(torch.cat(Ls, 0).contiguous())          # the reference calls model(Ls[i]) four times
        Es = [Eb[i * nb:(i + 1) * nb] for i in range(4)]
    else:
        Es = [test_split_fn(model, Ls[i], refield=refield, min_size=min_size, sf=sf, modulo=modulo) for i in range(4)]
    b, c = Es[0].shape[:2]
    E = torch.zeros(b, c, sf * h, sf * w, dtype=L.dtype, device=L.device)
    E[..., :h // 2 * sf, :w // 2 * sf] = Es[0][..., :h // 2 * sf, :w // 2 * sf]
    E[..., :h // 2 * sf, w // 2 * sf:w * sf] = Es[1][..., :h // 2 * sf, (-w + w // 2) * sf:]
    E[..., h // 2 * sf:h * sf, :w // 2 * sf] = Es[2][..., (-h + h // 2) * sf:, :w // 2 * sf]
    E[..., h // 2 * sf:h * sf, w // 2 * sf:w * sf] = Es[3][..., (-h + h // 2) * sf:, (-w + w // 2) * sf:]
    return E


def test_split_x8(model, L, refield=32, min_size=256, sf=1, modulo=1):
    """utils_model.test_split_x8 (:195-205): test_split_fn on each of the 8 transforms, mapped back, mean."""
    E_list = [test_split_fn(model, dihedral(L, i).contiguous(), refield=refield, min_size=min_size, sf=sf, modulo=modulo)
              for i in range(8)]
    E_list = [dihedral(E_list[i], dihedral_inverse_mode(i)) for i in range(8)]
    return torch.stack(E_list, dim=0).mean(dim=0, keepdim=False)


def test_mode(model, L, mode=0, refield=32, min_size=256, sf=1, modulo=1, ops=None):
    """utils_model.test_mode (:50-91): 0 normal, 1 pad, 2 split, 3 x8, 4 split + x8."""
    if mode == 0:
        return model(L)
    if mode == 1:
        return test_pad(model, L, modulo, sf)
    if mode == 2:
        return test_split_fn(model, L, refield, min_size, sf, modulo)
    if mode == 3:
        return test_x8(model, L, modulo, sf, ops=ops)
    if mode == 4:
        return test_split_x8(model, L, refield, min_size, sf, modulo)
    raise ValueError("test_mode: mode must be 0..4")
