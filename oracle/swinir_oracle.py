"""ORACLE (test infrastructure, not product code): CPU restatement of the reference's SwinIR forward pass.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
It restates, in functional PyTorch on plain state_dict tensors, the arithmetic of /root/reference
models/network_swinir.py (file:line cited per function).  It is pinned against the real reference modules by
tests/test_oracle_golden_cpu.py (its live-reference cases run wherever /root/reference is mounted) and against the committed golden
vectors in tests/golden/ (generated from the real reference by tests/golden/make_golden.py).

Differences in *form* (not in arithmetic) from the reference: the shifted-window mask and the relative-position
index are computed from closed forms instead of being read from buffers; windows are gathered with index
arithmetic instead of roll + view/permute.  Autograd works through every op, so gradients come from
torch.autograd exactly as they do for the reference.
"""
import math
from collections import OrderedDict

import torch
import torch.nn.functional as F


class SwinIRConfig:
    """Constructor arguments of models/network_swinir.py:646-652 that change the arithmetic."""

    def __init__(self, img_size=64, in_chans=3, embed_dim=96, depths=(6, 6, 6, 6), num_heads=(6, 6, 6, 6),
                 window_size=7, mlp_ratio=4.0, qkv_bias=True, qk_scale=None, upscale=2, img_range=1.0, upsampler="",
                 resi_connection="1conv", patch_norm=True, ape=False):
        self.img_size = (img_size, img_size) if isinstance(img_size, int) else tuple(img_size)
        self.in_chans = in_chans
        self.embed_dim = embed_dim
        self.depths = tuple(depths)
        self.num_heads = tuple(num_heads)
        self.window_size = window_size
        self.mlp_ratio = mlp_ratio
        self.qkv_bias = qkv_bias
        self.qk_scale = qk_scale
        self.upscale = upscale
        self.img_range = img_range
        self.upsampler = upsampler or ""
        self.resi_connection = resi_connection
        self.patch_norm = patch_norm
        self.ape = ape
        assert not ape, "absolute position embedding is not used by any KAIR config"

    def block_geometry(self):
        """(window, shift) actually used by blocks built for img_size (network_swinir.py:193-196, 383)."""
        ws = self.window_size
        if min(self.img_size) <= ws:
            return min(self.img_size), 0
        return ws, ws // 2


# --------------------------------------------------------------------------------------------- closed forms

def relative_position_index(ws):
    """network_swinir.py:92-103 == (yi - yj + ws-1) * (2ws-1) + (xi - xj + ws-1)."""
    ys, xs = torch.meshgrid(torch.arange(ws), torch.arange(ws), indexing="ij")
    ys, xs = ys.reshape(-1), xs.reshape(-1)
    return (ys[:, None] - ys[None, :] + ws - 1) * (2 * ws - 1) + (xs[:, None] - xs[None, :] + ws - 1)


def region_id(pos, size, ws, shift):
    """Region label along one axis in the shifted frame (network_swinir.py:220-225): [0,size-ws) -> 0,
    [size-ws, size-shift) -> 1, [size-shift, size) -> 2."""
    return (pos >= size - ws).long() + (pos >= size - shift).long()


def shift_mask(H, W, ws, shift):
    """network_swinir.py:216-237: (nW, ws*ws, ws*ws) additive mask with values {0, -100}."""
    hh = region_id(torch.arange(H), H, ws, shift)
    wwid = region_id(torch.arange(W), W, ws, shift)
    ids = (3 * hh[:, None] + wwid[None, :]).float()  # (H, W) in the shifted frame
    ids = ids.view(H // ws, ws, W // ws, ws).permute(0, 2, 1, 3).reshape(-1, ws * ws)
    diff = ids[:, None, :] - ids[:, :, None]
    return torch.where(diff != 0, torch.full_like(diff, -100.0), torch.zeros_like(diff))


def window_token_index(H, W, ws, shift):
    """Flat source-token index (h*W + w) of every token of every window, shape (nW, ws*ws).
    Equivalent to roll(-shift) followed by window_partition (network_swinir.py:249-256, 33-45)."""
    wi, wj = torch.meshgrid(torch.arange(H // ws), torch.arange(W // ws), indexing="ij")
    r, c = torch.meshgrid(torch.arange(ws), torch.arange(ws), indexing="ij")
    h = (wi.reshape(-1, 1) * ws + r.reshape(1, -1) + shift) % H
    w = (wj.reshape(-1, 1) * ws + c.reshape(1, -1) + shift) % W
    return h * W + w


# --------------------------------------------------------------------------------------------- state layout

def state_spec(cfg):
    """Ordered [(key, shape, dtype)] of SwinIR.state_dict() (network_swinir.py:646-764)."""
    C = cfg.embed_dim
    c = cfg.in_chans
    H, W = cfg.img_size
    ws, shift = cfg.block_geometry()
    hidden = int(C * cfg.mlp_ratio)
    spec = []

    def add(k, shape, dt=torch.float32):
        spec.append((k, tuple(shape), dt))

    def conv(prefix, cout, cin, k=3):
        add(prefix + ".weight", (cout, cin, k, k))
        add(prefix + ".bias", (cout,))

    conv("conv_first", C, c)
    if cfg.patch_norm:
        add("patch_embed.norm.weight", (C,))
        add("patch_embed.norm.bias", (C,))
    for i, (depth, nh) in enumerate(zip(cfg.depths, cfg.num_heads)):
        for j in range(depth):
            p = "layers.%d.residual_group.blocks.%d." % (i, j)
            if j % 2 == 1 and shift > 0:
                add(p + "attn_mask", ((H // ws) * (W // ws), ws * ws, ws * ws))
            add(p + "norm1.weight", (C,))
            add(p + "norm1.bias", (C,))
            add(p + "attn.relative_position_bias_table", ((2 * ws - 1) ** 2, nh))
            add(p + "attn.relative_position_index", (ws * ws, ws * ws), torch.int64)
            add(p + "attn.qkv.weight", (3 * C, C))
            if cfg.qkv_bias:
                add(p + "attn.qkv.bias", (3 * C,))
            add(p + "attn.proj.weight", (C, C))
            add(p + "attn.proj.bias", (C,))
            add(p + "norm2.weight", (C,))
            add(p + "norm2.bias", (C,))
            add(p + "mlp.fc1.weight", (hidden, C))
            add(p + "mlp.fc1.bias", (hidden,))
            add(p + "mlp.fc2.weight", (C, hidden))
            add(p + "mlp.fc2.bias", (C,))
        if cfg.resi_connection == "1conv":
            conv("layers.%d.conv" % i, C, C)
        else:
            conv("layers.%d.conv.0" % i, C // 4, C)
            conv("layers.%d.conv.2" % i, C // 4, C // 4, 1)
            conv("layers.%d.conv.4" % i, C, C // 4)
    add("norm.weight", (C,))
    add("norm.bias", (C,))
    if cfg.resi_connection == "1conv":
        conv("conv_after_body", C, C)
    else:
        conv("conv_after_body.0", C // 4, C)
        conv("conv_after_body.2", C // 4, C // 4, 1)
        conv("conv_after_body.4", C, C // 4)
    nf = 64
    if cfg.upsampler == "pixelshuffle":
        conv("conv_before_upsample.0", nf, C)
        s = cfg.upscale
        if (s & (s - 1)) == 0:
            for n in range(int(math.log2(s))):
                conv("upsample.%d" % (2 * n), 4 * nf, nf)
        elif s == 3:
            conv("upsample.0", 9 * nf, nf)
        conv("conv_last", c, nf)
    elif cfg.upsampler == "pixelshuffledirect":
        conv("upsample.0", cfg.upscale ** 2 * c, C)
    elif cfg.upsampler == "nearest+conv":
        conv("conv_before_upsample.0", nf, C)
        conv("conv_up1", nf, nf)
        conv("conv_up2", nf, nf)
        conv("conv_hr", nf, nf)
        conv("conv_last", c, nf)
    else:
        conv("conv_last", c, C)
    return spec


def deterministic_state(spec, seed, cfg=None):
    """Reference-independent, reproducible parameter values (CPU generator, one stream per key)."""
    sd = OrderedDict()
    for i, (k, shape, dt) in enumerate(spec):
        g = torch.Generator().manual_seed(seed * 1000003 + i)
        if k.endswith("relative_position_index"):
            ws = int(round(math.sqrt(shape[0])))
            sd[k] = relative_position_index(ws)
        elif k.endswith("attn_mask"):
            ws, shift = cfg.block_geometry()
            sd[k] = shift_mask(cfg.img_size[0], cfg.img_size[1], ws, shift)
        elif k.endswith("running_var"):
            sd[k] = torch.rand(shape, generator=g) + 0.5
        elif k.endswith("running_mean"):
            sd[k] = torch.randn(shape, generator=g) * 0.1
        elif k.endswith("num_batches_tracked"):
            sd[k] = torch.zeros(shape, dtype=torch.int64)
        elif "norm" in k.split(".")[-2] and k.endswith("weight") or (len(shape) == 1 and k.endswith("weight")):
            sd[k] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif k.endswith("bias"):
            sd[k] = 0.05 * torch.randn(shape, generator=g)
        elif k.endswith("relative_position_bias_table"):
            sd[k] = 0.2 * torch.randn(shape, generator=g)
        else:
            fan_in = 1
            for d in shape[1:]:
                fan_in *= d
            sd[k] = torch.randn(shape, generator=g) * (0.7 / math.sqrt(fan_in))
    return sd


# --------------------------------------------------------------------------------------------- forward

def window_attention(xw, sd, p, nh, ws, mask, qk_scale=None):
    """WindowAttention.forward (network_swinir.py:114-145); xw: (B_, N, C), mask: (nW, N, N) or None."""
    B_, N, C = xw.shape
    hd = C // nh
    scale = qk_scale or hd ** -0.5
    qkv = F.linear(xw, sd[p + "qkv.weight"], sd.get(p + "qkv.bias"))
    qkv = qkv.reshape(B_, N, 3, nh, hd).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0] * scale, qkv[1], qkv[2]
    attn = q @ k.transpose(-2, -1)
    idx = relative_position_index(ws).reshape(-1)
    bias = sd[p + "relative_position_bias_table"][idx].view(N, N, nh).permute(2, 0, 1)
    attn = attn + bias.unsqueeze(0)
    if mask is not None:
        nW = mask.shape[0]
        attn = attn.view(B_ // nW, nW, nh, N, N) + mask.to(attn.dtype)[None, :, None]
        attn = attn.view(-1, nh, N, N)
    attn = attn.softmax(dim=-1)
    out = (attn @ v).transpose(1, 2).reshape(B_, N, C)
    return F.linear(out, sd[p + "proj.weight"], sd[p + "proj.bias"])


def swin_block(x, x_size, sd, p, nh, ws, shift, cfg, drop_mask=None):
    """SwinTransformerBlock.forward (network_swinir.py:239-279). drop_mask: None, or the pair of per-sample DropPath
    scales (attention branch, MLP branch) -- the reference draws them independently (:276 and :277)."""
    H, W = x_size
    B, L, C = x.shape
    if min(x_size) <= ws:  # (:193-196) is decided at construction from img_size; tests use matching sizes
        pass
    shortcut = x
    xn = F.layer_norm(x, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"], 1e-5)
    idx = window_token_index(H, W, ws, shift).reshape(-1)           # (nW*N,)
    xw = xn[:, idx, :].reshape(-1, ws * ws, C)                       # (B*nW, N, C)
    mask = shift_mask(H, W, ws, shift) if shift > 0 else None
    aw = window_attention(xw, sd, p + "attn.", nh, ws, mask, cfg.qk_scale)
    out = torch.empty_like(xn)
    out[:, idx, :] = aw.reshape(B, -1, C)                            # window_reverse + roll(+shift)
    if drop_mask is not None:
        out = out * drop_mask[0].view(B, 1, 1)
    x = shortcut + out
    y = F.layer_norm(x, (C,), sd[p + "norm2.weight"], sd[p + "norm2.bias"], 1e-5)
    y = F.linear(F.gelu(F.linear(y, sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"])),
                 sd[p + "mlp.fc2.weight"], sd[p + "mlp.fc2.bias"])
    if drop_mask is not None:
        y = y * drop_mask[1].view(B, 1, 1)
    return x + y


def _resi_conv(x, sd, prefix, mode):
    """RSTB.conv / conv_after_body in '1conv' or '3conv' form (network_swinir.py:464-471, 728-736)."""
    if mode == "1conv":
        return F.conv2d(x, sd[prefix + ".weight"], sd[prefix + ".bias"], padding=1)
    x = F.leaky_relu(F.conv2d(x, sd[prefix + ".0.weight"], sd[prefix + ".0.bias"], padding=1), 0.2)
    x = F.leaky_relu(F.conv2d(x, sd[prefix + ".2.weight"], sd[prefix + ".2.bias"]), 0.2)
    return F.conv2d(x, sd[prefix + ".4.weight"], sd[prefix + ".4.bias"], padding=1)


def forward_features(x, sd, cfg, drop_masks=None):
    """SwinIR.forward_features (network_swinir.py:790-803) incl. RSTB.forward (:481-482)."""
    B, C, H, W = x.shape
    ws, shift_sz = cfg.block_geometry() if (H, W) == cfg.img_size else (cfg.window_size, cfg.window_size // 2)
    if min(cfg.img_size) <= cfg.window_size:  # geometry is frozen at construction (:193-196)
        ws, shift_sz = cfg.block_geometry()
    t = x.flatten(2).transpose(1, 2)
    if cfg.patch_norm:
        t = F.layer_norm(t, (C,), sd["patch_embed.norm.weight"], sd["patch_embed.norm.bias"], 1e-5)
    blk = 0
    for i, (depth, nh) in enumerate(zip(cfg.depths, cfg.num_heads)):
        t_in = t
        for j in range(depth):
            p = "layers.%d.residual_group.blocks.%d." % (i, j)
            shift = shift_sz if j % 2 == 1 else 0
            dm = None if drop_masks is None else drop_masks[blk]
            t = swin_block(t, (H, W), sd, p, nh, ws, shift, cfg, dm)
            blk += 1
        img = t.transpose(1, 2).reshape(B, C, H, W)
        img = _resi_conv(img, sd, "layers.%d.conv" % i, cfg.resi_connection)
        t = img.flatten(2).transpose(1, 2) + t_in
    t = F.layer_norm(t, (C,), sd["norm.weight"], sd["norm.bias"], 1e-5)
    return t.transpose(1, 2).reshape(B, C, H, W)


def swinir_forward(sd, x, cfg, drop_masks=None):
    """SwinIR.forward (network_swinir.py:805-839)."""
    H, W = x.shape[2:]
    ws = cfg.window_size
    ph, pw = (ws - H % ws) % ws, (ws - W % ws) % ws
    if ph or pw:
        x = F.pad(x, (0, pw, 0, ph), "reflect")
    if cfg.in_chans == 3:
        mean = torch.tensor((0.4488, 0.4371, 0.4040), dtype=x.dtype).view(1, 3, 1, 1)
    else:
        mean = torch.zeros(1, 1, 1, 1, dtype=x.dtype)
    x = (x - mean) * cfg.img_range

    def conv(t, name, pad=1):
        return F.conv2d(t, sd[name + ".weight"], sd[name + ".bias"], padding=pad)

    first = conv(x, "conv_first")
    body = _resi_conv(forward_features(first, sd, cfg, drop_masks), sd, "conv_after_body", cfg.resi_connection) + first
    if cfg.upsampler == "pixelshuffle":
        y = F.leaky_relu(conv(body, "conv_before_upsample.0"), 0.01)
        s = cfg.upscale
        if (s & (s - 1)) == 0:
            for n in range(int(math.log2(s))):
                y = F.pixel_shuffle(conv(y, "upsample.%d" % (2 * n)), 2)
        else:
            y = F.pixel_shuffle(conv(y, "upsample.0"), 3)
        y = conv(y, "conv_last")
    elif cfg.upsampler == "pixelshuffledirect":
        y = F.pixel_shuffle(conv(body, "upsample.0"), cfg.upscale)
    elif cfg.upsampler == "nearest+conv":
        y = F.leaky_relu(conv(body, "conv_before_upsample.0"), 0.01)
        y = F.leaky_relu(conv(F.interpolate(y, scale_factor=2, mode="nearest"), "conv_up1"), 0.2)
        y = F.leaky_relu(conv(F.interpolate(y, scale_factor=2, mode="nearest"), "conv_up2"), 0.2)
        y = conv(F.leaky_relu(conv(y, "conv_hr"), 0.2), "conv_last")
    else:
        y = x + conv(body, "conv_last")
    y = y / cfg.img_range + mean
    return y[:, :, :H * cfg.upscale, :W * cfg.upscale]


# --------------------------------------------------------------------------------------------- losses / metrics

def charbonnier(x, y, eps=1e-9):
    """models/loss.py:208-218."""
    d = x - y
    return torch.mean(torch.sqrt(d * d + eps))


def psnr_uint8(a, b):
    """utils/utils_image.py:296-300, 629-645 on float images in [0,1]."""
    a8 = (a.detach().clamp(0, 1) * 255.0).round().double()
    b8 = (b.detach().clamp(0, 1) * 255.0).round().double()
    mse = torch.mean((a8 - b8) ** 2).item()
    return float("inf") if mse == 0 else 20.0 * math.log10(255.0 / math.sqrt(mse))
