"""Fused tcgen05 window attention vs a plain PyTorch fp32 evaluation of network_swinir.py:124-142 (+ roll/partition)."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def _ref_attention(qkv, table, B, H, W, nH, shift, hd=32, ws=8):
    """qkv: fp32 [B*H*W, 3*nH*hd] (q pre-scaled). Returns ([B*H*W, nH*hd], P) computed window by window."""
    from oracle import swinir_oracle as so
    N = ws * ws
    idx = so.window_token_index(H, W, ws, shift).to(qkv.device)            # (nW, N)
    nW = idx.shape[0]
    x = qkv.view(B, H * W, 3, nH, hd)[:, idx.reshape(-1)]                    # (B, nW*N, 3, nH, hd)
    x = x.view(B * nW, N, 3, nH, hd).permute(2, 0, 3, 1, 4)                  # (3, B_, nH, N, hd)
    q, k, v = x[0], x[1], x[2]
    attn = q @ k.transpose(-2, -1)
    rpi = so.relative_position_index(ws).to(qkv.device).reshape(-1)
    bias = table[rpi].view(N, N, nH).permute(2, 0, 1)
    attn = attn + bias.unsqueeze(0)
    if shift > 0:
        mask = so.shift_mask(H, W, ws, shift).to(qkv.device)
        attn = attn.view(B, nW, nH, N, N) + mask[None, :, None]
        attn = attn.view(-1, nH, N, N)
    attn = attn.softmax(-1)
    o = (attn @ v).transpose(1, 2).reshape(B, nW * N, nH * hd)               # windows order
    out = torch.empty(B, H * W, nH * hd, device=qkv.device)
    out[:, idx.reshape(-1)] = o
    return out.view(B * H * W, nH * hd)


def _make_qkv(B, H, W, nH, seed, dev, ws=8):
    g = torch.Generator().manual_seed(seed)
    T = B * H * W
    qkv = torch.randn(T, 3 * nH * 32, generator=g) * 0.8
    qkv = qkv.view(T, 3, nH, 32)
    qkv[..., 30:] = 0          # head_dim 30 padded to 32
    qkv[:, 2, 0, 30] = 1.0     # ones column of v (head 0) used for the proj bias gradient
    qkv = qkv.view(T, -1)
    table = torch.randn((2 * ws - 1) ** 2, nH, generator=g) * 0.5
    return qkv.to(dev).bfloat16(), table.to(dev)


@pytest.mark.parametrize("B,H,W", [(2, 16, 24), (1, 24, 24), (1, 8, 8), (3, 32, 16)])
@pytest.mark.parametrize("shift", [0, 4])
def test_window_attn_fwd(B, H, W, shift):
    from kair_b200 import ops
    dev = torch.device("cuda:0")
    nH = 6
    qkv, table = _make_qkv(B, H, W, nH, 100 + H + W + shift, dev)
    out = torch.zeros(B * H * W, nH * 32, dtype=torch.bfloat16, device=dev)
    ops.window_attn_fwd(qkv, table, out, batch=B, H=H, W=W, num_heads=nH, shift=shift)
    torch.cuda.synchronize()
    ref = _ref_attention(qkv.float(), table, B, H, W, nH, shift)
    err = (out.float() - ref).abs().max().item()
    assert err < 3e-2, err
    assert (out.view(-1, nH, 32)[:, 0, 30].float() - 1.0).abs().max().item() < 2e-2


@pytest.mark.parametrize("B,H,W", [(2, 16, 24), (1, 24, 24), (1, 8, 8), (2, 64, 64)])
@pytest.mark.parametrize("shift", [0, 4])
def test_window_attn_bwd(B, H, W, shift):
    from kair_b200 import ops
    dev = torch.device("cuda:0")
    nH = 6
    qkv, table = _make_qkv(B, H, W, nH, 300 + H + W + shift, dev)
    g = torch.Generator().manual_seed(17)
    dout = (torch.randn(B * H * W, nH * 32, generator=g) * 0.5).to(dev).bfloat16()
    dqkv = torch.zeros_like(qkv)
    dtable = torch.zeros_like(table)
    ops.window_attn_bwd(qkv, table, dout, dqkv, dtable, batch=B, H=H, W=W, num_heads=nH, shift=shift)
    torch.cuda.synchronize()
    qf = qkv.float().requires_grad_(True)
    tf = table.clone().requires_grad_(True)
    ref = _ref_attention(qf, tf, B, H, W, nH, shift)
    (ref * dout.float()).sum().backward()
    gq = qf.grad
    e1 = (dqkv.float() - gq).abs().max().item() / (gq.abs().max().item() + 1e-9)
    e2 = (dtable - tf.grad).abs().max().item() / (tf.grad.abs().max().item() + 1e-9)
    assert e1 < 2e-2, ("dqkv", e1)
    assert e2 < 2e-2, ("dtable", e2)


# ------------------------------------------------------------------ windows smaller than 8 (CUDA-core kernels, attn_small.cu)
SMALL = [(7, 3, 2, 14, 21, 6), (7, 0, 1, 28, 28, 6), (7, 3, 1, 126, 126, 6), (4, 2, 2, 8, 12, 4), (5, 1, 1, 10, 15, 2)]


@pytest.mark.parametrize("ws,shift,B,H,W,nH", SMALL)
def test_small_window_attn_fwd(ws, shift, B, H, W, nH):
    """7x7 windows of the JPEG-artifact models (main_test_swinir.py:170-175): shift 3 = ws // 2, 49 tokens / window."""
    from kair_b200 import ops
    dev = torch.device("cuda:0")
    qkv, table = _make_qkv(B, H, W, nH, 500 + H + W + shift, dev, ws=ws)
    out = torch.zeros(B * H * W, nH * 32, dtype=torch.bfloat16, device=dev)
    ops.window_attn_fwd(qkv, table, out, batch=B, H=H, W=W, num_heads=nH, window=ws, shift=shift)
    torch.cuda.synchronize()
    ref = _ref_attention(qkv.float(), table, B, H, W, nH, shift, ws=ws)
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2, err
    assert (out.view(-1, nH, 32)[:, 0, 30].float() - 1.0).abs().max().item() < 1e-2


@pytest.mark.parametrize("ws,shift,B,H,W,nH", SMALL)
def test_small_window_attn_bwd(ws, shift, B, H, W, nH):
    from kair_b200 import ops
    dev = torch.device("cuda:0")
    qkv, table = _make_qkv(B, H, W, nH, 700 + H + W + shift, dev, ws=ws)
    g = torch.Generator().manual_seed(19)
    dout = (torch.randn(B * H * W, nH * 32, generator=g) * 0.5).to(dev).bfloat16()
    dqkv = torch.zeros_like(qkv)
    dtable = torch.zeros_like(table)
    ops.window_attn_bwd(qkv, table, dout, dqkv, dtable, batch=B, H=H, W=W, num_heads=nH, window=ws, shift=shift)
    torch.cuda.synchronize()
    qf = qkv.float().requires_grad_(True)
    tf = table.clone().requires_grad_(True)
    ref = _ref_attention(qf, tf, B, H, W, nH, shift, ws=ws)
    (ref * dout.float()).sum().backward()
    gq = qf.grad
    e1 = (dqkv.float() - gq).abs().max().item() / (gq.abs().max().item() + 1e-9)
    e2 = (dtable - tf.grad).abs().max().item() / (tf.grad.abs().max().item() + 1e-9)
    assert e1 < 1e-2, ("dqkv", e1)
    assert e2 < 1e-3, ("dtable", e2)
