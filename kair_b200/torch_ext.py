"""The torch C++ extension binding of the C ABI (csrc_torch/kair_b200_torch.cpp): `torch.ops.kair_b200.*`.

A second binding next to the ctypes one in `_lib.py`, in the style of the reference's own CUDA ops
(models/op/fused_act.py:10-72: a `torch.utils.cpp_extension` module wrapped by autograd.Functions).  The engines replay
their launch lists through ctypes by default; `KAIR_B200_TORCH_OPS=1` routes the replay through
`torch.ops.kair_b200.run_list` instead (same C entry point, same stream)."""
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libkair_b200_torch.so")
_loaded = False


def load():
    """Register the ops (once).  Raises if the extension has not been built (`python -m kair_b200.build`)."""
    global _loaded
    if not _loaded:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("kair_b200: %s not found. Build it with `python -m kair_b200.build`." % LIB_PATH)
        torch.ops.load_library(LIB_PATH)
        _loaded = True
    return torch.ops.kair_b200


class WindowAttention(torch.autograd.Function):
    """The fused shifted-window attention core as an autograd.Function over the torch ops
    (models/network_swinir.py:124-142 with the window partition / shift / mask of :33-62, 216-237, 249-272)."""

    @staticmethod
    def forward(ctx, qkv, table, batch, H, W, heads, window, shift):
        ops = load()
        ctx.save_for_backward(qkv, table)
        ctx.geom = (batch, H, W, heads, window, shift)
        return ops.window_attn_fwd(qkv, table, batch, H, W, heads, window, shift)

    @staticmethod
    def backward(ctx, dout):
        qkv, table = ctx.saved_tensors
        dqkv, dtable = load().window_attn_bwd(qkv, table, dout.contiguous(), *ctx.geom)
        return dqkv, dtable, None, None, None, None, None, None
