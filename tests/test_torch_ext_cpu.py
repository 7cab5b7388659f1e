"""The torch C++ extension binding (kair_b200/csrc_torch): loads on a CPU-only box and registers its ops; no compute."""
import pytest
import torch


def test_torch_extension_loads_and_registers_ops():
    from kair_b200 import _lib, torch_ext
    ops = torch_ext.load()
    assert ops.version() == _lib.load().kair_version()
    for name in ("window_attn_fwd", "window_attn_bwd", "pixel_loss", "run_list"):
        assert hasattr(ops, name), name
    # no CPU fallback: a CPU tensor is refused with a clear message
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        ops.window_attn_fwd(torch.zeros(64, 576), torch.zeros(225, 6), 1, 8, 8, 6, 8, 0)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        ops.pixel_loss(torch.zeros(4), torch.zeros(4), 1e-9, 0, 0.0)
