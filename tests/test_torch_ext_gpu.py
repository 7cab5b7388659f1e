"""torch.ops.kair_b200.* (the torch C++ extension binding) against the ctypes binding of the same C entry points."""
import pytest
import torch


@pytest.mark.gpu
@pytest.mark.parametrize("shift", [0, 4])
def test_window_attention_ops_match_ctypes_binding(shift):
    from kair_b200 import ops, torch_ext
    B, H, W, nH = 2, 24, 16, 6
    T = B * H * W
    g = torch.Generator().manual_seed(3)
    qkv = (torch.randn(T, 3 * nH * 32, generator=g) * 0.5).cuda().bfloat16()
    table = (torch.randn(225, nH, generator=g) * 0.3).cuda()
    dout = (torch.randn(T, nH * 32, generator=g) * 0.5).cuda().bfloat16()
    out = torch.empty(T, nH * 32, device="cuda", dtype=torch.bfloat16)
    ops.build_window_attn_fwd(qkv, table, out, batch=B, H=H, W=W, num_heads=nH, window=8, shift=shift).run(ops.stream_handle())
    dqkv, dtable = torch.zeros_like(qkv), torch.zeros_like(table)
    ops.window_attn_bwd(qkv, table, dout, dqkv, dtable, batch=B, H=H, W=W, num_heads=nH, shift=shift)
    t = torch_ext.load()
    assert torch.equal(t.window_attn_fwd(qkv, table, B, H, W, nH, 8, shift), out)
    dq2, dt2 = t.window_attn_bwd(qkv, table, dout, B, H, W, nH, 8, shift)
    assert torch.equal(dq2, dqkv)
    assert torch.allclose(dt2, dtable, rtol=1e-5, atol=1e-5)      # atomics: order of the partial sums
    # as an autograd.Function
    q = qkv.clone().requires_grad_(True)
    tb = table.clone().requires_grad_(True)
    y = torch_ext.WindowAttention.apply(q, tb, B, H, W, nH, 8, shift)
    y.backward(dout)
    assert torch.equal(y.detach(), out) and torch.equal(q.grad, dqkv)
    assert torch.allclose(tb.grad, dtable, rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
def test_pixel_loss_op_and_launch_list_replay():
    from kair_b200 import ops, torch_ext
    t = torch_ext.load()
    g = torch.Generator().manual_seed(1)
    x, y = torch.rand(3, 1, 40, 56, generator=g).cuda(), torch.rand(3, 1, 40, 56, generator=g).cuda()
    loss, dx = t.pixel_loss(x, y, 1e-9, 0, 0.0)
    ref = torch.sqrt((x - y) ** 2 + 1e-9)
    assert abs(loss.item() - ref.mean().item()) < 1e-6
    assert torch.allclose(dx, (x - y) / ref / x.numel(), rtol=1e-5, atol=1e-9)
    # a recorded launch list replayed through run_list: same result as the ctypes replay
    a = torch.randn(256, 64, device="cuda")
    b = torch.randn(256, 64, device="cuda")
    o1, o2 = torch.zeros_like(a), torch.zeros_like(a)
    l1 = ops.make_list([ops.build_axpby(a, n=64, a=2.0, y=b, b=-1.0, out_f32=o1)])
    l2 = ops.make_list([ops.build_axpby(a, n=64, a=2.0, y=b, b=-1.0, out_f32=o2)])
    l1.run(ops.stream_handle())
    l2.run_torch(a)
    torch.cuda.synchronize()
    assert torch.equal(o1, o2) and torch.allclose(o1, 2 * a - b)


@pytest.mark.gpu
def test_network_through_the_torch_op_replay(monkeypatch):
    """KAIR_B200_TORCH_OPS: a SwinIR training step with its launch lists replayed by torch.ops.kair_b200.run_list"""
    from kair_b200 import ops
    from kair_b200.network_swinir import SwinIR
    res = []
    for flag in (False, True):
        monkeypatch.setattr(ops, "USE_TORCH_OPS", flag)
        torch.manual_seed(0)
        m = SwinIR(upscale=1, in_chans=1, img_size=32, window_size=8, img_range=1.0, depths=[2], embed_dim=60,
                   num_heads=[6], mlp_ratio=2, upsampler="", resi_connection="1conv", drop_path_rate=0.0).cuda()
        x = torch.rand(2, 1, 32, 32, generator=torch.Generator().manual_seed(4)).cuda()
        y = m(x)
        y.square().mean().backward()
        res.append((y.detach().clone(), m.conv_first.weight.grad.clone()))
    assert torch.equal(res[0][0], res[1][0])
    assert torch.allclose(res[0][1], res[1][1], rtol=1e-4, atol=1e-6)   # split-K atomics
