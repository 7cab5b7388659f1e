"""kair_gemm_nt `swap` (3x3 convolution weight gradient with the operand roles swapped, result transposed) against the
plain order on the same inputs, and kair_unpack_conv_grad `transposed` against the plain unpack."""
import pytest
import torch


@pytest.mark.gpu
@pytest.mark.parametrize("geom", [(2, 32, 32), (1, 37, 45), (3, 8, 8), (16, 128, 128)])
@pytest.mark.parametrize("n1,dil", [(64, 1), (128, 1), (64, 2)])
def test_swap_matches_plain_order(geom, n1, dil):
    _swap_vs_plain(geom, n1, 64, dil)


@pytest.mark.gpu
@pytest.mark.parametrize("n2", [128, 192])
def test_swap_with_wider_inputs(n2):
    """the dense-block convolutions of RRDBNet: 64 output channels, 128 / 192 (padded) input channels"""
    _swap_vs_plain((2, 24, 40), 64, n2, 1)


def _swap_vs_plain(geom, n1, n2, dil):
    from kair_b200 import ops
    B, H, W = geom
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(B * 1000 + H + n1)
    T = B * H * W
    dout = (torch.randn(T, n1, generator=g) * 0.5).to(dev).bfloat16()
    x = (torch.randn(T, n2, generator=g) * 0.5).to(dev).bfloat16()
    st = ops.stream_handle()
    D = torch.zeros(n1, 9 * n2, device=dev)
    ops.build_gemm_nt(dout, x, D, n1=n1, n2=n2, bn=64, conv=geom, dilation=dil).run(st)
    Dt = torch.zeros(9 * n2, n1, device=dev)
    ops.build_gemm_nt(dout, x, Dt, n1=n1, n2=n2, bn=64, conv=geom, dilation=dil, swap=True).run(st)
    torch.cuda.synchronize()
    # same bf16 products, fp32 accumulation in a different order (split-K partials are added atomically in both)
    scale = D.abs().max().item()
    assert scale > 0
    assert (Dt.t() - D).abs().max().item() <= 2e-5 * scale + 1e-6 * (T ** 0.5)
    # unpack of both layouts gives the same (O, I, 3, 3) gradient
    O = n1 - 3
    I = n2 - 3
    w1 = torch.zeros(O, I, 3, 3, device=dev)
    w2 = torch.zeros_like(w1)
    ops.build_unpack_conv_grad(D[None], [w1], None, O=O, I=I, I_pad=n2).run(st)
    ops.build_unpack_conv_grad(Dt[None], [w2], None, O=O, I=I, I_pad=n2, transposed=True).run(st)
    torch.cuda.synchronize()
    assert torch.equal(w2, Dt.t().reshape(n1, 9, n2)[:O, :, :I].permute(0, 2, 1).reshape(O, I, 3, 3))
    assert (w1 - w2).abs().max().item() <= 2e-5 * scale + 1e-6 * (T ** 0.5)


@pytest.mark.gpu
def test_swap_rejects_other_shapes():
    from kair_b200 import ops
    dev = torch.device("cuda:0")
    a = torch.zeros(64, 64, dtype=torch.bfloat16, device=dev)
    b = torch.zeros(64, 96, dtype=torch.bfloat16, device=dev)
    D = torch.zeros(9 * 96, 64, device=dev)
    with pytest.raises(RuntimeError):
        ops.build_gemm_nt(a, b, D, n1=64, n2=96, bn=64, conv=(1, 8, 8), swap=True).run(ops.stream_handle())
