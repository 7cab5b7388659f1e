"""kair_gemm_nt `swap` (3x3 convolution weight gradient with the operand roles swapped, result transposed) against the
plain order on the same inputs, and kair_unpack_conv_grad `transposed` against the plain unpack."""
import pytest
import torch


@pytest.mark.gpu
@pytest.mark.parametrize("geom", [(2, 32, 32), (1, 37, 45), (3, 8, 8), (16, 128, 128)])
@pytest.mark.parametrize("n1,dil", [(64, 1), (128, 1), (64, 2)])
def test_swap_matches_plain_order(geom, n1, dil):
    from kair_b200 import ops
    B, H, W = geom
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(B * 1000 + H + n1)
    T = B * H * W
    dout = (torch.randn(T, n1, generator=g) * 0.5).to(dev).bfloat16()
    x = (torch.randn(T, 64, generator=g) * 0.5).to(dev).bfloat16()
    st = ops.stream_handle()
    D = torch.zeros(n1, 9 * 64, device=dev)
    ops.build_gemm_nt(dout, x, D, n1=n1, n2=64, bn=64, conv=geom, dilation=dil).run(st)
    Dt = torch.zeros(9 * 64, n1, device=dev)
    ops.build_gemm_nt(dout, x, Dt, n1=n1, n2=64, bn=64, conv=geom, dilation=dil, swap=True).run(st)
    torch.cuda.synchronize()
    # same bf16 products, fp32 accumulation in a different order (split-K partials are added atomically in both)
    scale = D.abs().max().item()
    assert scale > 0
    assert (Dt.t() - D).abs().max().item() <= 2e-5 * scale + 1e-6 * (T ** 0.5)
    # unpack of both layouts gives the same (O, I, 3, 3) gradient
    O = n1 - 3
    w1 = torch.zeros(O, 61, 3, 3, device=dev)
    w2 = torch.zeros_like(w1)
    ops.build_unpack_conv_grad(D[None], [w1], None, O=O, I=61, I_pad=64).run(st)
    ops.build_unpack_conv_grad(Dt[None], [w2], None, O=O, I=61, I_pad=64, transposed=True).run(st)
    torch.cuda.synchronize()
    assert torch.equal(w2, Dt.t().reshape(n1, 9, 64)[:O, :, :61].permute(0, 2, 1).reshape(O, 61, 3, 3))
    assert (w1 - w2).abs().max().item() <= 2e-5 * scale + 1e-6 * (T ** 0.5)


@pytest.mark.gpu
def test_swap_rejects_other_shapes():
    from kair_b200 import ops
    dev = torch.device("cuda:0")
    a = torch.zeros(64, 64, dtype=torch.bfloat16, device=dev)
    b = torch.zeros(64, 128, dtype=torch.bfloat16, device=dev)
    D = torch.zeros(9 * 128, 64, device=dev)
    with pytest.raises(RuntimeError):
        ops.build_gemm_nt(a, b, D, n1=64, n2=128, bn=64, conv=(1, 8, 8), swap=True).run(ops.stream_handle())
