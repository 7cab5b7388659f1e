"""kair_conv_chain (one persistent launch for a stack of 3x3 convolutions 64 -> 64) against kair_gemm_tn launched once
per layer over the same buffers: bit-identical, for ragged geometries, 1..17 layers, with and without bias, mixed
activations, twice in a row (the kernel re-arms its barrier words)."""
import pytest
import torch


def _run_pair(B, H, W, L, bias_on=True, seed=0):
    from kair_b200 import ops
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(seed)
    T = B * H * W
    x = (torch.randn(T, 64, generator=g) * 0.7).to(dev).bfloat16()
    Wp = (torch.randn(L, 64, 576, generator=g) * (1.4 / 24.0)).to(dev).bfloat16()
    bias = (torch.randn(L, 64, generator=g) * 0.1).to(dev) if bias_on else None
    slopes = [0.0, 0.2, 1.0]
    slope = torch.tensor([slopes[l % 3] for l in range(L)], dtype=torch.float32, device=dev)
    st = ops.stream_handle()
    # reference: one gemm_tn per layer
    ra = [x.clone(), torch.zeros_like(x)]
    for l in range(L):
        ops.build_gemm_tn(ra[l & 1], Wp[l], N=64, K=64, bn=64, conv=(B, H, W), bias=None if bias is None else bias[l],
                          act=ops.ACT_LRELU, act_slope=slopes[l % 3], outB=ra[(l + 1) & 1]).run(st)
    ref = ra[L & 1]
    ca = [x.clone(), torch.zeros_like(x)]
    sync = torch.zeros(2, dtype=torch.int32, device=dev)
    launch = ops.build_conv_chain(ca[0], ca[1], Wp, bias, slope, sync, L=L, conv=(B, H, W))
    outs = []
    for _ in range(2):
        ca[0].copy_(x)
        ca[1].zero_()
        launch.run(st)
        torch.cuda.synchronize()
        outs.append(ca[L & 1].clone())
        assert sync.tolist() == [0, 0]
    return ref, outs


@pytest.mark.gpu
@pytest.mark.parametrize("geom", [(1, 256, 256), (1, 37, 45), (3, 16, 8), (2, 100, 60), (1, 8, 8)])
@pytest.mark.parametrize("L", [1, 2, 15])
def test_chain_matches_per_layer_launches(geom, L):
    B, H, W = geom
    ref, outs = _run_pair(B, H, W, L, seed=L)
    assert torch.isfinite(ref.float()).all() and ref.float().abs().max() > 0
    for o in outs:
        assert torch.equal(o, ref)


@pytest.mark.gpu
def test_chain_without_bias_and_many_layers():
    ref, outs = _run_pair(1, 64, 64, 32, bias_on=False, seed=3)
    for o in outs:
        assert torch.equal(o, ref)


@pytest.mark.gpu
def test_chain_rejects_bad_arguments():
    from kair_b200 import ops
    dev = torch.device("cuda:0")
    a = torch.zeros(64, 64, dtype=torch.bfloat16, device=dev)
    b = torch.zeros_like(a)
    Wp = torch.zeros(33, 64, 576, dtype=torch.bfloat16, device=dev)
    slope = torch.zeros(33, device=dev)
    sync = torch.zeros(2, dtype=torch.int32, device=dev)
    with pytest.raises(RuntimeError, match="layers"):
        ops.build_conv_chain(a, b, Wp, None, slope, sync, L=33, conv=(1, 8, 8)).run(ops.stream_handle())


@pytest.mark.gpu
def test_dncnn_inference_uses_the_chain_and_matches_per_layer_plan(monkeypatch):
    """the module-level path: same DnCNN, plan with the chain vs plan without it"""
    from kair_b200 import convnet_engine
    from kair_b200.network_dncnn import DnCNN
    torch.manual_seed(0)
    x = torch.rand(1, 1, 96, 80).cuda()
    outs = []
    for limit in (2048, 0):
        monkeypatch.setattr(convnet_engine, "CHAIN_MAX_TILES", limit)
        torch.manual_seed(1)
        m = DnCNN(in_nc=1, out_nc=1, nc=64, nb=17, act_mode="R").cuda().eval()
        with torch.no_grad():
            outs.append(m(x).clone())
        assert bool(m.engine(x.device).plan(1, 96, 80, False, False).chain) == (limit > 0)
    assert torch.equal(outs[0], outs[1])
