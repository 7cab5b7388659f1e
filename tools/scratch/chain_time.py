"""kair_conv_chain vs per-layer kair_gemm_tn launches: time per layer and per tile (development aid; KB_CC_DBG ablations
need a --devel build)"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200 import ops
dev = torch.device("cuda:0")
L = 15
for B in (1, 16):
    H = W = 256
    T = B * H * W
    x = (torch.randn(T, 64, device=dev) * 0.5).bfloat16(); y = torch.zeros_like(x)
    Wp = (torch.randn(L, 64, 576, device=dev) * 0.05).bfloat16()
    bias = torch.zeros(L, 64, device=dev); slope = torch.zeros(L, device=dev)
    sync = torch.zeros(2, dtype=torch.int32, device=dev)
    ch = ops.build_conv_chain(x, y, Wp, bias, slope, sync, L=L, conv=(B, H, W))
    acts = (x, y)
    per = [ops.build_gemm_tn(acts[l & 1], Wp[l], N=64, K=64, bn=64, conv=(B, H, W), bias=bias[l], act=ops.ACT_LRELU,
                             act_slope=0.0, outB=acts[(l + 1) & 1]) for l in range(L)]
    st = ops.stream_handle()
    def timeit(fn, n=10):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record()
        for _ in range(n): fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / n * 1e3
    tc = timeit(lambda: ch.run(st))
    tp = timeit(lambda: [l.run(st) for l in per])
    tiles = B * 16 * 32
    per_sm = -(-tiles // 148)
    print("B=%2d tiles/SM %3d  chain %8.1f us (%.2f us/layer, %.2f us/tile)   per-layer launches %8.1f us (%.2f us/layer, %.2f us/tile)"
          % (B, per_sm, tc, tc / L, tc / L / per_sm, tp, tp / L, tp / L / per_sm))
