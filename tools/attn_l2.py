"""attention fwd time per window pair when qkv is L2-resident (small batch) vs streaming from HBM (development aid)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from kair_b200 import ops
from bench_kernels import timeit
dev = "cuda"
for B in (1, 2, 4, 8, 16):
    T = B * 128 * 128
    qkv = (torch.randn(T, 576, device=dev) * 0.5).bfloat16(); out = torch.empty(T, 192, device=dev, dtype=torch.bfloat16)
    table = torch.randn(225, 6, device=dev) * 0.1
    l = ops.build_window_attn_fwd(qkv, table, out, batch=B, H=128, W=128, num_heads=6, window=8, shift=4)
    us = timeit(lambda: l.run(ops.stream_handle()))
    pairs = T // 128
    print("B=%2d qkv %6.1f MB  %7.1f us  pairs/CTA %5.2f  us/pair/CTA %6.2f" % (B, T * 576 * 2 / 1e6, us, pairs / 148, us / (pairs / 148)))
