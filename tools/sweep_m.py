"""fixed-vs-per-tile cost of gemm_tn: time vs M for a few shapes (development aid)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kair_b200 import ops
from bench_kernels import timeit
dev = "cuda"
for N, K, out in ((576, 192, False), (576, 192, True), (192, 192, False), (192, 192, True)):
    for T in (18944, 37888, 75776, 151552, 303104):   # multiples of 148*128
        A = (torch.randn(T, K, device=dev) * 0.5).bfloat16()
        B = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
        O = torch.empty(T, N, device=dev, dtype=torch.bfloat16) if out else None
        l = ops.build_gemm_tn(A, B, N=N, K=K, bn=192, outB=O)
        us = timeit(lambda: l.run(ops.stream_handle()))
        tiles = (T // 128) * (N // 192)
        print("N%d K%d out=%d T=%6d tiles/CTA=%5.1f  %7.1f us" % (N, K, out, T, tiles / 148.0, us))
