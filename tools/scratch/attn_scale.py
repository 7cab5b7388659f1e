"""attention fwd / bwd time vs batch: slope = steady-state cost per window pair, intercept = fixed launch overhead"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200 import ops
H, W, nH = 128, 128, 6
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, n=10):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / n * 1e3
for B in (1, 2, 4, 8, 16, 32):
    T = B * H * W
    qkv = (torch.randn(T, 3 * nH * 32, device=dev) * 0.5).bfloat16()
    table = torch.randn(225, nH, device=dev) * 0.3
    dout = (torch.randn(T, nH * 32, device=dev) * 0.5).bfloat16()
    out = torch.empty(T, nH * 32, device=dev, dtype=torch.bfloat16)
    dqkv = torch.zeros_like(qkv); dtable = torch.zeros_like(table)
    f = ops.build_window_attn_fwd(qkv, table, out, batch=B, H=H, W=W, num_heads=nH, window=8, shift=4)
    st = ops.stream_handle()
    tf = timeit(lambda: f.run(st))
    tb = timeit(lambda: ops.window_attn_bwd(qkv, table, dout, dqkv, dtable, batch=B, H=H, W=W, num_heads=nH, shift=4))
    pairs = T // 128
    print("B=%2d pairs/CTA %6.2f  fwd %7.1f us  bwd %7.1f us   fwd GB/s %5.0f bwd GB/s %5.0f" % (
        B, pairs / 148, tf, tb, T * 768 * 2 / tf / 1e3, T * (576 * 2 + 192) * 2 / tb / 1e3))
    del qkv, dout, out, dqkv
