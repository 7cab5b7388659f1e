"""fixed cost of an attention launch: a one-window problem, and an empty kernel for reference"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200 import ops
nH = 6
dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, n=20, fl=True):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        if fl: flush.zero_()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / n * 1e3
x = torch.zeros(16, device=dev)
print("tiny torch kernel: %.1f us (flushed), %.1f us (not)" % (timeit(lambda: x.add_(1)), timeit(lambda: x.add_(1), fl=False)))
for (B, H, W) in ((1, 8, 8), (1, 16, 16), (1, 64, 64), (2, 128, 128)):
    T = B * H * W
    qkv = (torch.randn(T, 3 * nH * 32, device=dev) * 0.5).bfloat16()
    table = torch.randn(225, nH, device=dev) * 0.3
    dout = (torch.randn(T, nH * 32, device=dev) * 0.5).bfloat16()
    out = torch.empty(T, nH * 32, device=dev, dtype=torch.bfloat16)
    dqkv = torch.zeros_like(qkv); dtable = torch.zeros_like(table)
    f = ops.build_window_attn_fwd(qkv, table, out, batch=B, H=H, W=W, num_heads=nH, window=8, shift=4)
    st = ops.stream_handle()
    fb = lambda: ops.window_attn_bwd(qkv, table, dout, dqkv, dtable, batch=B, H=H, W=W, num_heads=nH, shift=4)
    print("B=%d %dx%d windows %4d: fwd %.1f us (no flush %.1f)  bwd %.1f us (no flush %.1f)" % (
        B, H, W, T // 64, timeit(lambda: f.run(st)), timeit(lambda: f.run(st), fl=False), timeit(fb), timeit(fb, fl=False)))
