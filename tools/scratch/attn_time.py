"""isolated attention fwd / bwd launch times at the cfg2 shape (production build; development aid)"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200 import ops
B, H, W, nH = 8, 128, 128, 6
dev = torch.device("cuda:0")
T = B * H * W
g = torch.Generator().manual_seed(0)
qkv = (torch.randn(T, 3 * nH * 32, generator=g) * 0.5).to(dev).bfloat16()
table = (torch.randn(225, nH, generator=g) * 0.3).to(dev)
dout = (torch.randn(T, nH * 32, generator=g) * 0.5).to(dev).bfloat16()
out = torch.empty(T, nH * 32, device=dev, dtype=torch.bfloat16)
dqkv = torch.zeros_like(qkv); dtable = torch.zeros_like(table)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn, n=20, do_flush=False):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        if do_flush: flush.zero_()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / n * 1e3
for shift in (0, 4):
    f = ops.build_window_attn_fwd(qkv, table, out, batch=B, H=H, W=W, num_heads=nH, window=8, shift=shift)
    st = ops.stream_handle()
    print("fwd shift", shift, "us/launch %.1f  (L2 flushed %.1f)" % (timeit(lambda: f.run(st)), timeit(lambda: f.run(st), do_flush=True)))
    fb = lambda: ops.window_attn_bwd(qkv, table, dout, dqkv, dtable, batch=B, H=H, W=W, num_heads=nH, shift=shift)
    print("bwd shift", shift, "us/launch %.1f  (L2 flushed %.1f)" % (timeit(fb), timeit(fb, do_flush=True)))
