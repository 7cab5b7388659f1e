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
out = torch.zeros(T, nH * 32, dtype=torch.bfloat16, device=dev)
for shift in (0, 4):
    for _ in range(2):
        ops.window_attn_fwd(qkv, table, out, batch=B, H=H, W=W, num_heads=nH, shift=shift)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(10):
        ops.window_attn_fwd(qkv, table, out, batch=B, H=H, W=W, num_heads=nH, shift=shift)
    b.record(); torch.cuda.synchronize()
    print("shift", shift, "us/launch", a.elapsed_time(b) * 100)
