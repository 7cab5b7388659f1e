import torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_kernels import timeit
x = torch.randn(512 * 1024 * 1024, device="cuda", dtype=torch.bfloat16)   # 1 GiB
y = torch.empty_like(x)
xf = x.view(torch.float32)
for name, fn, by in (("sum bf16 (read 1 GiB)", lambda: x.sum(), x.numel() * 2),
                     ("sum f32 view (read 1 GiB)", lambda: xf.sum(), x.numel() * 2),
                     ("max f32 view", lambda: xf.max(), x.numel() * 2),
                     ("copy (r+w 2 GiB)", lambda: y.copy_(x), x.numel() * 4),
                     ("fill (write 1 GiB)", lambda: y.zero_(), x.numel() * 2)):
    us = timeit(fn, n=5)
    print("%-28s %8.1f us  %6.0f GB/s" % (name, us, by / us / 1e3))
