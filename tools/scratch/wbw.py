import torch
x = torch.empty(1 << 30, dtype=torch.bfloat16, device="cuda")
y = torch.empty(1 << 30, dtype=torch.bfloat16, device="cuda")
def t(f, n=10):
    f(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        a.record(); f(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best
ms = t(lambda: x.zero_()); print("fill  %.1f GB/s" % (2 * (1 << 30) / ms / 1e6))
ms = t(lambda: y.copy_(x)); print("copy  %.1f GB/s (r+w)" % (4 * (1 << 30) / ms / 1e6))
ms = t(lambda: x.sum()); print("read  %.1f GB/s" % (2 * (1 << 30) / ms / 1e6))
xs = x[: 100 << 20]
ms = t(lambda: xs.zero_()); print("fill 200MB %.1f GB/s" % (2 * (100 << 20) / ms / 1e6))
