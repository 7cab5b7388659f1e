import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200 import ops
T = 131072
dev = torch.device("cuda:0")
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).bfloat16()
for (n1, n2, bn) in [(576, 192, 192), (192, 384, 192), (384, 192, 192), (192, 192, 192)]:
    A, B, D = bf(T, n1), bf(T, n2), torch.zeros(n1, n2, device=dev)
    f = lambda: ops.gemm_nt(A, B, D, n1=n1, n2=n2, bn=bn)
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record()
    for _ in range(10): f()
    b.record(); torch.cuda.synchronize()
    print("gemm_nt n1=%d n2=%d bn=%d: %.1f us" % (n1, n2, bn, a.elapsed_time(b) * 100), flush=True)
