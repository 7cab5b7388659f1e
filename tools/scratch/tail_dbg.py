import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200 import ops
from kair_b200.optim import FusedAdamEMA
from kair_b200.loss import L1Loss
from kair_b200.network_unet import UNetRes
cnt = [0]
orig = ops.adam_ema_set_grads
def counted(l, g, t):
    cnt[0] += 1
    return orig(l, g, t)
ops.adam_ema_set_grads = counted
model = UNetRes(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", bias=False).cuda().train()
x = torch.rand(16, 4, 128, 128).cuda(); t = torch.rand(16, 3, 128, 128).cuda()
for kind in ("fused", "torch", "fused-loss+torch-adam", "torch-loss+fused-adam"):
    opt = FusedAdamEMA(model.parameters(), lr=1e-4) if kind in ("fused", "torch-loss+fused-adam") else torch.optim.Adam(model.parameters(), lr=1e-4, fused=True)
    l1 = L1Loss() if kind.startswith("fused") else (lambda a, b: torch.mean(torch.abs(a - b)))
    def step():
        opt.zero_grad(set_to_none=True)
        loss = l1(model(x), t)
        loss.backward()
        opt.step()
    for _ in range(5): step()
    torch.cuda.synchronize(); cnt[0] = 0
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    t0 = time.perf_counter(); a.record()
    for _ in range(20): step()
    b.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
    print("%-24s gpu %.3f ms/step host %.3f ms/step  set_grads calls %d" % (kind, a.elapsed_time(b) / 20, (t1 - t0) * 50, cnt[0]), flush=True)
