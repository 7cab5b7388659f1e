import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200.optim import FusedAdamEMA
from kair_b200.network_unet import UNetRes
model = UNetRes(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", bias=False).cuda().train()
x = torch.rand(16, 4, 128, 128).cuda(); t = torch.rand(16, 3, 128, 128).cuda()
model(x)
eng = model.engine(x.device)
orig = eng.refresh_weights
marks = []
def timed_refresh(plan, st):
    a, b = torch.cuda.Event(True), torch.cuda.Event(True)
    a.record(); orig(plan, st); b.record()
    marks.append((a, b))
eng.refresh_weights = timed_refresh
for kind in ("fused", "torch", "fused"):
    opt = FusedAdamEMA(model.parameters(), lr=1e-4) if kind == "fused" else torch.optim.Adam(model.parameters(), lr=1e-4, fused=True)
    def step():
        opt.zero_grad(set_to_none=True)
        loss = torch.mean(torch.abs(model(x) - t))
        loss.backward()
        opt.step()
    for _ in range(5): step()
    torch.cuda.synchronize(); marks.clear()
    for i in range(20): step()
    torch.cuda.synchronize()
    print("%-6s refresh (pack) %.3f ms/step over %d calls" % (kind, sum(a.elapsed_time(b) for a, b in marks) / 20, len(marks)), flush=True)
    print("   versions sample:", [p._version for p in list(model.parameters())[:3]])
