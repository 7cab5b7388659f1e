import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200.optim import FusedAdamEMA
from kair_b200.network_unet import UNetRes
model = UNetRes(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", bias=False).cuda().train()
x = torch.rand(16, 4, 128, 128).cuda(); t = torch.rand(16, 3, 128, 128).cuda()
for kind in ("fused", "torch"):
    opt = FusedAdamEMA(model.parameters(), lr=1e-4) if kind == "fused" else torch.optim.Adam(model.parameters(), lr=1e-4, fused=True)
    ev = [[torch.cuda.Event(True) for _ in range(4)] for _ in range(20)]
    def step(e=None):
        opt.zero_grad(set_to_none=True)
        if e: e[0].record()
        out = model(x)
        loss = torch.mean(torch.abs(out - t))
        if e: e[1].record()
        loss.backward()
        if e: e[2].record()
        opt.step()
        if e: e[3].record()
    for _ in range(5): step()
    torch.cuda.synchronize()
    for i in range(20): step(ev[i])
    torch.cuda.synchronize()
    f = sum(e[0].elapsed_time(e[1]) for e in ev) / 20
    b = sum(e[1].elapsed_time(e[2]) for e in ev) / 20
    o = sum(e[2].elapsed_time(e[3]) for e in ev) / 20
    g = sum(ev[i][3].elapsed_time(ev[i + 1][0]) for i in range(19)) / 19
    print("%-6s fwd+loss %.3f  bwd %.3f  opt %.3f  gap-to-next %.3f ms" % (kind, f, b, o, g), flush=True)
