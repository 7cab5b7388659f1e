import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200.optim import FusedAdamEMA
from kair_b200.network_unet import UNetRes
from kair_b200.network_swinir import SwinIR
for name, mk in [("drunet", lambda: UNetRes(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", bias=False)),
                 ("swinir-m", lambda: SwinIR(upscale=1, in_chans=1, img_size=128, window_size=8, img_range=1.0, depths=[6] * 6,
                                             embed_dim=180, num_heads=[6] * 6, mlp_ratio=2, upsampler=""))]:
    for kind in ("torch", "fused", "fused+ema"):
        model = mk().cuda()
        ps = list(model.parameters())
        flat = torch.randn(sum(p.numel() for p in ps), device="cuda") * 0.01
        off = 0
        for p in ps:
            p.grad = flat[off:off + p.numel()].view(p.shape); off += p.numel()
        if kind == "torch":
            opt = torch.optim.Adam(ps, lr=1e-4, fused=True)
        else:
            ema = [p.detach().clone() for p in ps] if kind.endswith("ema") else None
            opt = FusedAdamEMA(ps, lr=1e-4, ema_params=ema, ema_decay=0.999 if ema else 0.0)
        for _ in range(3): opt.step()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(True), torch.cuda.Event(True)
        t0 = time.perf_counter(); a.record()
        for _ in range(20): opt.step()
        b.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
        print("%-9s %-10s gpu %.1f us/step  host %.1f us/step" % (name, kind, a.elapsed_time(b) * 50, (t1 - t0) * 5e4), flush=True)
