import os, sys, torch, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200 import profiling
from kair_b200.network_unet import UNetRes
from kair_b200.network_swinir import SwinIR
for name, model, x in [("drunet", UNetRes(in_nc=4, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", bias=False), torch.rand(2, 4, 64, 64)),
                       ("swinir-m", SwinIR(upscale=1, in_chans=1, img_size=128, window_size=8, img_range=1.0, depths=[6] * 6, embed_dim=180, num_heads=[6] * 6, mlp_ratio=2, upsampler=""), torch.rand(1, 1, 64, 64))]:
    model = model.cuda().train()
    y = model(x.cuda()); y.mean().backward()
    eng = model.engine(torch.device("cuda:0"))
    plan = next(iter(eng._plans.values()))
    packs = plan.Wt["pack"]
    sth = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    per = profiling.time_launches(packs, sth, reps=5)
    print(name, "pack launches:", len(packs), "total %.1f us" % (1e3 * sum(per)))
    for l, t in zip(packs, per):
        a = l.args
        extra = ""
        if l.name == "kair_pack_conv": extra = "L=%d O_pad=%d I_pad=%d taps=%d" % (a.L, a.O_pad, a.I_pad, a.taps)
        print("   %-20s %7.1f us  %s" % (l.name, t * 1e3, extra))
