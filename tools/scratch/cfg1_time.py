"""cfg1 (DnCNN-17, 1x1x256x256): each forward launch timed alone as 40 back-to-back copies inside one CUDA graph."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from kair_b200.network_dncnn import DnCNN

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
model = DnCNN(in_nc=1, out_nc=1, nc=64, nb=17, act_mode="R").cuda().eval()
x = torch.rand(B, 1, 256, 256, device="cuda")
with torch.no_grad():
    for _ in range(2):
        model(x)
eng = model.engine(torch.device("cuda:0"))
plan = next(iter(eng._plans.values()))
REP = 40
if len(sys.argv) > 2 and sys.argv[2] == "once":   # for ncu: every forward launch once, eagerly
    for l in plan.fwd:
        l.run(eng.ops.stream_handle())
    torch.cuda.synchronize()
    sys.exit(0)
side = torch.cuda.Stream()
for idx in (0, 1, 8, len(plan.fwd) - 1):
    l = plan.fwd[idx]
    with torch.cuda.stream(side):
        st = eng.ops.stream_handle()
        l.run(st)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            st = eng.ops.stream_handle()
            for _ in range(REP):
                l.run(st)
        g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record()
        e1.synchronize()
    print("%2d %-22s %7.2f us" % (idx, l.name, e0.elapsed_time(e1) * 1e3 / (5 * REP)))
