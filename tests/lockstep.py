"""TEST INFRASTRUCTURE: run an engine plan on the GPU kernels and on tests/emu_ops.py in lock-step, launch by launch,
and report the first launch whose written tensors diverge.  Used by the GPU tests and as a debugging tool:

    python tests/lockstep.py [B H W in_chans]
"""
import sys

import torch


class Recorder:
    """Proxy over an ops module that remembers, per built launch, every tensor argument (name -> tensor)."""

    def __init__(self, ops):
        self._ops = ops
        self.records = []

    def __getattr__(self, name):
        attr = getattr(self._ops, name)
        if not name.startswith("build_"):
            return attr

        def build(*args, **kwargs):
            launch = attr(*args, **kwargs)
            tensors = {}
            for i, a in enumerate(args):
                if isinstance(a, torch.Tensor):
                    tensors["arg%d" % i] = a
                elif isinstance(a, (list, tuple)) and a and isinstance(a[0], torch.Tensor):
                    for j, t in enumerate(a):
                        tensors["arg%d[%d]" % (i, j)] = t
            for k, a in kwargs.items():
                if isinstance(a, torch.Tensor):
                    tensors[k] = a
            desc = name[6:] + "(" + ",".join("%s=%s" % (k, v) for k, v in kwargs.items()
                                             if isinstance(v, (int, float, tuple))) + ")"
            self.records.append((launch, tensors, desc))
            return launch

        return build


def _cmp(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    if a.shape != b.shape:
        return float("inf"), float("inf")
    d = (a - b)
    return d.abs().max().item(), (d.norm() / (b.norm() + 1e-20)).item()


@torch.no_grad()
def run_lockstep(model_cpu, x, dout, drop=None, rel_tol=2e-2, verbose=False):
    """model_cpu: kair_b200.network_swinir.SwinIR on CPU.  Returns a list of (launch index, phase, description,
    tensor name, max-abs, rel-l2) for every tensor that diverged right after the launch that (may have) written it."""
    import copy

    import emu_ops
    from kair_b200 import ops as real_ops
    from kair_b200.swinir_engine import SwinIREngine

    dev = torch.device("cuda:0")
    model_gpu = copy.deepcopy(model_cpu).to(dev)
    rec_e, rec_g = Recorder(emu_ops), Recorder(real_ops)
    eng_e = SwinIREngine(model_cpu._spec, dict(model_cpu.named_parameters()), "cpu", ops=rec_e)
    eng_g = SwinIREngine(model_gpu._spec, dict(model_gpu.named_parameters()), dev, ops=rec_g)
    B, _, H, W = x.shape
    plan_e = eng_e.plan(B, H, W, True)
    plan_g = eng_g.plan(B, H, W, True)
    if drop is not None:
        plan_e.dp.copy_(drop)
        plan_g.dp.copy_(drop.to(dev))
    assert len(rec_e.records) == len(rec_g.records)
    index = {id(l): i for i, (l, _, _) in enumerate(rec_g.records)}
    index_e = {id(l): i for i, (l, _, _) in enumerate(rec_e.records)}
    st = real_ops.stream_handle()
    report = []

    def step(phase, le, lg):
        i = index[id(lg)]
        assert index_e[id(le)] == i
        le.run(None)
        lg.run(st)
        torch.cuda.synchronize()
        _, te, desc = rec_e.records[i]
        _, tg, _ = rec_g.records[i]
        for k in te:
            if te[k].dtype in (torch.int32, torch.int64):
                continue
            mx, rel = _cmp(tg[k], te[k])
            if verbose:
                print("%4d %-4s %-60s %-12s max %.3e rel %.3e" % (i, phase, desc[:60], k, mx, rel))
            if not rel < rel_tol:
                report.append((i, phase, desc, k, mx, rel))
                # re-synchronise so that one divergence does not cascade: copy the emulation's value over
                tg[k].copy_(te[k].to(dev))

    for le, lg in zip(eng_e._weights["pack"], eng_g._weights["pack"]):
        step("pack", le, lg)
    eng_e._weights["versions"] = eng_e._param_versions()
    eng_g._weights["versions"] = eng_g._param_versions()
    plan_e.img.copy_(x)
    plan_g.img.copy_(x.to(dev))
    for le, lg in zip(plan_e.fwd, plan_g.fwd):
        step("fwd", le, lg)
    plan_e.dout.copy_(dout)
    plan_g.dout.copy_(dout.to(dev))
    for p in (plan_e, plan_g):
        p.gflat.zero_()
        p.dwflat.zero_()
    for le, lg in zip(plan_e.bwd, plan_g.bwd):
        step("bwd", le, lg)
    return report


def _main():
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here)
    sys.path.insert(0, os.path.dirname(here))
    from kair_b200.network_swinir import SwinIR
    a = [int(v) for v in sys.argv[1:] if not v.startswith("-")]
    B, H, W, ci = (a + [2, 32, 32, 1][len(a):])[:4]
    torch.manual_seed(0)
    model = SwinIR(upscale=1, in_chans=ci, img_size=32, window_size=8, img_range=1.0, depths=[2, 2], embed_dim=180,
                   num_heads=[6, 6], mlp_ratio=2, upsampler="", resi_connection="1conv", drop_path_rate=0.0)
    g = torch.Generator().manual_seed(11)
    with torch.no_grad():
        for k, p in model.named_parameters():
            if "norm" in k and k.endswith("weight"):
                p.copy_(1.0 + 0.2 * torch.randn(p.shape, generator=g))
            elif k.endswith("bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif k.endswith("relative_position_bias_table"):
                p.copy_(0.3 * torch.randn(p.shape, generator=g))
            else:
                p.mul_(1.5)
    x = torch.rand(B, ci, H, W, generator=g)
    dout = torch.randn(B, ci, H, W, generator=g) / (B * H * W)
    rep = run_lockstep(model, x, dout, verbose="-v" in sys.argv)
    for r in rep:
        print("DIVERGED launch %d [%s] %s : %s max %.3e rel %.3e" % r)
    print("%d divergent tensors" % len(rep))


if __name__ == "__main__":
    _main()
