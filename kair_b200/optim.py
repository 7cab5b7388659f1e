"""`FusedAdamEMA`: torch.optim.Adam + the EMA of `ModelBase.update_E` (models/model_base.py:247-252 of the reference) as
ONE kernel launch per parameter group (`kair_adam_ema`, kair_b200/csrc/train_step.cu).  SURVEY.md section 8(f1).

Drop-in for the optimizer `ModelPlain.define_optimizer` builds (models/model_plain.py:139-152): same constructor
arguments, `param_groups` keys and `state_dict()` layout as `torch.optim.Adam` (state[p] = {step, exp_avg, exp_avg_sq}), so
`save_optimizer` / `load_optimizer` (model_base.py:218-238) and the lr schedulers keep working.  With `ema_params` (the
parameters of netE, in the order of netG's) and `ema_decay` > 0 the call that `optimize_parameters` makes to
`update_E` (model_plain.py:317-318) is folded into `step()`: e = e*decay + p_new*(1-decay).  There is no CPU fallback."""
import math

import torch
from torch.autograd.graph import increment_version

from . import ops as _real_ops


class FusedAdamEMA(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, *,
                 maximize=False, ema_params=None, ema_decay=0.0, ops=None):
        if amsgrad or maximize:
            raise NotImplementedError("kair_b200 FusedAdamEMA: amsgrad / maximize are not implemented (KAIR uses neither)")
        if not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0 or lr < 0 or eps < 0 or weight_decay < 0:
            raise ValueError("invalid Adam hyper-parameters")
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                        foreach=None, capturable=False, differentiable=False, fused=None)
        super().__init__(params, defaults)
        self._ops = ops or _real_ops
        self.ema_decay = float(ema_decay)
        self._ema = None
        if ema_params is not None and self.ema_decay > 0:
            ema = [e for e in ema_params]
            flat = [p for g in self.param_groups for p in g["params"]]
            if len(ema) != len(flat) or any(e.shape != p.shape for e, p in zip(ema, flat)):
                raise ValueError("ema_params must match the optimised parameters one to one (netE vs netG order)")
            self._ema = {id(p): e for p, e in zip(flat, ema)}
        self._launches = {}   # group index -> list of recorded sub-launches (see _build)

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._launches.clear()   # the moment tensors were replaced

    def _build(self, gi, group, ps):
        """Slow path (first step, or the set / storage of the tensors changed): create the optimizer state, split the
        parameters by step count (they all agree in KAIR's loop: every parameter gets a gradient every step) and record
        one launch per distinct count."""
        o = self._ops
        for p in ps:
            if p.grad.is_sparse or p.dtype != torch.float32:
                raise RuntimeError("kair_b200 FusedAdamEMA: dense fp32 parameters only")
            st = self.state[p]
            if len(st) == 0:
                st["step"] = torch.tensor(0.0, dtype=torch.float32)
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        b1, b2 = group["betas"]
        by_step = {}
        for p in ps:
            by_step.setdefault(float(self.state[p]["step"]), []).append(p)
        subs = []
        for step0, plist in by_step.items():
            gs = [p.grad if p.grad.is_contiguous() else p.grad.contiguous() for p in plist]
            ms = [self.state[p]["exp_avg"] for p in plist]
            vs = [self.state[p]["exp_avg_sq"] for p in plist]
            es = [self._ema[id(p)] for p in plist] if self._ema is not None else None
            launch = o.build_adam_ema([p.data for p in plist], gs, ms, vs, [e.data for e in es] if es else None,
                                      beta1=b1, beta2=b2, eps=group["eps"], weight_decay=group["weight_decay"],
                                      ema_decay=self.ema_decay if es else 0.0)
            subs.append(dict(launch=launch, params=plist, t=step0, gkey=tuple(g.data_ptr() for g in gs), gtables={},
                             steps=[self.state[p]["step"] for p in plist], bump=plist + ms + vs + (es or [])))
        ent = dict(ids=tuple(map(id, ps)), pptr=[p.data_ptr() for p in ps], subs=subs,
                   hyper=(b1, b2, group["eps"], group["weight_decay"]))
        self._launches[gi] = ent
        return ent

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        o = self._ops
        for gi, group in enumerate(self.param_groups):
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            b1, b2 = group["betas"]
            ent = self._launches.get(gi)
            if (ent is None or ent["ids"] != tuple(map(id, ps)) or len(ent["subs"]) != 1
                    or ent["hyper"] != (b1, b2, group["eps"], group["weight_decay"])
                    or any(p.data_ptr() != q for p, q in zip(ps, ent["pptr"]))):
                ent = self._build(gi, group, ps)
            for sub in ent["subs"]:
                gs = [p.grad if p.grad.is_contiguous() else p.grad.contiguous() for p in sub["params"]]
                gkey = tuple(g.data_ptr() for g in gs)
                if gkey != sub["gkey"]:
                    # autograd handed out new gradient tensors: refresh that one pointer table (stream-ordered)
                    o.adam_ema_set_grads(sub["launch"], gs, sub["gtables"])
                    sub["gkey"] = gkey
                    sub["grads"] = gs   # the table points into these
                sub["t"] += 1.0
                l = sub["launch"]
                l.args.step_size = group["lr"] / (1.0 - b1 ** sub["t"])
                l.args.bc2_sqrt = math.sqrt(1.0 - b2 ** sub["t"])
                l.run(o.stream_handle())
                # the kernel wrote through raw pointers: tell autograd (and the engines, which re-pack their bf16
                # operands when a parameter's version changes) that these tensors were modified in place
                increment_version(sub["bump"])
                torch._foreach_add_(sub["steps"], 1.0)
        return loss


@torch.no_grad()
def update_ema(ema_params, params, decay):
    """Stand-alone `update_E` (models/model_base.py:247-252) for callers that keep their own optimizer:
    e = e*decay + p*(1-decay) over all tensors with two multi-tensor (`torch._foreach`) calls instead of 2 launches per
    parameter.  `FusedAdamEMA(ema_params=...)` folds the same arithmetic into the optimizer kernel instead."""
    ema, ps = [e.data for e in ema_params], [p.data for p in params]
    torch._foreach_mul_(ema, decay)
    torch._foreach_add_(ema, ps, alpha=1.0 - decay)
