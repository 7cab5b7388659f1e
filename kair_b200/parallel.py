"""Data-parallel training over NCCL / NVLink for the drop-in modules.

The reference shards training by patch batch with `DistributedDataParallel` (models/model_base.py:113-119); that keeps
working unchanged around `kair_b200` modules.  Because every engine produces ALL parameter gradients of a step in one
flat fp32 buffer, the exchange can also be done as ONE all-reduce of that buffer issued from inside the backward pass
(46 MB for SwinIR-M, 131 MB for DRUNet), instead of DDP's per-parameter copy into bucket views (490 tiny copy kernels
per step for SwinIR-M, ~2 ms measured on B200) followed by bucketed all-reduces.  `FlatDataParallel` enables that:

    net = FlatDataParallel(model)            # instead of DistributedDataParallel(model, device_ids=[rank])
    loss = criterion(net(x), y); loss.backward(); optimizer.step()

Semantics are DDP's: parameters and buffers are broadcast from rank 0 once, at construction; gradients are averaged
over the group; buffers are NOT re-broadcast every step (DDP's broadcast_buffers=False: the kernels never read the
mask / index buffers, and BatchNorm statistics stay per-GPU like the reference's plain BatchNorm2d).  There is no
model-parallel exchange anywhere on this path.

Overlap with backward (model_base.py:113-119: DDP overlaps its bucket all-reduces with the backward pass): the SwinIR
engine marks the flat-buffer slice of an RSTB as final as soon as that RSTB's backward is done (the backward runs
from the last RSTB to the first), and `allreduce_slice` starts an asynchronous all-reduce of that slice on the process
group's own stream while the remaining backward kernels keep running; `allreduce_flat` at the end of the pass waits
for those, reduces what no slice covered (head / tail parameters, a few hundred KB) and applies the 1/world scale.
"""
import torch
import torch.distributed as dist
import torch.nn as nn


class FlatDataParallel(nn.Module):
    def __init__(self, module, process_group=None, overlap=True):
        """overlap=False: one all-reduce of the whole buffer after the backward pass (the round-1 behaviour)"""
        super().__init__()
        if not dist.is_available() or not dist.is_initialized():
            raise RuntimeError("FlatDataParallel needs an initialised torch.distributed process group")
        if not hasattr(module, "engine"):
            raise TypeError("FlatDataParallel wraps kair_b200 SwinIR / DnCNN / UNetRes modules")
        self.module = module
        self.process_group = process_group
        with torch.no_grad():
            for t in list(module.parameters()) + list(module.buffers()):
                dist.broadcast(t, 0, group=process_group)
        module._dp_group = (process_group, dist.get_world_size(process_group))
        module._dp_overlap = bool(overlap)

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)


def allreduce_slice(module, flat, lo, hi):
    """Start the all-reduce of flat[lo:hi] (final from here on) without waiting for it; no-op without a group."""
    dp = getattr(module, "_dp_group", None)
    if dp is None or dp[1] <= 1 or not getattr(module, "_dp_overlap", True):
        return
    work = dist.all_reduce(flat[lo:hi], op=dist.ReduceOp.SUM, group=dp[0], async_op=True)
    module.__dict__.setdefault("_dp_pending", []).append((lo, hi, work))


def allreduce_flat(module, flat):
    """Average `flat` (the engine's gradient buffer) over the module's data-parallel group, if it has one: waits for
    the slices already in flight (allreduce_slice), reduces the ranges no slice covered, scales by 1/world."""
    dp = getattr(module, "_dp_group", None)
    if dp is None:
        return
    group, world = dp
    if world <= 1:
        return
    pending = module.__dict__.pop("_dp_pending", [])
    covered = sorted((lo, hi) for lo, hi, _ in pending)
    pos, rest = 0, []
    for lo, hi in covered:
        if lo > pos:
            rest.append((pos, lo))
        pos = max(pos, hi)
    if pos < flat.numel():
        rest.append((pos, flat.numel()))
    works = [w for _, _, w in pending]
    for lo, hi in rest:
        works.append(dist.all_reduce(flat[lo:hi], op=dist.ReduceOp.SUM, group=group, async_op=True))
    for w in works:
        w.wait()
    flat.mul_(1.0 / world)
