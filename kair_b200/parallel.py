"""Data-parallel training over NCCL / NVLink for the drop-in modules.

The reference shards training by patch batch with `DistributedDataParallel` (models/model_base.py:113-119); that keeps
working unchanged around `kair_b200` modules.  Because every engine produces ALL parameter gradients of a step in one
flat fp32 buffer, the exchange can also be done as ONE all-reduce of that buffer issued from inside the backward pass
(46 MB for SwinIR-M, 131 MB for DRUNet), instead of DDP's per-parameter copy into bucket views (490 tiny copy kernels
per step for SwinIR-M, ~2 ms measured on B200) followed by bucketed all-reduces.  `FlatDataParallel` enables that:

    net = FlatDataParallel(model)            # instead of DistributedDataParallel(model, device_ids=[rank])
    loss = criterion(net(x), y); loss.backward(); optimizer.step()

Semantics are DDP's: parameters are broadcast from rank 0 at construction, gradients are averaged over the group,
buffers are not broadcast (the kernels never read the mask / index buffers; BatchNorm stays per-GPU like the
reference's plain BatchNorm2d).  There is no model-parallel exchange anywhere on this path.
"""
import torch
import torch.distributed as dist
import torch.nn as nn


class FlatDataParallel(nn.Module):
    def __init__(self, module, process_group=None):
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

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)


def allreduce_flat(module, flat):
    """Average `flat` (the engine's gradient buffer) over the module's data-parallel group, if it has one."""
    dp = getattr(module, "_dp_group", None)
    if dp is None:
        return
    group, world = dp
    if world > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.mul_(1.0 / world)
