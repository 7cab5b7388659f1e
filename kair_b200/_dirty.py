"""When do the engines have to re-pack their bf16 operands from the fp32 master parameters?

Tensor `_version` counters catch ordinary in-place updates (load_state_dict, init functions, foreach optimizers, update_E),
but fused optimizers -- `torch.optim.Adam(fused=True)`, apex -- update parameters without touching the counters.  A
global optimizer post-step hook therefore advances an epoch that is part of every engine's change key: any
`optimizer.step()` anywhere makes the next forward of every engine re-pack (a few hundred microseconds at most)."""
import torch

_EPOCH = [0]


def _after_step(optimizer, args, kwargs):
    _EPOCH[0] += 1


try:
    from torch.optim.optimizer import register_optimizer_step_post_hook
    register_optimizer_step_post_hook(_after_step)
except ImportError:  # pragma: no cover - torch < 2.0 has no global hook; version counters only
    pass


def optimizer_epoch():
    return _EPOCH[0]
