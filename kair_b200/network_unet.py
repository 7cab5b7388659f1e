"""Drop-in `UNetRes` (DRUNet) for models/network_unet.py:21-76 of the reference, computed by the sm_100a kernels.

Constructor arguments, forward(x0) contract and state_dict layout (m_head, m_down1..3, m_body, m_up3..1, m_tail;
ResBlock children under `.res`) are the reference's, so DRUNet checkpoints, `init_weights` (orthogonal,
select_network.py:385-433) and main_train_drunet.py work unchanged.  Only 'strideconv' / 'convtranspose' resampling
(the DRUNet configuration, options/train_drunet.json) has kernels; the pooling / pixelshuffle / upconv variants raise.
There is no CPU / PyTorch fallback.
"""
import torch
import torch.nn as nn

from . import parallel as _parallel

from . import basicblock as B
from .convnet_engine import ConvNetFunction, UNetResEngine, parse_unetres
from .network_dncnn import _tree_key


class UNetRes(nn.Module):
    def __init__(self, in_nc=3, out_nc=3, nc=[64, 128, 256, 512], nb=4, act_mode="R", downsample_mode="strideconv",
                 upsample_mode="convtranspose", bias=True):
        super().__init__()
        if downsample_mode != "strideconv":
            if downsample_mode in ("avgpool", "maxpool"):
                raise NotImplementedError("kair_b200 UNetRes: downsample_mode '%s' has no sm_100a kernel (DRUNet uses "
                                          "'strideconv')" % downsample_mode)
            raise NotImplementedError("downsample mode [{:s}] is not found".format(downsample_mode))
        if upsample_mode != "convtranspose":
            if upsample_mode in ("upconv", "pixelshuffle"):
                raise NotImplementedError("kair_b200 UNetRes: upsample_mode '%s' has no sm_100a kernel (DRUNet uses "
                                          "'convtranspose')" % upsample_mode)
            raise NotImplementedError("upsample mode [{:s}] is not found".format(upsample_mode))
        res = lambda c: B.ResBlock(c, c, bias=bias, mode="C" + act_mode + "C")
        self.m_head = B.conv(in_nc, nc[0], bias=bias, mode="C")
        self.m_down1 = B.sequential(*[res(nc[0]) for _ in range(nb)], B.downsample_strideconv(nc[0], nc[1], bias=bias, mode="2"))
        self.m_down2 = B.sequential(*[res(nc[1]) for _ in range(nb)], B.downsample_strideconv(nc[1], nc[2], bias=bias, mode="2"))
        self.m_down3 = B.sequential(*[res(nc[2]) for _ in range(nb)], B.downsample_strideconv(nc[2], nc[3], bias=bias, mode="2"))
        self.m_body = B.sequential(*[res(nc[3]) for _ in range(nb)])
        self.m_up3 = B.sequential(B.upsample_convtranspose(nc[3], nc[2], bias=bias, mode="2"), *[res(nc[2]) for _ in range(nb)])
        self.m_up2 = B.sequential(B.upsample_convtranspose(nc[2], nc[1], bias=bias, mode="2"), *[res(nc[1]) for _ in range(nb)])
        self.m_up1 = B.sequential(B.upsample_convtranspose(nc[1], nc[0], bias=bias, mode="2"), *[res(nc[0]) for _ in range(nb)])
        self.m_tail = B.conv(nc[0], out_nc, bias=bias, mode="C")
        self._engine = None
        self._engine_key = None
        self._ops = None   # test hook: tests/emu_ops.py replays the launch lists on CPU (host-logic / DDP tests)

    def engine(self, device):
        key = _tree_key(self, device)
        if self._engine is None or self._engine_key != key:
            self._engine = UNetResEngine(parse_unetres(self), dict(self.named_parameters()), device, ops=self._ops)
            self._engine_key = key
        self._engine.grad_hook = (lambda flat: _parallel.allreduce_flat(self, flat))
        return self._engine

    def forward(self, x0):
        if not x0.is_cuda and self._ops is None:
            raise RuntimeError("kair_b200 UNetRes runs on sm_100a GPUs only; got a %s tensor (there is no CPU fallback)"
                               % x0.device.type)
        x0 = x0.float().contiguous()
        eng = self.engine(x0.device)
        params = list(eng.P.values())
        train = torch.is_grad_enabled() and (x0.requires_grad or any(p.requires_grad for p in params))
        plan = eng.plan(x0.shape[0], x0.shape[2], x0.shape[3], train)
        if train:
            return ConvNetFunction.apply(x0, plan, *params)
        return plan.run_forward(x0)
