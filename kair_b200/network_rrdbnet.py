"""Drop-in RRDBNet_noUp (models/network_rrdbnet.py:35-125 of the reference): same constructor, forward(x) contract and
state_dict layout (conv_first, RRDB_trunk.{i}.RDB{1,2,3}.conv{1..5}, trunk_conv, HRconv, conv_last), so `define_G` with
net_type 'rrdbnet_no_up' (models/select_network.py) and existing checkpoints work unchanged.  The module tree is a
parameter container of stock nn.Conv2d children; forward hands it to kair_b200.rrdb_engine (dense blocks without
torch.cat, implicit-GEMM 3x3 convolutions on tcgen05, forward and backward)."""
import functools

import torch
import torch.nn as nn
import torch.nn.init as init

from . import parallel as _parallel
from .convnet_engine import ConvNetFunction
from .rrdb_engine import RRDBEngine, parse_rrdbnet


def initialize_weights(net_l, scale=1):
    """models/network_rrdbnet.py:8-25"""
    if not isinstance(net_l, list):
        net_l = [net_l]
    for net in net_l:
        for m in net.modules():
            if isinstance(m, nn.Conv2d):
                init.kaiming_normal_(m.weight, a=0, mode='fan_in')
                m.weight.data *= scale
                if m.bias is not None:
                    m.bias.data.zero_()


def _no_forward(self, *a, **k):
    raise RuntimeError("kair_b200: sub-modules are parameter containers; call the network's forward")


class ResidualDenseBlock_5C(nn.Module):
    def __init__(self, nf=64, gc=32, bias=True):
        super().__init__()
        self.conv1 = nn.Conv2d(nf, gc, 3, 1, 1, bias=bias)
        self.conv2 = nn.Conv2d(nf + gc, gc, 3, 1, 1, bias=bias)
        self.conv3 = nn.Conv2d(nf + 2 * gc, gc, 3, 1, 1, bias=bias)
        self.conv4 = nn.Conv2d(nf + 3 * gc, gc, 3, 1, 1, bias=bias)
        self.conv5 = nn.Conv2d(nf + 4 * gc, nf, 3, 1, 1, bias=bias)
        self.lrelu = nn.LeakyReLU(negative_slope=0.2, inplace=True)
        initialize_weights([self.conv1, self.conv2, self.conv3, self.conv4, self.conv5], 0.1)

    forward = _no_forward


class RRDB(nn.Module):
    def __init__(self, nf, gc=32):
        super().__init__()
        self.RDB1 = ResidualDenseBlock_5C(nf, gc)
        self.RDB2 = ResidualDenseBlock_5C(nf, gc)
        self.RDB3 = ResidualDenseBlock_5C(nf, gc)

    forward = _no_forward


class RRDBNet_noUp(nn.Module):
    def __init__(self, in_nc=3, out_nc=3, nf=64, nb=23, gc=32):
        super().__init__()
        block = functools.partial(RRDB, nf=nf, gc=gc)
        self.conv_first = nn.Conv2d(in_nc, nf, 3, 1, 1, bias=True)
        self.RRDB_trunk = nn.Sequential(*[block() for _ in range(nb)])
        self.trunk_conv = nn.Conv2d(nf, nf, 3, 1, 1, bias=True)
        self.HRconv = nn.Conv2d(nf, nf, 3, 1, 1, bias=True)
        self.conv_last = nn.Conv2d(nf, out_nc, 3, 1, 1, bias=True)
        self.lrelu = nn.LeakyReLU(negative_slope=0.2, inplace=True)
        if nf != 64 or gc != 32:
            raise NotImplementedError("kair_b200 RRDBNet_noUp: nf=64, gc=32 (the reference's and the fork's setting)")
        self._engine = None
        self._engine_key = None
        self._ops = None   # test hook: tests/emu_ops.py replays the launch lists on CPU

    def engine(self, device):
        params = dict(self.named_parameters())
        key = (str(device),) + tuple(p.data_ptr() for p in params.values())
        if self._engine is None or self._engine_key != key:
            children = dict(conv_first=self.conv_first, trunk_conv=self.trunk_conv, HRconv=self.HRconv,
                            conv_last=self.conv_last)
            self._engine = RRDBEngine(children, parse_rrdbnet(self), params, device, ops=self._ops)
            self._engine_key = key
        self._engine.grad_hook = (lambda flat: _parallel.allreduce_flat(self, flat))
        return self._engine

    def forward(self, x):
        if not x.is_cuda and self._ops is None:
            raise RuntimeError("kair_b200 RRDBNet_noUp runs on sm_100a GPUs only; got a %s tensor (there is no CPU "
                               "fallback)" % x.device.type)
        x = x.float().contiguous()
        eng = self.engine(x.device)
        params = list(eng.P.values())
        train = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in params))
        plan = eng.plan(x.shape[0], x.shape[2], x.shape[3], train)
        if train:
            return ConvNetFunction.apply(x, plan, *params)
        return plan.run_forward(x)
