"""Drop-in `DnCNN` (models/network_dncnn.py:40-71 of the reference) and `FDnCNN` (:126-148), computed by the sm_100a
kernels.

Constructor, forward(x) contract (x - model(x)) and state_dict layout are the reference's: `self.model` is the
flattened nn.Sequential of stock Conv2d / BatchNorm2d / ReLU children built by `basicblock.conv`, so
`model_zoo/dncnn_25.pth` loads with strict=True into `DnCNN(act_mode='R')` (main_test_dncnn.py:119-121),
`init_weights` (select_network.py:385-433) and `merge_bn` (utils_bnorm.py:32-63; main_train_dncnn.py:177-182) keep
working.  The children never run: forward() re-reads the tree (it may have been mutated by merge_bn) and hands it to
`DnCNNEngine`.  There is no CPU / PyTorch fallback.
"""
import torch
import torch.nn as nn

from . import parallel as _parallel

from . import basicblock as B
from .convnet_engine import ConvNetFunction, DnCNNEngine, parse_dncnn


def _tree_key(module, device):
    """Identity of the parameter / buffer storage and of the module structure the engine was built for."""
    key = [str(device)]
    for name, m in module.named_modules():
        key.append((name, type(m).__name__))
    for t in list(module.parameters()) + list(module.buffers()):
        key.append(t.data_ptr())
    return tuple(key)


class DnCNN(nn.Module):
    def __init__(self, in_nc=1, out_nc=1, nc=64, nb=17, act_mode="BR"):
        super().__init__()
        assert "R" in act_mode or "L" in act_mode, "Examples of activation function: R, L, BR, BL, IR, IL"
        bias = True
        m_head = B.conv(in_nc, nc, mode="C" + act_mode[-1], bias=bias)
        m_body = [B.conv(nc, nc, mode="C" + act_mode, bias=bias) for _ in range(nb - 2)]
        m_tail = B.conv(nc, out_nc, mode="C", bias=bias)
        self.model = B.sequential(m_head, *m_body, m_tail)
        self._engine = None
        self._engine_key = None
        self._ops = None   # test hook: tests/emu_ops.py replays the launch lists on CPU (host-logic / DDP tests)

    _residual = True   # forward(x) = x - model(x)

    def engine(self, device):
        key = _tree_key(self, device)
        if self._engine is None or self._engine_key != key:
            self._engine = DnCNNEngine(parse_dncnn(self.model), dict(self.named_parameters()), device, ops=self._ops,
                                       residual=self._residual)
            self._engine_key = key
        self._engine.grad_hook = (lambda flat: _parallel.allreduce_flat(self, flat))
        return self._engine

    def forward(self, x):
        if not x.is_cuda and self._ops is None:
            raise RuntimeError("kair_b200 DnCNN runs on sm_100a GPUs only; got a %s tensor (there is no CPU fallback)"
                               % x.device.type)
        x = x.float().contiguous()
        eng = self.engine(x.device)
        params = list(eng.P.values())
        train = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in params))
        plan = eng.plan(x.shape[0], x.shape[2], x.shape[3], train, self.training)
        if train:
            return ConvNetFunction.apply(x, plan, *params)
        return plan.run_forward(x)


class FDnCNN(DnCNN):
    """models/network_dncnn.py:126-148: DnCNN with a noise-level map as extra input channel, no BatchNorm by default
    and no residual learning -- forward(x) = model(x).  Same engine, the tail writes n instead of x - n."""
    _residual = False

    def __init__(self, in_nc=2, out_nc=1, nc=64, nb=20, act_mode="R"):
        super().__init__(in_nc=in_nc, out_nc=out_nc, nc=nc, nb=nb, act_mode=act_mode)


class IRCNN(DnCNN):
    """models/network_dncnn.py:77-116: the 7-layer dilated-convolution denoiser prior (dilations 1, 2, 3, 4, 3, 2, 1,
    ReLU, no BatchNorm), forward(x) = x - model(x).  Same engine as DnCNN: the dilated layers are the implicit-GEMM 3x3
    with the taps `dilation` pixels apart (kair_gemm_tn / kair_gemm_nt `dilation`), forward and backward."""

    def __init__(self, in_nc=1, out_nc=1, nc=64):
        nn.Module.__init__(self)
        L = [nn.Conv2d(in_nc, nc, 3, 1, 1, dilation=1, bias=True), nn.ReLU(inplace=True)]
        for d in (2, 3, 4, 3, 2):
            L += [nn.Conv2d(nc, nc, 3, 1, d, dilation=d, bias=True), nn.ReLU(inplace=True)]
        L.append(nn.Conv2d(nc, out_nc, 3, 1, 1, dilation=1, bias=True))
        self.model = B.sequential(*L)
        self._engine = None
        self._engine_key = None
        self._ops = None
