"""Drop-in FFDNet (models/network_ffdnet.py:27-72 of the reference): same constructor, forward(x, sigma) contract and
state_dict layout (`model.{0,2,..}` of the flattened B.sequential), computed by the sm_100a kernels.

forward = ReplicationPad2d to even size -> PixelUnShuffle(2) -> cat(noise-level map) -> conv3x3 + ReLU stack ->
PixelShuffle(2) -> crop.  Here the pad / unshuffle / cat is one gather into bf16 token rows (kair_ffd_pack), every
convolution -- including the first (4 in_nc + 1 -> nc) and the last (nc -> 4 out_nc), whose channel counts are padded
to 64 with zero weights -- is the implicit-GEMM 3x3 on tcgen05 (kair_gemm_tn conv mode), and PixelShuffle + crop write
the NCHW image directly (kair_pixel_shuffle_out).  Forward and backward (parameter gradients; the gradient w.r.t. the
noisy image is not produced: FFDNet is trained, not inverted).  nc = 96 (the colour model) runs on 128 padded channels.
"""
import torch
import torch.nn as nn

from . import basicblock as B
from . import parallel as _parallel
from .convnet_engine import (ConvNetFunction, _bn_tile, _EngineBase, _GradStore, _PlanBase, _unsupported,
                             parse_dncnn)
from .network_dncnn import _tree_key


def _rup(x, m=64):
    return (x + m - 1) // m * m


class FFDNetEngine(_EngineBase):
    def __init__(self, stages, P, device, ops=None):
        super().__init__(P, device, ops)
        self.stages = stages
        self.nc = stages[0].conv.out_channels
        self.ncp = _rup(self.nc)
        if any(s.bn is not None for s in stages):
            _unsupported("FFDNet with BatchNorm (act_mode 'R' / 'L' is what the reference configs use)")
        if any(s.dil != 1 for s in stages):
            _unsupported("dilated FFDNet convolutions")
        self._w = {}

    def weights(self, train):
        W = self._w.get(train) or self._w.get(True)
        if W is not None:
            return W
        o, dev, st = self.ops, self.device, self.stages
        bf = lambda *s: torch.zeros(s, dtype=torch.bfloat16, device=dev)
        f32 = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        head, body, tail = st[0], st[1:-1], st[-1]
        ncp, Lb = self.ncp, len(body)
        cin, cout = head.conv.in_channels, tail.conv.out_channels
        W = {"pack": [], "versions": None, "buffers": []}
        W["Wh"], W["bh"] = bf(1, ncp, 9 * 64), f32(1, ncp)
        W["Wb"], W["bb"], W["WbT"] = bf(Lb, ncp, 9 * ncp), f32(Lb, ncp), (bf(Lb, ncp, 9 * ncp) if train else None)
        W["Wt"], W["bt"], W["WtT"] = bf(1, 64, 9 * ncp), f32(1, 64), (bf(1, ncp, 9 * 64) if train else None)
        if cin > 63 or cout > 64:
            _unsupported("FFDNet with more than 15 image channels")
        W["pack"] = [
            o.build_pack_conv([head.conv.weight], [head.conv.bias], W["Wh"], None, W["bh"], O=self.nc, I=cin, I_pad=64),
            o.build_pack_conv([s.conv.weight for s in body], [s.conv.bias for s in body], W["Wb"], W["WbT"], W["bb"],
                              O=self.nc, I=self.nc, I_pad=ncp),
            o.build_pack_conv([tail.conv.weight], [tail.conv.bias], W["Wt"], W["WtT"], W["bt"], O=cout, I=self.nc,
                              I_pad=ncp),
        ]
        self._w[train] = W
        return W

    def refresh_weights(self, plan, stream):
        W = plan.Wt
        v = self._versions()
        if W["versions"] != v:
            for l in W["pack"]:
                l.run(stream)
            W["versions"] = v

    def plan(self, B_, h, w, train):
        return self._get_plan((B_, h, w, bool(train)), lambda: _FFDPlan(self, B_, h, w, bool(train)))


class _FFDPlan(_PlanBase):
    def __init__(self, eng, Bn, h, w, train):
        o, dev, st = eng.ops, eng.device, eng.stages
        self.Wt = Wt = eng.weights(train)
        head, body, tail = st[0], st[1:-1], st[-1]
        nc, ncp, Lb = eng.nc, eng.ncp, len(body)
        C = (head.conv.in_channels - 1) // 4
        co = tail.conv.out_channels // 4
        H2, W2 = (h + 1) // 2, (w + 1) // 2
        T, geo = Bn * H2 * W2, (Bn, H2, W2)
        f32 = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        bf = lambda *s: torch.zeros(s, dtype=torch.bfloat16, device=dev)
        self.img, self.sigma = f32(Bn, C, h, w), f32(Bn)
        self.out_full = f32(Bn, co, 2 * H2, 2 * W2)          # before the crop of network_ffdnet.py:68
        self.out = self.out_full
        self.crop = (h, w)
        zc = f32(co)
        bn = _bn_tile(ncp)
        slope = lambda s: (o.ACT_LRELU, s.slope) if s.slope is not None else (o.ACT_NONE, 0.0)
        xin = bf(T, 64)
        acts = [bf(T, ncp) for _ in range(Lb + 1)] if train else [bf(T, ncp), bf(T, ncp)]
        A = lambda i: acts[i] if train else acts[i % 2]
        tailo = f32(T, 64)
        fwd = [o.build_ffd_pack(self.img, self.sigma, xin)]
        a0, s0 = slope(head)
        fwd.append(o.build_gemm_tn(xin, Wt["Wh"][0], N=ncp, K=64, bn=bn, conv=geo, bias=Wt["bh"][0], act=a0, act_slope=s0,
                                   outB=A(0)))
        for i, s in enumerate(body):
            a_, s_ = slope(s)
            fwd.append(o.build_gemm_tn(A(i), Wt["Wb"][i], N=ncp, K=ncp, bn=bn, conv=geo, bias=Wt["bb"][i], act=a_,
                                       act_slope=s_, outB=A(i + 1)))
        fwd.append(o.build_gemm_tn(A(Lb), Wt["Wt"][0], N=64, K=ncp, bn=64, conv=geo, bias=Wt["bt"][0], outF=tailo))
        fwd.append(o.build_pixel_shuffle_out(tailo, self.out_full, zc, r=2, scale=1.0))
        self.fwd, self.bwd = fwd, None
        self.zero_before_bwd = []
        if not train:
            self._finish(eng)
            return
        # ================================================================================ backward
        self.dout = f32(Bn, co, 2 * H2, 2 * W2)              # the crop's zero padding is written by the caller
        self.grads = _GradStore(eng.P, dev)
        G = lambda p: self.grads.G[eng.name_of(p)]
        dWh, dWb, dWt = f32(1, ncp, 9 * 64), f32(Lb, ncp, 9 * ncp), f32(1, 64, 9 * ncp)
        self.zero_before_bwd = [dWh, dWb, dWt]
        g_t = bf(T, 64)
        d = [bf(T, ncp), bf(T, ncp)]
        bwd = [o.build_pixel_shuffle_out_bwd(self.dout, g_t, r=2, co=co, scale=1.0)]
        bwd.append(o.build_colsum(g_t, G(tail.conv.bias), C=4 * co))
        bwd.append(o.build_gemm_nt(g_t, A(Lb), dWt[0], n1=64, n2=ncp, bn=bn, conv=geo))
        last = body[-1] if Lb else head
        bwd.append(o.build_gemm_tn(g_t, Wt["WtT"][0], N=ncp, K=64, bn=bn, conv=geo, outB=d[0],
                                   aux=A(Lb) if last.slope is not None else None, act_slope=last.slope or 0.0))
        cur = 0
        for i in reversed(range(Lb)):
            s = body[i]
            bwd.append(o.build_colsum(d[cur], G(s.conv.bias), C=nc))
            bwd.append(o.build_gemm_nt(d[cur], A(i), dWb[i], n1=ncp, n2=ncp, bn=bn, conv=geo))
            prev = body[i - 1] if i > 0 else head
            bwd.append(o.build_gemm_tn(d[cur], Wt["WbT"][i], N=ncp, K=ncp, bn=bn, conv=geo, outB=d[cur ^ 1],
                                       aux=A(i) if prev.slope is not None else None, act_slope=prev.slope or 0.0))
            cur ^= 1
        bwd.append(o.build_colsum(d[cur], G(head.conv.bias), C=nc))
        bwd.append(o.build_gemm_nt(d[cur], xin, dWh[0], n1=ncp, n2=64, bn=64, conv=geo))
        bwd.append(o.build_unpack_conv_grad(dWh, [G(head.conv.weight)], None, O=nc, I=head.conv.in_channels, I_pad=64))
        if Lb:
            bwd.append(o.build_unpack_conv_grad(dWb, [G(s.conv.weight) for s in body], None, O=nc, I=nc, I_pad=ncp))
        bwd.append(o.build_unpack_conv_grad(dWt, [G(tail.conv.weight)], None, O=4 * co, I=nc, I_pad=ncp))
        self.bwd = bwd
        self._finish(eng)

    def run_forward(self, x):
        x, sigma = x
        self.sigma.copy_(sigma.reshape(-1))
        out = super().run_forward(x)
        return out[..., :self.crop[0], :self.crop[1]].contiguous()

    def run_backward(self, dout):
        full = torch.zeros_like(self.dout)
        full[..., :self.crop[0], :self.crop[1]] = dout
        return super().run_backward(full)


class _FFDFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, sigma, plan, *params):
        from .swinir_engine import mark_busy
        out = plan.run_forward((x, sigma))
        mark_busy(plan)
        ctx.plan, ctx.generation = plan, plan.generation
        return out

    @staticmethod
    def backward(ctx, dout):
        plan = ctx.plan
        if plan.generation != ctx.generation:
            raise RuntimeError("kair_b200: backward() called after two later forward() calls of the same shape "
                               "overwrote the saved activations (two sets are kept per shape)")
        plan.busy = False
        return (None, None, None) + tuple(plan.run_backward(dout.contiguous()))


class FFDNet(nn.Module):
    def __init__(self, in_nc=1, out_nc=1, nc=64, nb=15, act_mode="R"):
        super().__init__()
        assert "R" in act_mode or "L" in act_mode, "Examples of activation function: R, L, BR, BL, IR, IL"
        bias, sf = True, 2
        m_head = B.conv(in_nc * sf * sf + 1, nc, mode="C" + act_mode[-1], bias=bias)
        m_body = [B.conv(nc, nc, mode="C" + act_mode, bias=bias) for _ in range(nb - 2)]
        m_tail = B.conv(nc, out_nc * sf * sf, mode="C", bias=bias)
        self.model = B.sequential(m_head, *m_body, m_tail)
        self._engine, self._engine_key, self._ops = None, None, None

    def engine(self, device):
        key = _tree_key(self, device)
        if self._engine is None or self._engine_key != key:
            self._engine = FFDNetEngine(_parse(self.model), dict(self.named_parameters()), device, ops=self._ops)
            self._engine_key = key
        self._engine.grad_hook = (lambda flat: _parallel.allreduce_flat(self, flat))
        return self._engine

    def forward(self, x, sigma):
        if not x.is_cuda and self._ops is None:
            raise RuntimeError("kair_b200 FFDNet runs on sm_100a GPUs only; got a %s tensor (there is no CPU fallback)"
                               % x.device.type)
        if x.requires_grad and torch.is_grad_enabled():
            raise NotImplementedError("kair_b200 FFDNet: the gradient w.r.t. the input image is not produced")
        x = x.float().contiguous()
        sigma = sigma.float().reshape(x.shape[0]).contiguous()
        eng = self.engine(x.device)
        params = list(eng.P.values())
        train = torch.is_grad_enabled() and any(p.requires_grad for p in params)
        plan = eng.plan(x.shape[0], x.shape[2], x.shape[3], train)
        if train:
            return _FFDFunction.apply(x, sigma, plan, *params)
        return plan.run_forward((x, sigma))


def _parse(seq):
    """conv [+ activation] stages of FFDNet.model, without DnCNN's limits on the image-side channel counts"""
    from .convnet_engine import DnCNNStage, _act_slope, _is3x3
    stages = []
    for m in seq.children():
        if isinstance(m, nn.Conv2d):
            if not _is3x3(m):
                _unsupported("FFDNet convolution %r is not 3x3 / stride 1 / pad 1" % (m,))
            stages.append(DnCNNStage(m))
        elif isinstance(m, nn.BatchNorm2d):
            stages[-1].bn = m
        elif _act_slope(m) is not None:
            stages[-1].slope = _act_slope(m)
        else:
            _unsupported("module %s inside FFDNet.model" % type(m).__name__)
    if len(stages) < 2 or stages[-1].slope is not None or stages[0].conv.bias is None:
        _unsupported("FFDNet layout")
    return stages
