"""Launch-list engines for the conv nets of the hot path: DnCNN (conv3x3 [+BN] + ReLU stack, residual output) and
UNetRes / DRUNet (ResBlocks, 2x2 strided conv down, 2x2 transposed conv up), forward and backward.

Same construction as swinir_engine: one (batch, H, W) problem becomes a statically allocated HBM workspace plus two
ordered lists of pre-built kernel launches; a step replays the lists (one ctypes call per kernel, no allocation, no
host synchronisation).

HBM layout: activations are NHWC token matrices, t = (b*H + h)*W + w, channel counts are multiples of 64 (64 / 128 /
256 / 512), bf16 for every tensor-core operand (conv inputs, saved activations, gradients entering a GEMM), fp32 for
the residual stream of the ResBlocks, the U-Net skip features, BatchNorm inputs / statistics and all parameter
gradients.  3x3 convolutions are implicit GEMMs on tcgen05 (kair_gemm_tn conv mode: 9 shifted TMA boxes, zero padding
by TMA out-of-bounds fill) with bias / ReLU / residual fused in the epilogue; their input gradients are the same
kernel on the flipped weights with the ReLU mask of the previous layer fused in; weight gradients are kair_gemm_nt.
The strided conv / transposed conv are dense GEMMs over 2x2-regrouped rows (kair_space_to_depth / depth_to_space).

Reference arithmetic replaced (file:line in /root/reference): models/network_dncnn.py:61-71,
models/network_unet.py:25-76, models/basicblock.py:61-98, 211-223, 471-477, 495-501 and autograd's backward of them.
"""
import os
from collections import OrderedDict

import torch
import torch.nn as nn

from . import _dirty
from . import ops as _real_ops


# kair_conv_chain (the whole conv body as one persistent launch with grid barriers) serves inference plans of at most this
# many 8 x 16 pixel tiles (env KAIR_B200_CHAIN_MAX_TILES).  Default 0 = OFF: measured on B200 it does not beat the per-layer
# launches -- 151 vs 148 us for the 15 body layers of DnCNN at batch 1, 1316 vs 1270 us at batch 16 -- because a
# 128 x 64 x 16 UMMA costs ~77 cycles whatever feeds it (1.45 us per tile with the fill, the epilogue and the stores
# ablated away), and PDL + graph replay already hide the launch gaps the chain was meant to remove.
CHAIN_MAX_TILES = int(os.environ.get("KAIR_B200_CHAIN_MAX_TILES", "0"))


def _conv_tile(n):
    """N tile of the 3x3 implicit GEMMs: weights of >= 128-channel layers are streamed through a shared-memory ring,
    where 128-wide tiles leave room for a deep enough ring"""
    return _bn_tile(n) if n < 256 else (128 if n % 128 == 0 else _bn_tile(n))


def _bn_tile(n):
    """N tile of the tcgen05 GEMMs for an n-column output (n is a multiple of 64)."""
    for bn in (256, 192, 128, 64):
        if n % bn == 0:
            return bn
    raise NotImplementedError("kair_b200: %d output channels (must be a multiple of 64)" % n)


def _unsupported(msg):
    raise NotImplementedError("kair_b200 conv nets: " + msg + " (see DESIGN.md, 'coverage')")


def _act_slope(m):
    if isinstance(m, nn.ReLU):
        return 0.0
    if isinstance(m, nn.LeakyReLU):
        return float(m.negative_slope)
    return None


def _is3x3(c, dilated=False):
    """3x3, stride 1, 'same' zero padding; dilated=True also accepts nn.Conv2d(.., padding=d, dilation=d) (IRCNN,
    models/network_dncnn.py:99-111)"""
    if not (isinstance(c, nn.Conv2d) and c.kernel_size == (3, 3) and c.stride == (1, 1) and c.groups == 1):
        return False
    d = c.dilation[0]
    if c.dilation != (d, d) or c.padding != (d, d):
        return False
    return d == 1 or (dilated and 1 <= d <= 8)


class _GradStore:
    """Parameter gradients in the parameters' own layouts: one flat fp32 buffer with one view per parameter."""

    def __init__(self, P, dev):
        sizes = [p.numel() for p in P.values()]
        self.flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
        self.G, off = {}, 0
        for (k, p), n in zip(P.items(), sizes):
            self.G[k] = self.flat[off:off + n].view(p.shape)
            off += n
        self.P = P

    def snapshot(self):
        flat = self.flat.clone()
        grads, off = [], 0
        for p in self.P.values():
            n = p.numel()
            grads.append(flat[off:off + n].view(p.shape))
            off += n
        return grads


class _EngineBase:
    def __init__(self, P, device, ops):
        self.P = P
        self.device = torch.device(device)
        self.ops = ops or _real_ops
        self._plans = {}
        self._names = {id(p): k for k, p in P.items()}
        self.grad_hook = None   # callable(flat fp32 gradient buffer), e.g. parallel.allreduce_flat bound to the module

    def name_of(self, p):
        return self._names[id(p)]

    def _versions(self, extra=()):
        # + the global optimizer-step epoch: fused optimizers do not touch the version counters (kair_b200/_dirty.py)
        return (_dirty.optimizer_epoch(),) + tuple(p._version for p in self.P.values()) + tuple(b._version for b in extra)

    def _get_plan(self, key, make):
        from .swinir_engine import pick_plan
        return pick_plan(self._plans, key, 4, make)   # two activation sets per shape: see pick_plan


class _PlanBase:
    """Execution of a built plan: forward list, backward list, gradient store, buffers zeroed before backward."""

    def _finish(self, eng):
        self.eng = eng
        self.generation = 0

    def _lists(self):
        """the launch lists as C arrays (one FFI call per pass), built on first use"""
        ls = getattr(self, "_launch_lists", None)
        if ls is None:
            o = self.eng.ops
            ls = self._launch_lists = (o.make_list(self.fwd), o.make_list(self.bwd) if self.bwd is not None else None)
        return ls

    def run_forward(self, x):
        o = self.eng.ops
        with o.device_guard(self.img.device):   # the launchers use the current device and its current stream
            st = o.stream_handle(self.img.device)
            self.eng.refresh_weights(self, st)
            self.img.copy_(x)
            self._lists()[0].run(st)
        # kair_bn_finalize wrote running_mean / running_var / num_batches_tracked through raw pointers: advance their
        # version counters like an in-place torch op would, so that the folded eval-mode weights (keyed on these
        # versions, refresh_weights) and GraphedInference see the change even when no optimizer step follows
        for b in getattr(self, "bn_buffers", ()):
            torch.autograd.graph.increment_version(b)
        self.generation += 1
        return self.out.clone()

    def run_backward(self, dout):
        o = self.eng.ops
        with o.device_guard(self.img.device):
            st = o.stream_handle(self.img.device)
            self.dout.copy_(dout)
            self.grads.flat.zero_()
            for t in self.zero_before_bwd:
                t.zero_()
            self._lists()[1].run(st)
            if self.eng.grad_hook is not None:
                self.eng.grad_hook(self.grads.flat)
        return self.grads.snapshot()


class ConvNetFunction(torch.autograd.Function):
    """forward/backward pair over a whole conv net (the reference's plugin convention, models/op/fused_act.py:21-72)."""

    @staticmethod
    def forward(ctx, x, plan, *params):
        from .swinir_engine import mark_busy
        out = plan.run_forward(x)
        mark_busy(plan)
        ctx.plan = plan
        ctx.generation = plan.generation
        return out

    @staticmethod
    def backward(ctx, dout):
        plan = ctx.plan
        if plan.generation != ctx.generation:
            raise RuntimeError("kair_b200: backward() called after two later forward() calls of the same shape "
                               "overwrote the saved activations (two sets are kept per shape)")
        plan.busy = False
        grads = plan.run_backward(dout.contiguous())
        dx = None
        if ctx.needs_input_grad[0]:
            from .swinir_engine import input_grad
            dx = input_grad(plan, dout)
        return (dx, None) + tuple(grads)


# ====================================================================================================== DnCNN

class DnCNNStage:
    def __init__(self, conv):
        self.conv, self.bn, self.slope = conv, None, None
        self.dil = conv.dilation[0]


def parse_dncnn(seq):
    """Group the children of DnCNN.model (network_dncnn.py:61-67, possibly after merge_bn, utils_bnorm.py:32-63) into
    conv [+BN] [+activation] stages."""
    children = list(seq.children()) if isinstance(seq, nn.Sequential) else [seq]
    stages = []
    for m in children:
        if isinstance(m, nn.Conv2d):
            if not _is3x3(m, dilated=True):
                _unsupported("DnCNN / IRCNN convolution %r is not 3x3 / stride 1 / padding == dilation" % (m,))
            stages.append(DnCNNStage(m))
        elif isinstance(m, nn.BatchNorm2d):
            if not stages or stages[-1].bn is not None or stages[-1].slope is not None:
                _unsupported("BatchNorm2d must directly follow a convolution")
            if not (m.affine and m.track_running_stats and m.momentum is not None):
                _unsupported("BatchNorm2d without affine / running statistics / momentum")
            stages[-1].bn = m
        elif _act_slope(m) is not None:
            if not stages or stages[-1].slope is not None:
                _unsupported("activation must follow a convolution")
            stages[-1].slope = _act_slope(m)
        else:
            _unsupported("module %s inside DnCNN.model" % type(m).__name__)
    if len(stages) < 3:
        _unsupported("DnCNN needs at least 3 convolutions")
    head, tail = stages[0], stages[-1]
    nc = head.conv.out_channels
    if head.bn is not None or tail.bn is not None or tail.slope is not None:
        _unsupported("DnCNN head with BatchNorm / tail with BatchNorm or activation")
    if head.conv.in_channels > 4 or tail.conv.out_channels > 4:
        _unsupported("DnCNN with more than 4 image channels")
    if head.dil != 1 or tail.dil != 1:
        _unsupported("dilated first / last convolution (the thin image-side kernels are dense 3x3)")
    if nc % 64 or nc > 256:
        _unsupported("DnCNN width nc=%d (must be 64, 128, 192 or 256)" % nc)
    for s in stages[1:-1]:
        if s.conv.in_channels != nc or s.conv.out_channels != nc:
            _unsupported("DnCNN body convolutions must be nc -> nc")
    if tail.conv.in_channels != nc:
        _unsupported("DnCNN tail convolution must read nc channels")
    return stages


class DnCNNEngine(_EngineBase):
    def __init__(self, stages, P, device, ops=None, residual=True):
        """residual: DnCNN returns x - model(x) (network_dncnn.py:69-71); FDnCNN returns model(x) (:146-148)."""
        super().__init__(P, device, ops)
        self.stages = stages
        self.residual = bool(residual)
        if self.residual and stages[0].conv.in_channels != stages[-1].conv.out_channels:
            _unsupported("residual DnCNN output needs in_nc == out_nc")
        self.nc = stages[0].conv.out_channels
        self.Lb = len(stages) - 2
        self._w = {}

    # ------------------------------------------------------------------------------------------ packed weights
    def weights(self, train, fold):
        """Packed bf16 operands of the body convolutions.  fold: eval-mode BatchNorm folded into weights and bias."""
        key = (bool(train), bool(fold))
        W = self._w.get(key)
        if W is not None:
            return W
        o, dev, nc, Lb = self.ops, self.device, self.nc, self.Lb
        body = self.stages[1:-1]
        W = {"fold": fold, "versions": None, "buffers": []}
        W["Wp"] = torch.zeros(Lb, nc, 9 * nc, dtype=torch.bfloat16, device=dev)
        W["WpT"] = torch.zeros(Lb, nc, 9 * nc, dtype=torch.bfloat16, device=dev) if train else None
        W["bp"] = torch.zeros(Lb, nc, dtype=torch.float32, device=dev)
        pack = []
        osc = osh = None
        if fold and any(s.bn is not None for s in body):
            sc = torch.ones(Lb, nc, dtype=torch.float32, device=dev)
            sh = torch.zeros(Lb, nc, dtype=torch.float32, device=dev)
            for i, s in enumerate(body):
                if s.bn is not None:
                    pack.append(o.build_bn_fold(s.bn.weight, s.bn.bias, s.bn.running_mean, s.bn.running_var, sc[i],
                                                sh[i], C=nc, eps=s.bn.eps))
                    W["buffers"] += [s.bn.running_mean, s.bn.running_var]
            osc, osh = [sc[i] for i in range(Lb)], [sh[i] for i in range(Lb)]
        biases = [s.conv.bias for s in body]
        if any(b is None for b in biases):
            if not all(b is None for b in biases):
                _unsupported("DnCNN body convolutions with and without bias mixed")
            biases = None
        pack.append(o.build_pack_conv([s.conv.weight for s in body], biases, W["Wp"], W["WpT"], W["bp"], O=nc, I=nc,
                                      I_pad=nc, oscale=osc, oshift=osh))
        W["pack"] = pack
        self._w[key] = W
        return W

    def refresh_weights(self, plan, stream):
        W = plan.Wt
        v = self._versions(W["buffers"])
        if W["versions"] != v:
            for l in W["pack"]:
                l.run(stream)
            W["versions"] = v

    def plan(self, B, H, W, train, bn_batch):
        key = (B, H, W, bool(train), bool(bn_batch))
        return self._get_plan(key, lambda: _DnCNNPlan(self, B, H, W, bool(train), bool(bn_batch)))


class _DnCNNPlan(_PlanBase):
    def __init__(self, eng, B, H, W, train, bn_batch):
        o, dev, nc, Lb = eng.ops, eng.device, eng.nc, eng.Lb
        stages = eng.stages
        head, body, tail = stages[0], stages[1:-1], stages[-1]
        has_bn = any(s.bn is not None for s in body)
        if train and has_bn and not bn_batch:
            _unsupported("autograd through eval-mode BatchNorm (call model.train(), or merge_bn first)")
        fold = has_bn and not bn_batch
        self.Wt = eng.weights(train, fold)
        Wt = self.Wt
        T = B * H * W
        geo = (B, H, W)
        cin, cout = head.conv.in_channels, tail.conv.out_channels
        f32 = lambda *sh: torch.zeros(sh, dtype=torch.float32, device=dev)
        bf = lambda *sh: torch.zeros(sh, dtype=torch.bfloat16, device=dev)
        act_of = lambda s: (o.ACT_LRELU, s.slope) if s.slope is not None else (o.ACT_NONE, 0.0)
        self.img, self.out = f32(B, cin, H, W), f32(B, cout, H, W)
        zc_in, zc_out = f32(cin), f32(cout)
        bn = _bn_tile(nc)
        # activations: every stage output is kept when training, two ping-pong buffers otherwise
        n_act = Lb + 1
        acts = [bf(T, nc) for _ in range(n_act)] if train else [bf(T, nc), bf(T, nc)]
        A = lambda i: acts[i] if train else acts[i % 2]
        bn_live = bn_batch and has_bn
        zs = {}
        if bn_live:
            zs = {i: f32(T, nc) for i, s in enumerate(body) if s.bn is not None} if train else None
            zshared = None if train else f32(T, nc)
            self.bn_acc = torch.zeros(2, nc, dtype=torch.float64, device=dev)
            self.bn_mean, self.bn_rstd = f32(Lb, nc), f32(Lb, nc)

        fwd = []
        a0, s0 = act_of(head)
        self.bn_buffers = []
        fwd.append(o.build_conv_in_fwd(self.img, head.conv.weight, head.conv.bias, zc_in, 1.0, O=nc, O_pad=nc, pitch=nc,
                                       out_bf16=A(0), act=a0, act_slope=s0))
        # opt-in (CHAIN_MAX_TILES): the whole body as ONE persistent launch with a grid barrier between layers
        # (kair_conv_chain) instead of one launch per layer
        tiles = B * ((H + 15) // 16) * ((W + 7) // 8)
        chain = (not train and not bn_live and nc == 64 and 1 <= Lb <= 32 and all(s.dil == 1 for s in body)
                 and tiles <= CHAIN_MAX_TILES and hasattr(o, "build_conv_chain"))
        self.chain = chain
        if chain:
            self.chain_slope = torch.tensor([1.0 if s.slope is None else float(s.slope) for s in body],
                                            dtype=torch.float32, device=dev)
            self.chain_sync = torch.zeros(2, dtype=torch.int32, device=dev)
            fwd.append(o.build_conv_chain(acts[0], acts[1], Wt["Wp"], Wt["bp"], self.chain_slope, self.chain_sync,
                                          L=Lb, conv=geo))
        for i, s in enumerate(body if not chain else ()):
            act, slope = act_of(s)
            if s.bn is not None and bn_live:
                z = zs[i] if train else zshared
                fwd.append(o.build_gemm_tn(A(i), Wt["Wp"][i], N=nc, K=nc, bn=bn, conv=geo, bias=Wt["bp"][i], outF=z,
                                           dilation=s.dil))
                fwd.append(o.build_bn_stats(z, self.bn_acc, C=nc))
                fwd.append(o.build_bn_finalize(self.bn_acc, self.bn_mean[i], self.bn_rstd[i], s.bn.running_mean,
                                               s.bn.running_var, s.bn.num_batches_tracked, C=nc, T=T, eps=s.bn.eps,
                                               momentum=s.bn.momentum))
                self.bn_buffers += [s.bn.running_mean, s.bn.running_var, s.bn.num_batches_tracked]
                fwd.append(o.build_bn_apply(z, self.bn_mean[i], self.bn_rstd[i], s.bn.weight, s.bn.bias, C=nc, act=act,
                                            act_slope=slope, out_bf16=A(i + 1)))
            else:
                fwd.append(o.build_gemm_tn(A(i), Wt["Wp"][i], N=nc, K=nc, bn=bn, conv=geo, bias=Wt["bp"][i], act=act,
                                           act_slope=slope, outB=A(i + 1), dilation=s.dil))
        if eng.residual:
            # x - n  ==  -(conv(feat) + b - x)
            fwd.append(o.build_conv_out_fwd(A(Lb), tail.conv.weight, tail.conv.bias, self.out, zc_out, C=nc, pitch=nc,
                                            out_scale=-1.0, skip=self.img, skip_mean=zc_out, skip_scale=-1.0))
        else:
            fwd.append(o.build_conv_out_fwd(A(Lb), tail.conv.weight, tail.conv.bias, self.out, zc_out, C=nc, pitch=nc,
                                            out_scale=1.0))
        self.fwd, self.bwd = fwd, None
        self.zero_before_bwd = []
        if not train:
            self._finish(eng)
            return

        # ================================================================================ backward
        self.dout = f32(B, cout, H, W)
        self.grads = _GradStore(eng.P, dev)
        G = lambda p: self.grads.G[eng.name_of(p)]
        # 64 -> 64 layers: weight gradients with the operand roles swapped (kair_gemm_nt `swap`: every row of the 128-row
        # UMMA is useful), which leaves the packed gradient transposed, [9*nc, nc]
        swap = nc == 64
        dWp = f32(Lb, 9 * nc, nc) if swap else f32(Lb, nc, 9 * nc)
        self.zero_before_bwd = [dWp]
        d = [bf(T, nc), bf(T, nc)]   # ping-pong gradients w.r.t. stage outputs / pre-activations
        # the gradient entering a live BatchNorm stays fp32: its backward subtracts projections of dy, which would
        # amplify bf16 rounding by the ratio |dy| / |dz|
        df = f32(T, nc) if bn_live else None
        live = lambda s: s.bn is not None and bn_live
        bwd = []
        # thin tail / head convolutions as tensor-core GEMMs over the 9-tap expansion of the <= 4 image channels
        # (columns tap*7 + c, K padded to 64): d(out) expanded with dir=-1 serves the tail's dgrad and wgrad
        dsh, clT, dWcl, csum = bf(T, 64), bf(nc, 64), f32(nc, 64), f32(64)
        xcol, dWcf = bf(T, 64), f32(1, nc, 64)
        self.zero_before_bwd += [dWcl, csum, dWcf]
        bwd.append(o.build_im2col3x3(self.dout, dsh, cpad=7, direction=-1,
                                     scale=-1.0 if eng.residual else 1.0))   # out = x - n  /  out = n
        bwd.append(o.build_conv_out_pack(mode=0, co=cout, C=nc, cpad=7, W=tail.conv.weight, Wt=clT))
        bwd.append(o.build_gemm_nt(A(Lb), dsh, dWcl, n1=nc, n2=64, bn=64))
        bwd.append(o.build_conv_out_pack(mode=1, co=cout, C=nc, cpad=7, D=dWcl, dW=G(tail.conv.weight)))
        if tail.conv.bias is not None:
            bwd.append(o.build_colsum(dsh, csum, C=64))
            bwd.append(o.build_copy(G(tail.conv.bias), csum[4 * 7:4 * 7 + cout]))
        # gradient w.r.t. the last body output; the activation mask is applied here unless that stage has a live BN
        last = body[-1]
        if live(last):
            bwd.append(o.build_gemm_tn(dsh, clT, N=nc, K=64, bn=bn, outF=df))
        else:
            bwd.append(o.build_gemm_tn(dsh, clT, N=nc, K=64, bn=bn, outB=d[0],
                                       aux=A(Lb) if last.slope is not None else None, act_slope=last.slope or 0.0))
        cur = 0
        for i in reversed(range(Lb)):
            s = body[i]
            act, slope = act_of(s)
            if live(s):
                g = d[cur]
                bwd.append(o.build_bn_bwd_reduce(zs[i], A(i + 1), df, self.bn_mean[i], self.bn_rstd[i], G(s.bn.weight),
                                                 G(s.bn.bias), C=nc, act=act, act_slope=slope))
                bwd.append(o.build_bn_bwd_apply(zs[i], A(i + 1), df, self.bn_mean[i], self.bn_rstd[i], s.bn.weight,
                                                G(s.bn.weight), G(s.bn.bias), g, C=nc, act=act, act_slope=slope,
                                                dbias=G(s.conv.bias) if s.conv.bias is not None else None))
            else:
                g = d[cur]
                if s.conv.bias is not None:
                    bwd.append(o.build_colsum(g, G(s.conv.bias), C=nc))
            bwd.append(o.build_gemm_nt(g, A(i), dWp[i], n1=nc, n2=nc, bn=bn, conv=geo, dilation=s.dil, swap=swap))
            prev = body[i - 1] if i > 0 else head
            if live(prev):
                bwd.append(o.build_gemm_tn(g, Wt["WpT"][i], N=nc, K=nc, bn=bn, conv=geo, outF=df, dilation=s.dil))
            else:
                bwd.append(o.build_gemm_tn(g, Wt["WpT"][i], N=nc, K=nc, bn=bn, conv=geo, outB=d[cur ^ 1],
                                           aux=A(i) if prev.slope is not None else None, act_slope=prev.slope or 0.0,
                                           dilation=s.dil))
            cur ^= 1
        bwd.append(o.build_im2col3x3(self.img, xcol, cpad=7, direction=1, mean=zc_in, scale=1.0, ones_col=4 * 7 + cin))
        bwd.append(o.build_gemm_nt(d[cur], xcol, dWcf[0], n1=nc, n2=64, bn=64))
        bwd.append(o.build_unpack_conv_grad(dWcf, [G(head.conv.weight)],
                                            [G(head.conv.bias)] if head.conv.bias is not None else None, O=nc, I=cin,
                                            I_pad=7, ones_col=cin, center_tap=4, dwp_pitch=64))
        bwd.append(o.build_unpack_conv_grad(dWp, [G(s.conv.weight) for s in body], None, O=nc, I=nc, I_pad=nc,
                                            transposed=swap))
        self.bwd = bwd
        self._dx_spec = dict(dhead=d[cur], W=head.conv.weight, C=nc, pitch=nc, scale=1.0, skip=bool(eng.residual))
        self._finish(eng)


# ====================================================================================================== UNetRes

class ResSpec:
    def __init__(self, block):
        mods = list(block.res.children())
        if len(mods) != 3 or not _is3x3(mods[0]) or not _is3x3(mods[2]) or _act_slope(mods[1]) is None:
            _unsupported("ResBlock must be conv3x3 - ReLU/LeakyReLU - conv3x3 (mode 'CRC' / 'CLC')")
        self.c1, self.c2, self.slope = mods[0], mods[2], _act_slope(mods[1])


def parse_unetres(model):
    """Read the module tree of UNetRes (network_unet.py:25-57) into per-level ResBlock lists and the resampling
    convolutions.  Levels: 0 (full resolution) .. 3."""
    def split(seq, first_is_up):
        mods = list(seq.children())
        if first_is_up:
            up, blocks = mods[0], mods[1:]
            if not (isinstance(up, nn.ConvTranspose2d) and up.kernel_size == (2, 2) and up.stride == (2, 2)
                    and up.padding == (0, 0)):
                _unsupported("UNetRes upsample_mode other than 'convtranspose' (2x2, stride 2)")
            return up, [ResSpec(b) for b in blocks]
        blocks, down = mods[:-1], mods[-1]
        if not (isinstance(down, nn.Conv2d) and down.kernel_size == (2, 2) and down.stride == (2, 2)
                and down.padding == (0, 0)):
            _unsupported("UNetRes downsample_mode other than 'strideconv' (2x2, stride 2)")
        return down, [ResSpec(b) for b in blocks]

    if not _is3x3(model.m_head) or not _is3x3(model.m_tail):
        _unsupported("UNetRes head / tail must be single 3x3 convolutions")
    net = {"head": model.m_head, "tail": model.m_tail, "down": [], "up": [None, None, None]}
    for name in ("m_down1", "m_down2", "m_down3"):
        net["down"].append(split(getattr(model, name), False))
    body = model.m_body
    net["body"] = [ResSpec(b) for b in (body.children() if isinstance(body, nn.Sequential) else [body])]
    for lvl, name in ((2, "m_up3"), (1, "m_up2"), (0, "m_up1")):
        net["up"][lvl] = split(getattr(model, name), True)
    if not net["body"] or any(not blocks for _, blocks in net["down"]) or any(not blocks for _, blocks in net["up"]):
        _unsupported("UNetRes with nb == 0 (every stage needs at least one ResBlock)")
    nc = [net["head"].out_channels] + [d.out_channels for d, _ in net["down"]]
    for c in nc:
        if c % 64:
            _unsupported("UNetRes channel count %d (must be a multiple of 64)" % c)
    if net["head"].in_channels > 4 or net["tail"].out_channels > 4:
        _unsupported("UNetRes with more than 4 image channels")
    net["nc"] = nc
    return net


class UNetResEngine(_EngineBase):
    def __init__(self, net, P, device, ops=None):
        super().__init__(P, device, ops)
        self.net = net
        self.nc = net["nc"]
        # ResBlocks per level in execution order: down path, (body), up path
        self.level_blocks = []
        for lvl in range(4):
            blocks = list(net["down"][lvl][1]) if lvl < 3 else list(net["body"])
            self.n_down = [len(net["down"][l][1]) for l in range(3)]
            if lvl < 3:
                blocks += list(net["up"][lvl][1])
            self.level_blocks.append(blocks)
        self._w = {}

    def weights(self, train):
        W = self._w.get(bool(train))
        if W is not None:
            return W
        o, dev, nc = self.ops, self.device, self.nc
        bf = lambda *sh: torch.zeros(sh, dtype=torch.bfloat16, device=dev)
        f32 = lambda *sh: torch.zeros(sh, dtype=torch.float32, device=dev)
        W = {"versions": None, "buffers": [], "pack": [], "res": [], "down": [], "up": []}
        i32 = lambda v: torch.tensor(v, dtype=torch.int32, device=dev)
        for lvl in range(4):
            C = nc[lvl]
            convs = [c for b in self.level_blocks[lvl] for c in (b.c1, b.c2)]
            n = len(convs)
            R = {"Wp": bf(n, C, 9 * C), "WpT": bf(n, C, 9 * C) if train else None, "bp": None, "convs": convs}
            biases = [c.bias for c in convs]
            if all(b is not None for b in biases) and n:
                R["bp"] = f32(n, C)
            elif any(b is not None for b in biases):
                _unsupported("ResBlock convolutions with and without bias mixed")
            if n:
                W["pack"].append(o.build_pack_conv([c.weight for c in convs], biases if R["bp"] is not None else None,
                                                   R["Wp"], R["WpT"], R["bp"], O=C, I=C, I_pad=C))
            W["res"].append(R)
        for lvl in range(3):
            Ci, Co = nc[lvl], nc[lvl + 1]
            # strided conv weight (Co, Ci, 2, 2) seen as a linear layer (Co, Ci*4) over 2x2-regrouped rows:
            # packed column tap*Ci + i  <-  flat input index i*4 + tap
            down = self.net["down"][lvl][0]
            D = {"Wp": bf(1, Co, 4 * Ci), "WpT": bf(1, 4 * Ci, Co) if train else None,
                 "bp": f32(1, Co) if down.bias is not None else None, "conv": down,
                 "rows": i32(list(range(Co))), "cols": i32([(k % Ci) * 4 + k // Ci for k in range(4 * Ci)])}
            W["pack"].append(o.build_pack_linear([down.weight.detach().view(Co, Ci * 4)],
                                                 [down.bias] if down.bias is not None else None, None, None,
                                                 D["rows"], D["cols"], None, None, D["Wp"], D["WpT"], D["bp"],
                                                 out_f=Co, in_f=4 * Ci))
            W["down"].append(D)
            # transposed conv weight (Cc, Cf, 2, 2) (coarse -> fine channels) seen as a linear layer (Cc, Cf*4):
            # its packed form Wp' [Cc][tap*Cf + o] is the input-gradient operand, the transpose WpT' the forward one
            up = self.net["up"][lvl][0]
            Cc, Cf = nc[lvl + 1], nc[lvl]
            U = {"Wp": bf(1, Cc, 4 * Cf), "WpT": bf(1, 4 * Cf, Cc), "conv": up,
                 "rows": i32(list(range(Cc))), "cols": i32([(k % Cf) * 4 + k // Cf for k in range(4 * Cf)]),
                 "bp": f32(4 * Cf) if up.bias is not None else None}
            W["pack"].append(o.build_pack_linear([up.weight.detach().view(Cc, Cf * 4)], None, None, None, U["rows"], U["cols"],
                                                 None, None, U["Wp"], U["WpT"], None, out_f=Cc, in_f=4 * Cf))
            W["up"].append(U)
        self._w[bool(train)] = W
        return W

    def refresh_weights(self, plan, stream):
        W = plan.Wt
        v = self._versions()
        if W["versions"] != v:
            for l in W["pack"]:
                l.run(stream)
            for U in W["up"]:
                if U["bp"] is not None:  # the transposed-conv bias applies to all four taps of the regrouped row
                    U["bp"].copy_(U["conv"].bias.detach().repeat(4))
            W["versions"] = v

    def plan(self, B, H, W, train):
        return self._get_plan((B, H, W, bool(train)), lambda: _UNetResPlan(self, B, H, W, bool(train)))


class _UNetResPlan(_PlanBase):
    def __init__(self, eng, B, H, W, train):
        if H % 8 or W % 8:
            raise RuntimeError("kair_b200 UNetRes: input %dx%d must be a multiple of 8 (three 2x2 stride-2 stages; the "
                               "reference fails on such sizes too)" % (H, W))
        o, dev, nc, net = eng.ops, eng.device, eng.nc, eng.net
        self.Wt = eng.weights(train)
        Wt = self.Wt
        f32 = lambda *sh: torch.zeros(sh, dtype=torch.float32, device=dev)
        bf = lambda *sh: torch.zeros(sh, dtype=torch.bfloat16, device=dev)
        head, tail = net["head"], net["tail"]
        cin, cout = head.in_channels, tail.out_channels
        self.img, self.out = f32(B, cin, H, W), f32(B, cout, H, W)
        zc, zc_out = f32(cin), f32(cout)
        geos = [(B, H >> l, W >> l) for l in range(4)]
        Ts = [g[0] * g[1] * g[2] for g in geos]
        bns = [_conv_tile(c) for c in nc]

        # fp32 residual streams: two rotating buffers per level + the skip features kept for the up path
        rot = [[f32(Ts[l], nc[l]), f32(Ts[l], nc[l])] for l in range(4)]
        feats = [f32(Ts[l], nc[l]) for l in range(4)]
        saved = []   # per ResBlock (x_b, h_b) in execution order, for the backward pass

        def new_b(l):
            return bf(Ts[l], nc[l])

        # bf16 activation buffers: unique per use when training, small rotating pools otherwise
        pools = [[bf(Ts[l], nc[l]) for _ in range(3)] for l in range(4)] if not train else None
        pool_i = [0, 0, 0, 0]

        def act_b(l):
            if train:
                return new_b(l)
            pool_i[l] = (pool_i[l] + 1) % 3
            return pools[l][pool_i[l]]

        fwd = []
        res_idx = [0, 0, 0, 0]

        def res_block(l, spec, xf, xb):
            """x + conv2(act(conv1(x))); returns (yf, yb).  xf is an fp32 [T,C] buffer, xb its bf16 copy."""
            R = Wt["res"][l]
            k = res_idx[l]
            res_idx[l] += 2
            hb, yb = act_b(l), act_b(l)
            yf = rot[l][0] if xf is not rot[l][0] else rot[l][1]
            b1 = R["bp"][k] if R["bp"] is not None else None
            b2 = R["bp"][k + 1] if R["bp"] is not None else None
            fwd.append(o.build_gemm_tn(xb, R["Wp"][k], N=nc[l], K=nc[l], bn=bns[l], conv=geos[l], bias=b1,
                                       act=o.ACT_LRELU, act_slope=spec.slope, outB=hb))
            fwd.append(o.build_gemm_tn(hb, R["Wp"][k + 1], N=nc[l], K=nc[l], bn=bns[l], conv=geos[l], bias=b2,
                                       resid=xf, outF=yf, outB=yb))
            saved.append((l, k, spec, xb, hb))
            return yf, yb

        xb = act_b(0)
        fwd.append(o.build_conv_in_fwd(self.img, head.weight, head.bias, zc, 1.0, O=nc[0], O_pad=nc[0], pitch=nc[0],
                                       out_f32=feats[0], out_bf16=xb))
        xf = feats[0]
        down_in = []   # regrouped inputs of the strided convs (kept for their weight gradients)
        for l in range(3):
            for spec in net["down"][l][1]:
                xf, xb = res_block(l, spec, xf, xb)
            D = Wt["down"][l]
            a4 = bf(Ts[l + 1], 4 * nc[l])
            down_in.append(a4)
            nb_ = act_b(l + 1)
            fwd.append(o.build_space_to_depth(xb, a4, batch=B, H=geos[l][1], W=geos[l][2], C=nc[l]))
            fwd.append(o.build_gemm_tn(a4, D["Wp"][0], N=nc[l + 1], K=4 * nc[l], bn=bns[l + 1],
                                       bias=D["bp"][0] if D["bp"] is not None else None, outF=feats[l + 1], outB=nb_))
            xf, xb = feats[l + 1], nb_
        for spec in net["body"]:
            xf, xb = res_block(3, spec, xf, xb)
        up_in = [None, None, None]
        for l in (2, 1, 0):
            U = Wt["up"][l]
            sb = act_b(l + 1)
            up_in[l] = sb
            fwd.append(o.build_add_inplace(xf, feats[l + 1], sb))
            y4 = f32(Ts[l + 1], 4 * nc[l])
            fwd.append(o.build_gemm_tn(sb, U["WpT"][0], N=4 * nc[l], K=nc[l + 1], bn=_bn_tile(4 * nc[l]), bias=U["bp"],
                                       outF=y4))
            xf, xb = rot[l][0], act_b(l)
            fwd.append(o.build_depth_to_space(y4, batch=B, H=geos[l][1], W=geos[l][2], C=nc[l], out_f32=xf, out_bf16=xb))
            for spec in net["up"][l][1]:
                xf, xb = res_block(l, spec, xf, xb)
        tail_in = act_b(0)
        fwd.append(o.build_add_inplace(xf, feats[0], tail_in))
        fwd.append(o.build_conv_out_fwd(tail_in, tail.weight, tail.bias, self.out, zc_out, C=nc[0], pitch=nc[0]))
        self.fwd, self.bwd = fwd, None
        self.zero_before_bwd = []
        if not train:
            self._finish(eng)
            return

        # ================================================================================ backward
        self.dout = f32(B, cout, H, W)
        self.grads = _GradStore(eng.P, dev)
        G = lambda p: self.grads.G[eng.name_of(p)] if p is not None else None
        swp = [nc[l] == 64 for l in range(4)]   # 64-channel levels: swapped-role weight gradients (see _DnCNNPlan)
        dWres = [f32(len(Wt["res"][l]["convs"]), 9 * nc[l], nc[l]) if swp[l] else f32(len(Wt["res"][l]["convs"]), nc[l], 9 * nc[l])
                 for l in range(4)]
        dWdown = [f32(1, nc[l + 1], 4 * nc[l]) for l in range(3)]
        dWup = [f32(1, nc[l + 1], 4 * nc[l]) for l in range(3)]
        self.zero_before_bwd = dWres + dWdown + dWup
        gf = [[f32(Ts[l], nc[l]) for _ in range(2)] for l in range(4)]   # rotating fp32 gradient streams
        gb = [[bf(Ts[l], nc[l]) for _ in range(2)] for l in range(4)]
        gskip = [f32(Ts[l], nc[l]) for l in range(4)]                    # gradients of the skip additions
        gskip_b = bf(Ts[0], nc[0])
        dh = [bf(Ts[l], nc[l]) for l in range(4)]
        bwd = []

        def res_block_bwd(entry, cf, cb):
            """(cf, cb): fp32 / bf16 gradient w.r.t. the block output; returns the gradient w.r.t. its input."""
            l, k, spec, xb_, hb_ = entry
            R = Wt["res"][l]
            c1, c2 = R["convs"][k], R["convs"][k + 1]
            nf = gf[l][0] if cf is not gf[l][0] else gf[l][1]
            nb2 = gb[l][0] if cb is not gb[l][0] else gb[l][1]
            bwd.append(o.build_gemm_nt(cb, hb_, dWres[l][k + 1], n1=nc[l], n2=nc[l], bn=_bn_tile(nc[l]), conv=geos[l],
                                       swap=swp[l]))
            if c2.bias is not None:
                bwd.append(o.build_colsum(cb, G(c2.bias), C=nc[l]))
            bwd.append(o.build_gemm_tn(cb, R["WpT"][k + 1], N=nc[l], K=nc[l], bn=bns[l], conv=geos[l], outB=dh[l],
                                       aux=hb_, act_slope=spec.slope))
            bwd.append(o.build_gemm_nt(dh[l], xb_, dWres[l][k], n1=nc[l], n2=nc[l], bn=_bn_tile(nc[l]), conv=geos[l],
                                       swap=swp[l]))
            if c1.bias is not None:
                bwd.append(o.build_colsum(dh[l], G(c1.bias), C=nc[l]))
            bwd.append(o.build_gemm_tn(dh[l], R["WpT"][k], N=nc[l], K=nc[l], bn=bns[l], conv=geos[l], resid=cf, outF=nf,
                                       outB=nb2))
            return nf, nb2

        # tail: d(tail_in) feeds both the up-path output and the head feature (skip)
        dsh, clT, dWcl, csum = bf(Ts[0], 64), bf(nc[0], 64), f32(nc[0], 64), f32(64)
        xcol, dWcf = bf(Ts[0], 64), f32(1, nc[0], 64)
        self.zero_before_bwd += [dWcl, csum, dWcf]
        bwd.append(o.build_im2col3x3(self.dout, dsh, cpad=7, direction=-1, scale=1.0))
        bwd.append(o.build_conv_out_pack(mode=0, co=cout, C=nc[0], cpad=7, W=tail.weight, Wt=clT))
        bwd.append(o.build_gemm_nt(tail_in, dsh, dWcl, n1=nc[0], n2=64, bn=64))
        bwd.append(o.build_conv_out_pack(mode=1, co=cout, C=nc[0], cpad=7, D=dWcl, dW=G(tail.weight)))
        if tail.bias is not None:
            bwd.append(o.build_colsum(dsh, csum, C=64))
            bwd.append(o.build_copy(G(tail.bias), csum[4 * 7:4 * 7 + cout]))
        bwd.append(o.build_gemm_tn(dsh, clT, N=nc[0], K=64, bn=bns[0], outF=gskip[0], outB=gskip_b))
        cf, cb = gskip[0], gskip_b
        entries = list(saved)
        # index ranges of `saved`: down path blocks per level, body, up path blocks per level (3 -> 1)
        pos = len(entries)
        for l in (0, 1, 2):
            n_up = len(net["up"][l][1])
            for e in reversed(entries[pos - n_up:pos]):
                cf, cb = res_block_bwd(e, cf, cb)
            pos -= n_up
            # transposed conv backward: regroup the fine gradient, contract with the packed weight
            U = Wt["up"][l]
            g4 = bf(Ts[l + 1], 4 * nc[l])
            bwd.append(o.build_space_to_depth(cb, g4, batch=B, H=geos[l][1], W=geos[l][2], C=nc[l]))
            if U["conv"].bias is not None:
                bwd.append(o.build_colsum(cb, G(U["conv"].bias), C=nc[l]))
            bwd.append(o.build_gemm_nt(up_in[l], g4, dWup[l][0], n1=nc[l + 1], n2=4 * nc[l], bn=_bn_tile(4 * nc[l])))
            bwd.append(o.build_gemm_tn(g4, U["Wp"][0], N=nc[l + 1], K=4 * nc[l], bn=bns[l + 1], outF=gskip[l + 1],
                                       outB=gb[l + 1][0]))
            cf, cb = gskip[l + 1], gb[l + 1][0]
        for e in reversed(entries[pos - len(net["body"]):pos]):
            cf, cb = res_block_bwd(e, cf, cb)
        pos -= len(net["body"])
        for l in (2, 1, 0):
            # total gradient of feats[l+1] = path through the blocks above + the skip addition
            bwd.append(o.build_add_inplace(cf, gskip[l + 1], cb))
            D = Wt["down"][l]
            bwd.append(o.build_gemm_nt(cb, down_in[l], dWdown[l][0], n1=nc[l + 1], n2=4 * nc[l], bn=_bn_tile(4 * nc[l])))
            if D["conv"].bias is not None:
                bwd.append(o.build_colsum(cb, G(D["conv"].bias), C=nc[l + 1]))
            d4 = f32(Ts[l + 1], 4 * nc[l])
            bwd.append(o.build_gemm_tn(cb, D["WpT"][0], N=4 * nc[l], K=nc[l + 1], bn=_bn_tile(4 * nc[l]), outF=d4))
            cf, cb = gf[l][0], gb[l][0]
            bwd.append(o.build_depth_to_space(d4, batch=B, H=geos[l][1], W=geos[l][2], C=nc[l], out_f32=cf, out_bf16=cb))
            n_dn = len(net["down"][l][1])
            for e in reversed(entries[pos - n_dn:pos]):
                cf, cb = res_block_bwd(e, cf, cb)
            pos -= n_dn
        assert pos == 0
        bwd.append(o.build_add_inplace(cf, gskip[0], cb))
        bwd.append(o.build_im2col3x3(self.img, xcol, cpad=7, direction=1, mean=zc, scale=1.0, ones_col=4 * 7 + cin))
        bwd.append(o.build_gemm_nt(cb, xcol, dWcf[0], n1=nc[0], n2=64, bn=64))
        bwd.append(o.build_unpack_conv_grad(dWcf, [G(head.weight)], [G(head.bias)] if head.bias is not None else None,
                                            O=nc[0], I=cin, I_pad=7, ones_col=cin, center_tap=4, dwp_pitch=64))
        for l in range(4):
            convs = Wt["res"][l]["convs"]
            if convs:
                bwd.append(o.build_unpack_conv_grad(dWres[l], [G(c.weight) for c in convs], None, O=nc[l], I=nc[l],
                                                    I_pad=nc[l], transposed=swp[l]))
        for l in range(3):
            D, U = Wt["down"][l], Wt["up"][l]
            Ci, Co = nc[l], nc[l + 1]
            bwd.append(o.build_unpack_linear_grad(dWdown[l], [D["conv"].weight.detach().view(Co, Ci * 4)], None, None,
                                                  [G(D["conv"].weight).view(Co, Ci * 4)], None, None, None, D["rows"],
                                                  D["cols"], None, out_f=Co, in_f=4 * Ci, ones_col=-1))
            bwd.append(o.build_unpack_linear_grad(dWup[l], [U["conv"].weight.detach().view(Co, Ci * 4)], None, None,
                                                  [G(U["conv"].weight).view(Co, Ci * 4)], None, None, None, U["rows"],
                                                  U["cols"], None, out_f=Co, in_f=4 * Ci, ones_col=-1))
        self.bwd = bwd
        self._dx_spec = dict(dhead=cf, W=head.weight, C=nc[0], pitch=nc[0], scale=1.0, skip=False)
        self._finish(eng)
