"""Launch-list engine for RRDBNet_noUp (models/network_rrdbnet.py:35-125 of the reference; the fork trains it as
`rrdbnet_no_up`, SPECToptions/network_configs.json:16-31), forward and backward, on the sm_100a kernels.

A ResidualDenseBlock_5C convolves growing concatenations cat(x, x1, .., xk) (:49-56).  Here a dense block owns ONE token
buffer D [T][256] bf16 whose columns are the concatenation itself -- x in [0, 64), x1 in [64, 96), x2 in [96, 128),
x3 in [128, 160), x4 in [160, 192) -- so torch.cat never happens: conv_k is the implicit-GEMM 3x3 (kair_gemm_tn conv
mode) over the first 64 + 32(k-1) columns (K rounded up to a multiple of 64 with zero weight columns) and its
LeakyReLU(0.2) epilogue writes x_k straight into its column slot (as a 64-column store whose upper 32 columns are
exact zeros and are overwritten by x_{k+1}).  conv5's epilogue is "x5 * 0.2 + x" (acc_scale = 0.2, fp32 residual) and
writes the next block's x (fp32) and its bf16 copy into the next block's buffer.  "out * 0.2 + x" of an RRDB (:67-72) is
one kair_axpby launch.

Backward keeps the gradient of the concatenation in an fp32 buffer dD [T][192]: conv5's input-gradient GEMM fills it,
then for k = 4..1 the slot of x_k is cut out, multiplied by LeakyReLU'(x_k) (kair_axpby with the saved activation as
mask), fed to conv_k's weight-gradient GEMM (kair_gemm_nt) and to its input-gradient GEMM, whose epilogue adds the
running dD (fp32 residual operand) -- the "+=" of autograd's cat backward without atomics.
"""
import torch
import torch.nn as nn

from .convnet_engine import _EngineBase, _GradStore, _PlanBase, _unsupported

NF, GC, DP = 64, 32, 256          # features, growth channels, pitch of a dense buffer
SLOT = [NF + GC * k for k in range(5)]                  # column of x1..x4 (index 0..3), 192 = end
KPAD = [64, 128, 128, 192, 192]                          # K of conv1..conv5 (inputs 64, 96, 128, 160, 192)


def parse_rrdbnet(net):
    convs = ("conv1", "conv2", "conv3", "conv4", "conv5")
    blocks = []
    for rrdb in net.RRDB_trunk:
        for name in ("RDB1", "RDB2", "RDB3"):
            rdb = getattr(rrdb, name)
            cs = [getattr(rdb, c) for c in convs]
            for k, c in enumerate(cs):
                want_in, want_out = NF + GC * k, (GC if k < 4 else NF)
                if not (isinstance(c, nn.Conv2d) and c.kernel_size == (3, 3) and c.padding == (1, 1) and
                        c.in_channels == want_in and c.out_channels == want_out and c.bias is not None):
                    _unsupported("RRDBNet dense block convolution %r (nf=64, gc=32, bias=True is supported)" % (c,))
            blocks.append(cs)
    for c in (net.conv_first, net.trunk_conv, net.HRconv, net.conv_last):
        if c.bias is None:
            _unsupported("RRDBNet head / tail convolutions without bias")
    if net.conv_first.in_channels > 4 or net.conv_last.out_channels > 4 or net.conv_first.out_channels != NF:
        _unsupported("RRDBNet with nf != 64 or more than 4 image channels")
    return blocks


class RRDBEngine(_EngineBase):
    def __init__(self, net_children, blocks, P, device, ops=None):
        super().__init__(P, device, ops)
        self.net = net_children          # dict: conv_first, trunk_conv, HRconv, conv_last
        self.blocks = blocks             # [3 * nb][5] convolutions
        self._w = {}

    def weights(self, train):
        W = self._w.get(train) or self._w.get(True)
        if W is not None:
            return W
        o, dev = self.ops, self.device
        bf = lambda *s: torch.zeros(s, dtype=torch.bfloat16, device=dev)
        f32 = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        L = len(self.blocks)
        W = {"train": train, "pack": [], "versions": None, "buffers": []}
        for k in range(5):
            O = GC if k < 4 else NF
            W["Wp%d" % k], W["bp%d" % k] = bf(L, 64, 9 * KPAD[k]), f32(L, 64)
            W["WpT%d" % k] = bf(L, KPAD[k], 9 * 64) if train else None
            cs = [blk[k] for blk in self.blocks]
            W["pack"].append(o.build_pack_conv([c.weight for c in cs], [c.bias for c in cs], W["Wp%d" % k],
                                               W["WpT%d" % k], W["bp%d" % k], O=O, I=NF + GC * k, I_pad=KPAD[k]))
        tw = [self.net["trunk_conv"], self.net["HRconv"]]
        W["Wt"], W["bt"], W["WtT"] = bf(2, 64, 9 * 64), f32(2, 64), (bf(2, 64, 9 * 64) if train else None)
        W["pack"].append(o.build_pack_conv([c.weight for c in tw], [c.bias for c in tw], W["Wt"], W["WtT"], W["bt"],
                                           O=NF, I=NF, I_pad=NF))
        self._w[train] = W
        return W

    def refresh_weights(self, plan, stream):
        W = plan.Wt
        v = self._versions()
        if W["versions"] != v:
            for l in W["pack"]:
                l.run(stream)
            W["versions"] = v

    def plan(self, B, H, W, train):
        return self._get_plan((B, H, W, bool(train)), lambda: _RRDBPlan(self, B, H, W, bool(train)))


class _RRDBPlan(_PlanBase):
    def __init__(self, eng, B, H, W, train):
        o, dev = eng.ops, eng.device
        self.Wt = Wt = eng.weights(train)
        blocks, net = eng.blocks, eng.net
        L = len(blocks)
        T, geo, tps = B * H * W, (B, H, W), H * W
        head, trunk, hr, tail = net["conv_first"], net["trunk_conv"], net["HRconv"], net["conv_last"]
        cin, cout = head.in_channels, tail.out_channels
        f32 = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        bf = lambda *s: torch.zeros(s, dtype=torch.bfloat16, device=dev)
        self.img, self.out = f32(B, cin, H, W), f32(B, cout, H, W)
        zc_in, zc_out = f32(cin), f32(cout)
        c02 = torch.full((B,), 0.2, dtype=torch.float32, device=dev)     # "x5 * 0.2" as a per-sample epilogue scale
        # one dense buffer per block when training (saved for the backward pass), two alternating ones otherwise;
        # +1: the buffer that receives the trunk output
        nD = L + 1 if train else 2
        Ds = [bf(T, DP) for _ in range(nD)]
        D = lambda i: Ds[i if train else i % 2]
        fea = f32(T, NF)
        xs = [f32(T, NF), f32(T, NF), f32(T, NF)]     # residual stream (fp32): two rotating + the RRDB input
        feat2, hrb = bf(T, NF), bf(T, NF)

        fwd = []
        fwd.append(o.build_conv_in_fwd(self.img, head.weight, head.bias, zc_in, 1.0, O=NF, O_pad=NF, pitch=NF, out_f32=fea))
        fwd.append(o.build_axpby(fea, n=NF, out_bf16=D(0)))
        x_cur = fea
        for i in range(L):
            Di, Dn = D(i), D(i + 1)
            if i % 3 == 0:
                x_rrdb = x_cur                       # input of this RRDB (kept until its closing axpby)
            for k in range(4):
                fwd.append(o.build_gemm_tn(Di, Wt["Wp%d" % k][i], N=64, K=KPAD[k], bn=64, conv=geo, bias=Wt["bp%d" % k][i],
                                           act=o.ACT_LRELU, act_slope=0.2, outB=Di[:, SLOT[k]:]))
            # conv5: x5 * 0.2 + x -> the next block's x (fp32) and its bf16 copy in the next dense buffer
            x_new = [b_ for b_ in xs if b_ is not x_cur and b_ is not x_rrdb][0]
            last_of_rrdb = i % 3 == 2
            fwd.append(o.build_gemm_tn(Di, Wt["Wp4"][i], N=64, K=KPAD[4], bn=64, conv=geo, bias=Wt["bp4"][i], resid=x_cur,
                                       outF=x_new, outB=None if last_of_rrdb else Dn, acc_scale=c02, tokens_per_sample=tps))
            x_cur = x_new
            if last_of_rrdb:
                # out * 0.2 + x of the RRDB
                x_out = [b_ for b_ in xs if b_ is not x_cur and b_ is not x_rrdb][0]
                fwd.append(o.build_axpby(x_cur, n=NF, a=0.2, y=x_rrdb, b=1.0, out_f32=x_out, out_bf16=Dn))
                x_cur = x_out
        # trunk_conv + "fea + trunk", HRconv + LeakyReLU, conv_last
        fwd.append(o.build_gemm_tn(D(L), Wt["Wt"][0], N=64, K=64, bn=64, conv=geo, bias=Wt["bt"][0], resid=fea, outB=feat2))
        fwd.append(o.build_gemm_tn(feat2, Wt["Wt"][1], N=64, K=64, bn=64, conv=geo, bias=Wt["bt"][1], act=o.ACT_LRELU,
                                   act_slope=0.2, outB=hrb))
        fwd.append(o.build_conv_out_fwd(hrb, tail.weight, tail.bias, self.out, zc_out, C=NF, pitch=NF, out_scale=1.0))
        self.fwd, self.bwd = fwd, None
        self.zero_before_bwd = []
        if not train:
            self._finish(eng)
            return

        # ================================================================================ backward
        self.dout = f32(B, cout, H, W)
        self.grads = _GradStore(eng.P, dev)
        G = lambda p: self.grads.G[eng.name_of(p)]
        # every convolution here has 64 (padded) output channels: weight gradients with the operand roles swapped
        # (kair_gemm_nt `swap`: M = (tap, input channel)), packed gradients transposed [9 * I_pad, 64]
        dW = [f32(L, 9 * KPAD[k], 64) for k in range(5)]
        dWt = f32(2, 9 * 64, 64)
        dsh, clT, dWcl, csum = bf(T, 64), bf(NF, 64), f32(NF, 64), f32(64)
        xcol, dWcf = bf(T, 64), f32(1, NF, 64)
        self.zero_before_bwd = dW + [dWt, dWcl, csum, dWcf]
        bwd = []
        # conv_last (thin): 9-tap expansion of d(out) serves its weight gradient and its input gradient
        bwd.append(o.build_im2col3x3(self.dout, dsh, cpad=7, direction=-1, scale=1.0))
        bwd.append(o.build_conv_out_pack(mode=0, co=cout, C=NF, cpad=7, W=tail.weight, Wt=clT))
        bwd.append(o.build_gemm_nt(hrb, dsh, dWcl, n1=NF, n2=64, bn=64))
        bwd.append(o.build_conv_out_pack(mode=1, co=cout, C=NF, cpad=7, D=dWcl, dW=G(tail.weight)))
        bwd.append(o.build_colsum(dsh, csum, C=64))
        bwd.append(o.build_copy(G(tail.bias), csum[4 * 7:4 * 7 + cout]))
        g_hr, g_f2 = bf(T, NF), bf(T, NF)
        bwd.append(o.build_gemm_tn(dsh, clT, N=NF, K=64, bn=64, outB=g_hr, aux=hrb, act_slope=0.2))   # * lrelu'(HRconv)
        # HRconv
        bwd.append(o.build_colsum(g_hr, G(hr.bias), C=NF))
        bwd.append(o.build_gemm_nt(g_hr, feat2, dWt[1], n1=NF, n2=NF, bn=64, conv=geo, swap=True))
        # d(feat2) = d(fea + trunk): fp32 copy in Z (gradient of the residual stream, [T][192] with zero upper columns)
        Z = [f32(T, 192), f32(T, 192)]
        dDs = [f32(T, 192), f32(T, 192)]
        gsave, dfea = f32(T, NF), f32(T, NF)
        g5, gk = bf(T, NF), bf(T, NF)
        bwd.append(o.build_gemm_tn(g_hr, Wt["WtT"][1], N=NF, K=NF, bn=64, conv=geo, outF=dfea, outB=g_f2))
        # trunk_conv: fea receives d(feat2) directly (dfea) and, at the very end, the gradient through the trunk
        bwd.append(o.build_colsum(g_f2, G(trunk.bias), C=NF))
        bwd.append(o.build_gemm_nt(g_f2, D(L), dWt[0], n1=NF, n2=NF, bn=64, conv=geo, swap=True))
        z = 0
        bwd.append(o.build_gemm_tn(g_f2, Wt["WtT"][0], N=NF, K=NF, bn=64, conv=geo, outF=Z[z][:, :NF]))
        for i in reversed(range(L)):
            Di = D(i)
            if i % 3 == 2:
                # RRDB output = 0.2 * y3 + x0: keep g_out for the skip, continue with 0.2 * g_out
                bwd.append(o.build_axpby(Z[z], n=NF, out_f32=gsave))
                bwd.append(o.build_axpby(Z[z], n=NF, a=0.2, out_f32=Z[z]))
            # out = 0.2 * x5 + x:  d(x5) = 0.2 * g,  the skip stays in Z[z][:, :64]
            bwd.append(o.build_axpby(Z[z], n=NF, a=0.2, out_bf16=g5))
            bwd.append(o.build_colsum(g5, G(blocks[i][4].bias), C=NF))
            bwd.append(o.build_gemm_nt(g5, Di, dW[4][i], n1=64, n2=KPAD[4], bn=64, conv=geo, swap=True))
            d = 0
            bwd.append(o.build_gemm_tn(g5, Wt["WpT4"][i], N=KPAD[4], K=64, bn=KPAD[4], conv=geo, resid=Z[z], outF=dDs[d]))
            for k in (3, 2, 1, 0):
                # gradient of x_{k+1}'s pre-activation: its slot of dD times LeakyReLU'
                bwd.append(o.build_axpby(dDs[d][:, SLOT[k]:], n=GC, out_bf16=gk, mask=Di[:, SLOT[k]:], mask_slope=0.2, n_pad=NF))
                bwd.append(o.build_colsum(gk, G(blocks[i][k].bias), C=GC))
                bwd.append(o.build_gemm_nt(gk, Di, dW[k][i], n1=64, n2=KPAD[k], bn=64, conv=geo, swap=True))
                if k > 0:
                    bwd.append(o.build_gemm_tn(gk, Wt["WpT%d" % k][i], N=KPAD[k], K=64, bn=KPAD[k], conv=geo, resid=dDs[d],
                                               outF=dDs[d ^ 1]))
                    d ^= 1
                else:
                    # conv1 closes the block: its input gradient + the running dD[:, :64] is d(x) of this block
                    bwd.append(o.build_gemm_tn(gk, Wt["WpT0"][i], N=64, K=64, bn=64, conv=geo, resid=dDs[d][:, :NF],
                                               outF=Z[z ^ 1][:, :NF]))
                    z ^= 1
            if i % 3 == 0:
                bwd.append(o.build_axpby(Z[z], n=NF, y=gsave, b=1.0, out_f32=Z[z]))     # + skip of the RRDB
        # fea = conv_first(x) feeds the trunk and the global skip
        dfirst_b = bf(T, NF)
        bwd.append(o.build_axpby(Z[z], n=NF, y=dfea, b=1.0, out_f32=dfea, out_bf16=dfirst_b))
        bwd.append(o.build_im2col3x3(self.img, xcol, cpad=7, direction=1, mean=zc_in, scale=1.0, ones_col=4 * 7 + cin))
        bwd.append(o.build_gemm_nt(dfirst_b, xcol, dWcf[0], n1=NF, n2=64, bn=64))
        bwd.append(o.build_unpack_conv_grad(dWcf, [G(head.weight)], [G(head.bias)], O=NF, I=cin, I_pad=7, ones_col=cin,
                                            center_tap=4, dwp_pitch=64))
        for k in range(5):
            cs = [blk[k] for blk in blocks]
            bwd.append(o.build_unpack_conv_grad(dW[k], [G(c.weight) for c in cs], None, O=GC if k < 4 else NF,
                                                I=NF + GC * k, I_pad=KPAD[k], transposed=True))
        bwd.append(o.build_unpack_conv_grad(dWt, [G(trunk.weight), G(hr.weight)], None, O=NF, I=NF, I_pad=NF,
                                            transposed=True))
        self.bwd = bwd
        self._dx_spec = dict(dhead=dfea, W=head.weight, C=NF, pitch=NF, scale=1.0, skip=False)
        self._finish(eng)
