"""Launch-list engine for SwinIR (forward and backward) on the sm_100a kernels.

The engine turns one (batch, H, W) problem into a *plan*: a statically allocated HBM workspace plus two ordered
lists of pre-built kernel launches (forward, backward).  Running a step is replaying a list -- one ctypes call per
kernel, no tensor allocation, no host synchronisation -- so the plan is also CUDA-graph capturable.

Data layout in HBM: activations are token-major NHWC, t = (b*H + h)*W + w, channels padded
C -> Cp (180 -> 192); bf16 for every tensor-core operand, fp32 for the residual stream and all statistics.
q/k/v and the attention output use a per-head pitch of 32 (head_dim 30 -> 32).  Spare padded columns carry a
constant 1.0 ("ones column") so that each weight-gradient GEMM also produces the bias gradient.

Reference arithmetic replaced (file:line in /root/reference): models/network_swinir.py:805-839 (SwinIR.forward),
:790-803 (forward_features), :481-482 (RSTB.forward), :239-279 (SwinTransformerBlock.forward), :114-145
(WindowAttention.forward), :24-30 (Mlp.forward), and autograd's backward of all of them.
"""
import math
import os

import torch

from . import _dirty
from . import ops as _real_ops

# fc1 + GELU + fc2 + residual (+ LayerNorm) as one launch (kair_mlp_fused) instead of two kair_gemm_tn launches;
# KAIR_B200_FUSE_MLP=0 keeps the unfused pair (A/B measurements; the results are bit-identical)
FUSE_MLP = os.environ.get("KAIR_B200_FUSE_MLP", "1") != "0"


def _tile(n, fused=False):
    """N tile of a GEMM with n output columns.  The fused epilogues (GELU, GELU', LayerNorm forward / backward) are
    instantiated at 192 (SwinIR-M, C=180) and 64 (SwinIR-S, C=60); plain LINEAR and the weight-gradient GEMMs at
    64 / 128 / 192 / 256."""
    for bn in ((256, 192, 64) if fused else (256, 192, 128, 64)):
        if n % bn == 0:
            return bn
    raise NotImplementedError("kair_b200 SwinIR: no GEMM tile for %d columns" % n)


def _round_up(x, m):
    return (x + m - 1) // m * m


_forward_clock = [0]


def pick_plan(plans, key, max_shapes, make):
    """The plan serving `key`.  A training plan owns ONE set of saved activations, so while its backward is pending
    (`busy`, set by the autograd Function) a second forward of the same shape gets a second plan instance instead of
    overwriting them: `y1 = m(x1); y2 = m(x2); (l(y1) + l(y2)).backward()` works like it does for the reference.  At
    most two instances per shape are kept; a third pending forward reuses the older one, whose own backward then
    raises (generation check) instead of returning gradients of the wrong activations."""
    slots = plans.get(key)
    if slots is None:
        if len(plans) >= max_shapes:   # bound the workspace kept alive by stale shapes
            plans.pop(next(iter(plans)))
        slots = plans[key] = []
    for p in slots:
        if not getattr(p, "busy", False):
            return p
    if len(slots) < 2:
        p = make()
        p.busy = False
        slots.append(p)
        return p
    return min(slots, key=lambda q: getattr(q, "busy_since", 0))


def mark_busy(plan):
    _forward_clock[0] += 1
    plan.busy, plan.busy_since = True, _forward_clock[0]


class SwinIRSpec:
    """The constructor arguments of models/network_swinir.py:646-652 that shape the computation."""

    def __init__(self, *, in_chans, embed_dim, depths, num_heads, window_size, mlp_ratio, qk_scale, img_range,
                 upscale, upsampler, resi_connection, patch_norm, qkv_bias):
        self.in_chans = in_chans
        self.C = embed_dim
        self.depths = tuple(depths)
        self.num_heads = tuple(num_heads)
        self.ws = window_size
        self.hidden = int(embed_dim * mlp_ratio)
        self.qk_scale = qk_scale
        self.img_range = float(img_range)
        self.upscale = upscale
        self.upsampler = upsampler or ""
        self.resi_connection = resi_connection
        self.patch_norm = patch_norm
        self.qkv_bias = qkv_bias

    def check_supported(self):
        """Loud errors for configurations the sm_100a kernels do not cover yet (no fallback path exists)."""
        def bad(msg):
            raise NotImplementedError("kair_b200 SwinIR: " + msg + " (see DESIGN.md, 'coverage')")
        if not 2 <= self.ws <= 8:
            bad("window_size=%d; window attention kernels exist for 2..8 (8: tcgen05, smaller: CUDA cores)" % self.ws)
        if len(set(self.num_heads)) != 1:
            bad("num_heads must be the same in every RSTB")
        nH = self.num_heads[0]
        if nH % 2 or nH > 8 or self.C % nH:
            bad("num_heads=%d must be even, <= 8 and divide embed_dim" % nH)
        hd = self.C // nH
        if hd >= 32:
            bad("head_dim=%d must be < 32 (one padded column carries the proj bias gradient)" % hd)
        if _round_up(self.C + 1, 64) not in (64, 192, 256):
            bad("embed_dim=%d: the LayerNorm epilogues are instantiated for 64, 192 and 256 padded channels "
                "(embed_dim 1..63 and 128..255, e.g. SwinIR-S 60, SwinIR-M 180, SwinIR-L 240)" % self.C)
        if self.resi_connection not in ("1conv", "3conv"):
            bad("resi_connection='%s'" % self.resi_connection)
        if self.upsampler not in ("", "pixelshuffle", "pixelshuffledirect", "nearest+conv"):
            bad("upsampler='%s'" % self.upsampler)
        if self.upsampler == "pixelshuffledirect" and not (1 <= self.upscale <= 8):
            bad("pixelshuffledirect upscale=%d" % self.upscale)
        if self.upsampler == "pixelshuffle" and not (self.upscale in (2, 3, 4, 8)):
            bad("pixelshuffle upscale=%d" % self.upscale)
        if self.in_chans > 4:
            bad("in_chans=%d > 4" % self.in_chans)


class _Blk:
    """Activations of one Swin block kept for the backward pass (shared across blocks when not training)."""

    def __init__(self, T, g, dev):
        bf = lambda n: torch.zeros(T, n, dtype=torch.bfloat16, device=dev)
        self.xh1 = bf(g.Cp)
        self.r1 = torch.zeros(T, dtype=torch.float32, device=dev)
        self.qkv = bf(g.QP)
        self.att = bf(g.HP)
        self.xh2 = bf(g.Cp)
        self.r2 = torch.zeros(T, dtype=torch.float32, device=dev)
        self.u = bf(g.Fp)
        self.g = bf(g.Fp)


class SwinIREngine:
    def __init__(self, spec, params, device, ops=None):
        """params: ordered {state_dict key: nn.Parameter} of the (reference-layout) module."""
        spec.check_supported()
        self.spec = spec
        self.ops = ops or _real_ops
        self.device = torch.device(device)
        self.P = params
        s = spec
        self.C, self.nH = s.C, s.num_heads[0]
        self.hd = s.C // self.nH
        self.Cp = _round_up(s.C + 1, 64)
        self.HP = self.nH * 32
        self.QP = 3 * self.HP
        self.Fp = _round_up(s.hidden + 1, 64)
        self.L = sum(s.depths)
        self.nl = len(s.depths)
        self.scale = s.qk_scale or self.hd ** -0.5
        self.mean = torch.tensor((0.4488, 0.4371, 0.4040) if s.in_chans == 3 else (0.0,) * s.in_chans,
                                 dtype=torch.float32, device=self.device)
        self._plans = {}
        self._weights = None
        self._plans_fp32 = {}
        self._weights_fp32 = None
        self.grad_hook = None   # callable(flat fp32 gradient buffer), e.g. parallel.allreduce_flat bound to the module
        self._build_maps()

    # ------------------------------------------------------------------------------------------ packing maps
    def _build_maps(self):
        C, Cp, nH, hd, HP, QP, Fp, hid = self.C, self.Cp, self.nH, self.hd, self.HP, self.QP, self.Fp, self.spec.hidden
        dev = self.device
        i32 = lambda v: torch.tensor(v, dtype=torch.int32, device=dev)
        f32 = lambda v: torch.tensor(v, dtype=torch.float32, device=dev)
        self.map_c = i32([c if c < C else -1 for c in range(Cp)])
        self.map_f = i32([c if c < hid else -1 for c in range(Fp)])
        head = [(c // 32) * hd + c % 32 if c % 32 < hd else -1 for c in range(HP)]
        self.map_h = i32(head)
        qkv_rows, qkv_scale, qkv_const = [], [], []
        for t3 in range(3):
            for c in range(HP):
                src = head[c]
                qkv_rows.append(t3 * C + src if src >= 0 else -1)
                qkv_scale.append(self.scale if t3 == 0 else 1.0)
                # v of head 0 carries a constant 1.0 in its first padded column: softmax rows sum to one, so the
                # attention output holds 1.0 there and the proj weight-gradient GEMM yields d(proj.bias)
                qkv_const.append(1.0 if (t3 == 2 and c == hd) else 0.0)
        self.map_qkv, self.scale_qkv, self.const_qkv = i32(qkv_rows), f32(qkv_scale), f32(qkv_const)

    # ------------------------------------------------------------------------------------------ packed weights
    def _block_names(self):
        names = []
        for i, d in enumerate(self.spec.depths):
            for j in range(d):
                names.append("layers.%d.residual_group.blocks.%d." % (i, j))
        return names

    def _weight_state(self, train):
        """Packed bf16 operands (+ the launches that refresh them from the fp32 master parameters)."""
        if self._weights is not None and (self._weights["train"] or not train):
            return self._weights
        o, P, dev = self.ops, self.P, self.device
        L, nl, Cp, HP, QP, Fp, C = self.L, self.nl, self.Cp, self.HP, self.QP, self.Fp, self.C
        bf = lambda *s: torch.zeros(s, dtype=torch.bfloat16, device=dev)
        f32 = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        blk = self._block_names()
        W = {"train": train}
        W["qkv"], W["qkvT"], W["bqkv"] = bf(L, QP, Cp), (bf(L, Cp, QP) if train else None), f32(L, QP)
        W["proj"], W["projT"], W["bproj"] = bf(L, Cp, HP), (bf(L, HP, Cp) if train else None), f32(L, Cp)
        W["fc1"], W["fc1T"], W["bfc1"] = bf(L, Fp, Cp), (bf(L, Cp, Fp) if train else None), f32(L, Fp)
        W["fc2"], W["fc2T"], W["bfc2"] = bf(L, Cp, Fp), (bf(L, Fp, Cp) if train else None), f32(L, Cp)
        three = self.spec.resi_connection == "3conv"
        Cq, Cqp = C // 4, _round_up(C // 4 + 1, 64)
        Kc = Cqp if three else Cp   # input channels of the convolution that closes an RSTB / the body
        W["conv"], W["convT"], W["bconv"] = bf(nl + 1, Cp, 9 * Kc), (bf(nl + 1, Kc, 9 * Cp) if train else None), \
            f32(nl + 1, Cp)
        get = lambda suffix: [P[b + suffix] for b in blk]
        qb = get("attn.qkv.bias") if self.spec.qkv_bias else None
        conv_names = ["layers.%d.conv" % i for i in range(nl)] + ["conv_after_body"]
        W["conv_names"] = conv_names
        W["pack"] = [
            o.build_pack_linear(get("attn.qkv.weight"), qb, get("norm1.weight"), get("norm1.bias"), self.map_qkv,
                                self.map_c, self.scale_qkv, self.const_qkv, W["qkv"], W["qkvT"], W["bqkv"],
                                out_f=3 * C, in_f=C),
            o.build_pack_linear(get("attn.proj.weight"), get("attn.proj.bias"), None, None, self.map_c, self.map_h,
                                None, None, W["proj"], W["projT"], W["bproj"], out_f=C, in_f=C),
            o.build_pack_linear(get("mlp.fc1.weight"), get("mlp.fc1.bias"), get("norm2.weight"), get("norm2.bias"),
                                self.map_f, self.map_c, None, None, W["fc1"], W["fc1T"], W["bfc1"],
                                out_f=self.spec.hidden, in_f=C),
            o.build_pack_linear(get("mlp.fc2.weight"), get("mlp.fc2.bias"), None, None, self.map_c, self.map_f,
                                None, None, W["fc2"], W["fc2T"], W["bfc2"], out_f=C, in_f=self.spec.hidden),
        ]
        pw = lambda sfx: [P[n + sfx + ".weight"] for n in conv_names]
        pb = lambda sfx: [P[n + sfx + ".bias"] for n in conv_names]
        if not three:
            W["pack"].append(o.build_pack_conv(pw(""), pb(""), W["conv"], W["convT"], W["bconv"], O=C, I=C, I_pad=Cp))
        else:
            # '3conv' (network_swinir.py:467-471, 733-737): conv3x3 C -> C/4, LeakyReLU(0.2), conv1x1, LeakyReLU(0.2),
            # conv3x3 C/4 -> C; the last one keeps the single-conv epilogues ("conv"), the first two are c3a / c3b
            W["c3a"], W["c3aT"], W["bc3a"] = bf(nl + 1, Cqp, 9 * Cp), (bf(nl + 1, Cp, 9 * Cqp) if train else None), \
                f32(nl + 1, Cqp)
            W["c3b"], W["c3bT"], W["bc3b"] = bf(nl + 1, Cqp, Cqp), (bf(nl + 1, Cqp, Cqp) if train else None), \
                f32(nl + 1, Cqp)
            W["pack"] += [
                o.build_pack_conv(pw(".0"), pb(".0"), W["c3a"], W["c3aT"], W["bc3a"], O=Cq, I=C, I_pad=Cp),
                o.build_pack_conv(pw(".2"), pb(".2"), W["c3b"], W["c3bT"], W["bc3b"], O=Cq, I=Cq, I_pad=Cqp, taps=1),
                o.build_pack_conv(pw(".4"), pb(".4"), W["conv"], W["convT"], W["bconv"], O=C, I=Cq, I_pad=Cqp),
            ]
        # reconstruction heads (network_swinir.py:740-760): num_feat = 64
        up = self.spec.upsampler
        if up == "pixelshuffledirect":
            # UpsampleOneStep (network_swinir.py:594-609): one conv embed_dim -> upscale^2 * num_out_ch, rows padded to 64
            no = self.spec.upscale ** 2 * self.spec.in_chans
            nop = _round_up(no, 64)
            W["psd"], W["bpsd"], W["psdT"] = bf(1, nop, 9 * Cp), f32(1, nop), (bf(1, Cp, 9 * nop) if train else None)
            W["pack"].append(o.build_pack_conv([P["upsample.0.weight"]], [P["upsample.0.bias"]],
                                               W["psd"], W["psdT"], W["bpsd"], O=no, I=C, I_pad=Cp))
        elif up:
            nf = 64
            W["cbu"], W["bcbu"], W["cbuT"] = bf(1, nf, 9 * Cp), f32(1, nf), (bf(1, Cp, 9 * nf) if train else None)
            W["pack"].append(o.build_pack_conv([P["conv_before_upsample.0.weight"]], [P["conv_before_upsample.0.bias"]],
                                               W["cbu"], W["cbuT"], W["bcbu"], O=nf, I=C, I_pad=Cp))
            if up == "pixelshuffle":
                s = self.spec.upscale
                r = 3 if s == 3 else 2
                n = 1 if s == 3 else int(math.log2(s))
                names = ["upsample.%d" % (2 * i) for i in range(n)]
                W["ups"], W["bups"], W["ups_r"] = bf(n, r * r * nf, 9 * nf), f32(n, r * r * nf), r
            else:
                names = ["conv_up1", "conv_up2", "conv_hr"]
                W["ups"], W["bups"] = bf(3, nf, 9 * nf), f32(3, nf)
            W["ups_names"] = names
            W["upsT"] = bf(W["ups"].shape[0], nf, 9 * W["ups"].shape[1]) if train else None
            W["pack"].append(o.build_pack_conv([P[n_ + ".weight"] for n_ in names], [P[n_ + ".bias"] for n_ in names],
                                               W["ups"], W["upsT"], W["bups"], O=W["ups"].shape[1], I=nf, I_pad=nf))
        W["versions"] = None
        self._weights = W
        self._plans = {}  # plans hold pointers into the packed operands
        return W

    def _param_versions(self):
        # + the global optimizer-step epoch: fused optimizers do not touch the version counters (kair_b200/_dirty.py)
        return (_dirty.optimizer_epoch(),) + tuple(p._version for p in self.P.values())

    def refresh_weights(self, W, stream):
        """Re-pack when any master parameter changed (optimizer step, load_state_dict)."""
        v = self._param_versions()
        if W["versions"] != v:
            for l in W["pack"]:
                l.run(stream)
            W["versions"] = v

    # ------------------------------------------------------------------------------------------ plans
    def plan(self, B, H, W, train, precision="bf16", terms=3):
        """precision 'bf16': tensor-core operands in bf16 (2e-2 max-abs contract); 'fp32': the fp32-accurate split
        mode of swinir_engine_fp32.py (1e-4 forward / 1e-3 gradient contract)."""
        if precision == "fp32":
            from . import swinir_engine_fp32 as f32m
            Wt = f32m.weight_state_fp32(self, train, terms)
            return pick_plan(self._plans_fp32, (B, H, W, bool(train), terms), 2,
                             lambda: f32m.PlanF32(self, Wt, B, H, W, bool(train), terms))
        if precision != "bf16":
            raise ValueError("kair_b200: precision must be 'bf16' or 'fp32', got %r" % (precision,))
        Wt = self._weight_state(train)
        return pick_plan(self._plans, (B, H, W, bool(train)), 4, lambda: _Plan(self, Wt, B, H, W, bool(train)))


class _Plan:
    def __init__(self, eng, Wt, B, H, W, train):
        s, o, dev = eng.spec, eng.ops, eng.device
        if H % s.ws or W % s.ws or min(H, W) <= s.ws:
            raise NotImplementedError("kair_b200 SwinIR: feature map %dx%d must be a multiple of and larger than the "
                                      "window (%d)" % (H, W, s.ws))
        self.eng, self.Wt, self.B, self.H, self.W, self.train = eng, Wt, B, H, W, train
        self.generation = 0
        T = B * H * W
        C, Cp, HP, QP, Fp, L, nl = eng.C, eng.Cp, eng.HP, eng.QP, eng.Fp, eng.L, eng.nl
        P = eng.P
        f32 = lambda *sh: torch.zeros(sh, dtype=torch.float32, device=dev)
        bf = lambda *sh: torch.zeros(sh, dtype=torch.bfloat16, device=dev)
        geo = (B, H, W)
        tps = H * W

        up = s.upsampler
        sc = s.upscale if up else 1
        self.img = f32(B, s.in_chans, H, W)
        self.out = f32(B, s.in_chans, H * sc, W * sc)
        first = f32(T, Cp)
        mean0, rstd0 = f32(T), f32(T)
        xs = [f32(T, Cp), f32(T, Cp)]
        tin = [f32(T, Cp), f32(T, Cp)]
        tfin = f32(T, Cp)
        meanF, rstdF = f32(T), f32(T)
        yb, resb = bf(T, Cp), bf(T, Cp)
        if train:
            blocks = [_Blk(T, eng, dev) for _ in range(L)]
            x2b = [bf(T, Cp) for _ in range(nl)]
        else:
            shared = _Blk(T, eng, dev)
            alt = _Blk.__new__(_Blk)  # the next block's xh1 / r1 must not alias the current one's
            alt.__dict__.update(shared.__dict__)
            alt.xh1, alt.r1 = bf(T, Cp), f32(T)
            blocks = [shared if l % 2 == 0 else alt for l in range(L)]
            x2b = [bf(T, Cp)] * nl
        # DropPath per-sample scales: rows [0, L) attention branch, [L, 2L) MLP branch
        self.dp = torch.ones(2 * L, B, dtype=torch.float32, device=dev)
        self.dp_active = False
        dpa = lambda l: self.dp[l] if train else None
        dpm = lambda l: self.dp[L + l] if train else None

        three = s.resi_connection == "3conv"
        Cq, Cqp = C // 4, _round_up(C // 4 + 1, 64)
        Kc = Cqp if three else Cp
        mids = {}   # '3conv': conv index -> (a1, a2), the two LeakyReLU outputs (kept per conv when training)
        mid_shared = (bf(T, Cqp), bf(T, Cqp)) if three and not train else None

        def conv_operand(i, src, lst):
            """bf16 input of the convolution that closes RSTB i (i == nl: conv_after_body); '3conv' runs its first two
            layers here (the ones column Cq of a1 / a2 carries the next layer's bias gradient)"""
            if not three:
                return src
            a1, a2 = mid_shared if mid_shared is not None else (bf(T, Cqp), bf(T, Cqp))
            mids[i] = (a1, a2)
            lst.append(o.build_gemm_tn(src, Wt["c3a"][i], N=Cqp, K=Cp, bn=_tile(Cqp), conv=geo, bias=Wt["bc3a"][i],
                                       act=o.ACT_LRELU, act_slope=0.2, outB=a1, ones_col=Cq))
            lst.append(o.build_gemm_tn(a1, Wt["c3b"][i], N=Cqp, K=Cqp, bn=_tile(Cqp), bias=Wt["bc3b"][i],
                                       act=o.ACT_LRELU, act_slope=0.2, outB=a2, ones_col=Cq))
            return a2

        cf_w, cf_b = P["conv_first.weight"], P["conv_first.bias"]
        cl_w, cl_b = (P["conv_last.weight"], P["conv_last.bias"]) if up != "pixelshuffledirect" else (None, None)
        tables = [P[b + "attn.relative_position_bias_table"] for b in eng._block_names()]
        mean_t = eng.mean
        rng = s.img_range

        fuse_mlp = FUSE_MLP and getattr(eng, "fuse_mlp", True) and Fp == 2 * Cp and Cp in (64, 192, 256)
        self.fuse_mlp = fuse_mlp

        # ================================================================================ forward
        fwd = []
        fwd.append(o.build_conv_in_fwd(self.img, cf_w, cf_b, mean_t, rng, O=C, O_pad=Cp, pitch=Cp, out_f32=first))
        if s.patch_norm:
            fwd.append(o.build_layernorm_fwd(first, P["patch_embed.norm.weight"], P["patch_embed.norm.bias"], C=C,
                                             y_f32=tin[0], mean=mean0, rstd=rstd0))
            t0 = tin[0]
        else:
            t0 = first
        fwd.append(o.build_layernorm_fwd(t0, None, None, C=C, xhat_bf16=blocks[0].xh1, rstd=blocks[0].r1, ones_col=C))
        l = 0
        t_cur = t0
        for i, depth in enumerate(s.depths):
            xin = t_cur
            for j in range(depth):
                S = blocks[l]
                shift = s.ws // 2 if j % 2 == 1 else 0
                fwd.append(o.build_gemm_tn(S.xh1, Wt["qkv"][l], N=QP, K=Cp, bn=_tile(QP), bias=Wt["bqkv"][l], outB=S.qkv))
                fwd.append(o.build_window_attn_fwd(S.qkv, tables[l], S.att, batch=B, H=H, W=W, num_heads=eng.nH,
                                                   window=s.ws, shift=shift))
                fwd.append(o.build_gemm_tn(S.att, Wt["proj"][l], N=Cp, K=HP, bn=_tile(Cp, True), mode=o.EPI_RESID_LN,
                                           bias=Wt["bproj"][l], resid=xin, outF=xs[0], outB=S.xh2, rowstat=S.r2,
                                           n_valid=C, ones_col=C, acc_scale=dpa(l), tokens_per_sample=tps))
                if fuse_mlp:
                    # the MLP half in one launch: u / g stay on chip (training plans keep one bf16 copy of each for
                    # the backward pass), the epilogue is the residual add + the next block's LayerNorm
                    nxt = blocks[l + 1] if j < depth - 1 else None
                    fwd.append(o.build_mlp_fused(S.xh2, Wt["fc1"][l], Wt["fc2"][l], N=Cp, F=Fp, bias1=Wt["bfc1"][l],
                                                 bias2=Wt["bfc2"][l], resid=xs[0],
                                                 outF=xs[1] if nxt is not None else None,
                                                 outB=nxt.xh1 if nxt is not None else x2b[i],
                                                 outU=S.u if train else None, outG=S.g if train else None,
                                                 rowstat=nxt.r1 if nxt is not None else None, acc_scale=dpm(l),
                                                 tokens_per_sample=tps, ln=nxt is not None, n_valid=C, ones_col=C,
                                                 ones_col_g=s.hidden))
                    if nxt is not None:
                        xin = xs[1]
                    l += 1
                    continue
                fwd.append(o.build_gemm_tn(S.xh2, Wt["fc1"][l], N=Fp, K=Cp, bn=_tile(Fp, True), mode=o.EPI_GELU,
                                           bias=Wt["bfc1"][l], outB=S.g, outB2=S.u, ones_col=s.hidden))
                if j < depth - 1:
                    nxt = blocks[l + 1]
                    fwd.append(o.build_gemm_tn(S.g, Wt["fc2"][l], N=Cp, K=Fp, bn=_tile(Cp, True), mode=o.EPI_RESID_LN,
                                               bias=Wt["bfc2"][l], resid=xs[0], outF=xs[1], outB=nxt.xh1,
                                               rowstat=nxt.r1, n_valid=C, ones_col=C, acc_scale=dpm(l),
                                               tokens_per_sample=tps))
                    xin = xs[1]
                else:
                    fwd.append(o.build_gemm_tn(S.g, Wt["fc2"][l], N=Cp, K=Fp, bn=_tile(Cp), mode=o.EPI_LINEAR,
                                               bias=Wt["bfc2"][l], resid=xs[0], outB=x2b[i], ones_col=C,
                                               acc_scale=dpm(l), tokens_per_sample=tps))
                l += 1
            if i < nl - 1:
                nxt = blocks[l]
                t_next = tin[1] if t_cur is tin[0] else tin[0]
                fwd.append(o.build_gemm_tn(conv_operand(i, x2b[i], fwd), Wt["conv"][i], N=Cp, K=Kc, bn=_tile(Cp, True), conv=geo, mode=o.EPI_RESID_LN,
                                           bias=Wt["bconv"][i], resid=t_cur, outF=t_next, outB=nxt.xh1,
                                           rowstat=nxt.r1, n_valid=C, ones_col=C))
                t_cur = t_next
            else:
                fwd.append(o.build_gemm_tn(conv_operand(i, x2b[i], fwd), Wt["conv"][i], N=Cp, K=Kc, bn=_tile(Cp), conv=geo, mode=o.EPI_LINEAR,
                                           bias=Wt["bconv"][i], resid=t_cur, outF=tfin))
        fwd.append(o.build_layernorm_fwd(tfin, P["norm.weight"], P["norm.bias"], C=C, y_bf16=yb, mean=meanF,
                                         rstd=rstdF, ones_col=C))
        fwd.append(o.build_gemm_tn(conv_operand(nl, yb, fwd), Wt["conv"][nl], N=Cp, K=Kc, bn=_tile(Cp), conv=geo, mode=o.EPI_LINEAR,
                                   bias=Wt["bconv"][nl], resid=first, outB=resb, ones_col=C))
        if not up:
            fwd.append(o.build_conv_out_fwd(resb, cl_w, cl_b, self.out, mean_t, C=C, pitch=Cp, out_scale=1.0 / rng,
                                            skip=self.img, skip_mean=mean_t, skip_scale=rng))
        elif up == "pixelshuffledirect":
            no = Wt["psd"].shape[1]
            u = bf(T, no)
            head = None
            fwd.append(o.build_gemm_tn(resb, Wt["psd"][0], N=no, K=Cp, bn=_tile(no), conv=geo, bias=Wt["bpsd"][0], outB=u))
            fwd.append(o.build_pixel_shuffle_out(u, self.out, mean_t, r=s.upscale, scale=1.0 / rng))
        else:
            # conv_before_upsample + LeakyReLU(0.01), then the upsampling convolutions on 64 features
            nf = 64
            f = f0 = bf(T, nf)
            fwd.append(o.build_gemm_tn(resb, Wt["cbu"][0], N=nf, K=Cp, bn=64, conv=geo, bias=Wt["bcbu"][0],
                                       act=o.ACT_LRELU, act_slope=0.01, outB=f))
            h, w = H, W
            head = []   # per upsampling conv: (conv input, its (B, h, w), activation output feeding the resampler)
            if up == "pixelshuffle":
                r = Wt["ups_r"]
                for i in range(Wt["ups"].shape[0]):
                    u = bf(B * h * w, r * r * nf)
                    head.append((f, (B, h, w), None))
                    fwd.append(o.build_gemm_tn(f, Wt["ups"][i], N=r * r * nf, K=nf, bn=(192 if r == 3 else 256),
                                               conv=(B, h, w), bias=Wt["bups"][i], outB=u))
                    f = bf(B * h * r * w * r, nf)
                    fwd.append(o.build_pixel_shuffle(u, f, batch=B, H=h, W=w, C=nf, r=r))
                    h, w = h * r, w * r
            else:
                for i in range(2):
                    n_ = bf(B * 4 * h * w, nf)
                    fwd.append(o.build_upsample_nearest2(f, n_, batch=B, H=h, W=w, C=nf))
                    h, w = 2 * h, 2 * w
                    head.append((n_, (B, h, w), f))
                    f = bf(B * h * w, nf)
                    fwd.append(o.build_gemm_tn(n_, Wt["ups"][i], N=nf, K=nf, bn=64, conv=(B, h, w), bias=Wt["bups"][i],
                                               act=o.ACT_LRELU, act_slope=0.2, outB=f))
                f2 = bf(B * h * w, nf)
                head.append((f, (B, h, w), None))
                fwd.append(o.build_gemm_tn(f, Wt["ups"][2], N=nf, K=nf, bn=64, conv=(B, h, w), bias=Wt["bups"][2],
                                           act=o.ACT_LRELU, act_slope=0.2, outB=f2))
                f = f2
            fwd.append(o.build_conv_out_fwd(f, cl_w, cl_b, self.out, mean_t, C=nf, pitch=nf, out_scale=1.0 / rng))
        self.fwd = fwd
        self.bwd = None
        if not train:
            return

        # ================================================================================ backward
        self.dout = f32(B, s.in_chans, H * sc, W * sc)
        # gradients in the parameters' own layouts: one flat fp32 buffer, views per parameter
        sizes = [p.numel() for p in P.values()]
        self.gflat = f32(sum(sizes))
        G, off = {}, 0
        for (k, p), n in zip(P.items(), sizes):
            G[k] = self.gflat[off:off + n].view(p.shape)
            off += n
        self.G = G
        # packed weight gradients (fp32, accumulated by the token-contraction GEMMs)
        n_qkv, n_proj, n_fc = L * QP * Cp, L * Cp * HP, L * Fp * Cp
        n_conv = (nl + 1) * Cp * 9 * Kc
        self.dwflat = f32(n_qkv + n_proj + 2 * n_fc + n_conv)
        o0 = 0
        dWqkv = self.dwflat[o0:o0 + n_qkv].view(L, QP, Cp); o0 += n_qkv
        dWproj = self.dwflat[o0:o0 + n_proj].view(L, Cp, HP); o0 += n_proj
        dWfc1 = self.dwflat[o0:o0 + n_fc].view(L, Fp, Cp); o0 += n_fc
        dWfc2 = self.dwflat[o0:o0 + n_fc].view(L, Cp, Fp); o0 += n_fc
        dWconv = self.dwflat[o0:o0 + n_conv].view(nl + 1, Cp, 9 * Kc)
        if three:
            dWa, dWb = f32(nl + 1, Cqp, 9 * Cp), f32(nl + 1, Cqp, Cqp)
            da1, da2 = bf(T, Cqp), bf(T, Cqp)

        def conv_backward(i, g, src, **out):
            """weight gradients of closing convolution i and its input gradient (epilogue outputs given by **out)"""
            if not three:
                bwd.append(o.build_gemm_nt(g, src, dWconv[i], n1=Cp, n2=Cp, bn=_tile(Cp), conv=geo))
                bwd.append(o.build_gemm_tn(g, Wt["convT"][i], N=Cp, K=Cp, bn=_tile(Cp), conv=geo, **out))
                return
            a1, a2 = mids[i]
            bwd.append(o.build_gemm_nt(g, a2, dWconv[i], n1=Cp, n2=Cqp, bn=_tile(Cqp), conv=geo))
            bwd.append(o.build_gemm_tn(g, Wt["convT"][i], N=Cqp, K=Cp, bn=_tile(Cqp), conv=geo, outB=da2, aux=a2,
                                       act_slope=0.2))
            bwd.append(o.build_gemm_nt(da2, a1, dWb[i], n1=Cqp, n2=Cqp, bn=_tile(Cqp)))
            bwd.append(o.build_gemm_tn(da2, Wt["c3bT"][i], N=Cqp, K=Cqp, bn=_tile(Cqp), outB=da1, aux=a1, act_slope=0.2))
            bwd.append(o.build_gemm_nt(da1, src, dWa[i], n1=Cqp, n2=Cp, bn=_tile(Cp), conv=geo))
            bwd.append(o.build_gemm_tn(da1, Wt["c3aT"][i], N=Cp, K=Cqp, bn=_tile(Cp), conv=geo, **out))
        dresf, dresb, dyb = f32(T, Cp), bf(T, Cp), bf(T, Cp)
        R = [(f32(T, Cp), bf(T, Cp)) for _ in range(3)]  # rotating (fp32, bf16) gradient buffers
        du, datt, dqkv = bf(T, Fp), bf(T, HP), bf(T, QP)
        blk_names = eng._block_names()

        cn = Wt["conv_names"]
        keys = list(P.keys())
        offs, acc_ = {}, 0
        for k_, n_ in zip(keys, sizes):
            offs[k_] = (acc_, acc_ + n_)
            acc_ += n_

        def rstb_range(i):
            """[lo, hi) of RSTB i's parameters in the flat gradient buffer (contiguous in state_dict order)"""
            ks = [k_ for k_ in keys if k_.startswith("layers.%d." % i)]
            lo, hi = offs[ks[0]][0], offs[ks[-1]][1]
            assert hi - lo == sum(offs[k_][1] - offs[k_][0] for k_ in ks)
            return lo, hi

        def unpack_convs(c0, c1):
            names_ = cn[c0:c1]
            if not three:
                return [o.build_unpack_conv_grad(dWconv[c0:c1], [G[n + ".weight"] for n in names_],
                                                 [G[n + ".bias"] for n in names_], O=C, I=C, I_pad=Cp, ones_col=C,
                                                 center_tap=4)]
            gw = lambda sfx: [G[n + sfx + ".weight"] for n in names_]
            gb = lambda sfx: [G[n + sfx + ".bias"] for n in names_]
            return [o.build_unpack_conv_grad(dWa[c0:c1], gw(".0"), gb(".0"), O=Cq, I=C, I_pad=Cp, ones_col=C, center_tap=4),
                    o.build_unpack_conv_grad(dWb[c0:c1], gw(".2"), gb(".2"), O=Cq, I=Cq, I_pad=Cqp, taps=1, ones_col=Cq,
                                             center_tap=0),
                    o.build_unpack_conv_grad(dWconv[c0:c1], gw(".4"), gb(".4"), O=C, I=Cq, I_pad=Cqp, ones_col=Cq,
                                             center_tap=4)]

        def unpack_rstb(i, l0, l1):
            """packed weight gradients of blocks [l0, l1) and of the convolution closing RSTB i -> dW, db, dgamma, dbeta"""
            names_ = blk_names[l0:l1]
            gl = lambda suffix: [G[bn_ + suffix] for bn_ in names_]
            pl = lambda suffix: [P[bn_ + suffix] for bn_ in names_]
            return [
                o.build_unpack_linear_grad(dWqkv[l0:l1], pl("attn.qkv.weight"), pl("norm1.weight"), pl("norm1.bias"),
                                           gl("attn.qkv.weight"), gl("attn.qkv.bias") if s.qkv_bias else None,
                                           gl("norm1.weight"), gl("norm1.bias"), eng.map_qkv, eng.map_c, eng.scale_qkv,
                                           out_f=3 * C, in_f=C, ones_col=C),
                o.build_unpack_linear_grad(dWproj[l0:l1], pl("attn.proj.weight"), None, None, gl("attn.proj.weight"),
                                           gl("attn.proj.bias"), None, None, eng.map_c, eng.map_h, None, out_f=C, in_f=C,
                                           ones_col=eng.hd),
                o.build_unpack_linear_grad(dWfc1[l0:l1], pl("mlp.fc1.weight"), pl("norm2.weight"), pl("norm2.bias"),
                                           gl("mlp.fc1.weight"), gl("mlp.fc1.bias"), gl("norm2.weight"), gl("norm2.bias"),
                                           eng.map_f, eng.map_c, None, out_f=s.hidden, in_f=C, ones_col=C),
                o.build_unpack_linear_grad(dWfc2[l0:l1], pl("mlp.fc2.weight"), None, None, gl("mlp.fc2.weight"),
                                           gl("mlp.fc2.bias"), None, None, eng.map_c, eng.map_f, None, out_f=C,
                                           in_f=s.hidden, ones_col=s.hidden)] + unpack_convs(i, i + 1)

        marks = []   # (number of backward launches after which the slice is final, lo, hi) per RSTB
        bwd = []
        inv = 1.0 / rng
        ci = s.in_chans
        unpack_head = []
        colsum = lambda x, out, n: [o.build_colsum(x[:, c0:min(c0 + 512, n)], out[c0:min(c0 + 512, n)],
                                                   C=min(c0 + 512, n) - c0) for c0 in range(0, n, 512)]
        if not up:
            # conv_last backward as GEMMs over the 9-tap expansion of d(out) (columns tap*7 + c, K padded to 64)
            dsh, clT, dWcl = bf(T, 64), bf(Cp, 64), f32(Cp, 64)
            self._thin = (dWcl,)
            bwd.append(o.build_im2col3x3(self.dout, dsh, cpad=7, direction=-1, scale=inv))
            bwd.append(o.build_conv_out_pack(mode=0, co=ci, C=C, cpad=7, W=cl_w, Wt=clT))
            bwd.append(o.build_gemm_tn(dsh, clT, N=Cp, K=64, bn=_tile(Cp), outF=dresf, outB=dresb))
            bwd.append(o.build_gemm_nt(resb, dsh, dWcl, n1=Cp, n2=64, bn=64))
            bwd.append(o.build_conv_out_pack(mode=1, co=ci, C=C, cpad=7, D=dWcl, dW=G["conv_last.weight"],
                                             db=G["conv_last.bias"], ones_row=C))
        elif up == "pixelshuffledirect":
            # d(out) un-shuffled into the conv's output channels; bias gradient from the ones column of `res`
            nop = Wt["psd"].shape[1]
            dps, dWpsd = bf(T, nop), f32(1, nop, 9 * Cp)
            self._thin = (dWpsd,)
            bwd.append(o.build_pixel_shuffle_out_bwd(self.dout, dps, r=s.upscale, co=ci, scale=inv))
            bwd.append(o.build_gemm_nt(dps, resb, dWpsd[0], n1=nop, n2=Cp, bn=_tile(Cp), conv=geo))
            bwd.append(o.build_gemm_tn(dps, Wt["psdT"][0], N=Cp, K=nop, bn=_tile(Cp), conv=geo, outF=dresf, outB=dresb))
            unpack_head.append(o.build_unpack_conv_grad(dWpsd, [G["upsample.0.weight"]], [G["upsample.0.bias"]],
                                                        O=s.upscale ** 2 * ci, I=C, I_pad=Cp, ones_col=C, center_tap=4))
        else:
            # conv_last on 64 features at the output resolution (thin: 9-tap expansion of d(out)), then the upsampling
            # convolutions in reverse; LeakyReLU masks ride on the dgrad epilogues / the nearest-x2 adjoint
            nf = 64
            names = Wt["ups_names"]
            Th = B * H * sc * W * sc
            dsh, clT, dWcl, csum = bf(Th, 64), bf(nf, 64), f32(nf, 64), f32(64)
            n_up, O_up = Wt["ups"].shape[0], Wt["ups"].shape[1]
            dWups, dWcbu = f32(n_up, O_up, 9 * nf), f32(1, nf, 9 * Cp)
            self._thin = (dWcl, csum, dWups, dWcbu)
            bwd.append(o.build_im2col3x3(self.dout, dsh, cpad=7, direction=-1, scale=inv))
            bwd.append(o.build_conv_out_pack(mode=0, co=ci, C=nf, cpad=7, W=cl_w, Wt=clT))
            bwd.append(o.build_gemm_nt(f, dsh, dWcl, n1=nf, n2=64, bn=64))
            bwd.append(o.build_conv_out_pack(mode=1, co=ci, C=nf, cpad=7, D=dWcl, dW=G["conv_last.weight"]))
            bwd.append(o.build_colsum(dsh, csum, C=64))
            bwd.append(o.build_copy(G["conv_last.bias"], csum[4 * 7:4 * 7 + ci]))
            g = bf(Th, nf)
            lrelu_out = up == "nearest+conv"   # conv_last reads lrelu(conv_hr(.)) there, a bare PixelShuffle otherwise
            bwd.append(o.build_gemm_tn(dsh, clT, N=nf, K=64, bn=64, outB=g, aux=f if lrelu_out else None,
                                       act_slope=0.2 if lrelu_out else 0.0))
            for i in reversed(range(n_up)):
                xin, (_, h_, w_), pre = head[i]
                if up == "pixelshuffle":
                    r = Wt["ups_r"]
                    du_ = bf(B * h_ * w_, O_up)
                    bwd.append(o.build_pixel_shuffle_bwd(g, du_, batch=B, H=h_, W=w_, C=nf, r=r))
                    g = du_
                bwd += colsum(g, G[names[i] + ".bias"], O_up)
                bwd.append(o.build_gemm_nt(g, xin, dWups[i], n1=O_up, n2=nf, bn=64, conv=(B, h_, w_)))
                gin = bf(B * h_ * w_, nf)
                if up == "pixelshuffle":
                    # the first conv reads f0 = LeakyReLU(0.01)(conv_before_upsample), later ones a bare PixelShuffle
                    bwd.append(o.build_gemm_tn(g, Wt["upsT"][i], N=nf, K=O_up, bn=64, conv=(B, h_, w_), outB=gin,
                                               aux=f0 if i == 0 else None, act_slope=0.01 if i == 0 else 0.0))
                    g = gin
                elif pre is None:
                    # conv_hr reads lrelu(conv_up2(.)) directly
                    bwd.append(o.build_gemm_tn(g, Wt["upsT"][i], N=nf, K=O_up, bn=64, conv=(B, h_, w_), outB=gin,
                                               aux=xin, act_slope=0.2))
                    g = gin
                else:
                    bwd.append(o.build_gemm_tn(g, Wt["upsT"][i], N=nf, K=O_up, bn=64, conv=(B, h_, w_), outB=gin))
                    g = bf(B * (h_ // 2) * (w_ // 2), nf)
                    bwd.append(o.build_upsample_nearest2_bwd(gin, g, batch=B, H=h_ // 2, W=w_ // 2, C=nf, mask=pre,
                                                             mask_slope=0.01 if i == 0 else 0.2))
            # g = d(f0), masked; conv_before_upsample
            bwd.append(o.build_gemm_nt(g, resb, dWcbu[0], n1=nf, n2=Cp, bn=_tile(Cp), conv=geo))
            bwd.append(o.build_gemm_tn(g, Wt["cbuT"][0], N=Cp, K=nf, bn=_tile(Cp), conv=geo, outF=dresf, outB=dresb))
            unpack_head.append(o.build_unpack_conv_grad(dWcbu, [G["conv_before_upsample.0.weight"]],
                                                        [G["conv_before_upsample.0.bias"]], O=nf, I=C, I_pad=Cp,
                                                        ones_col=C, center_tap=4))
            unpack_head.append(o.build_unpack_conv_grad(dWups, [G[n_ + ".weight"] for n_ in names], None, O=O_up, I=nf,
                                                        I_pad=nf))
        if three:
            self._thin += (dWa, dWb)
        conv_backward(nl, dresb, yb, outB=dyb)
        dt = 2
        bwd.append(o.build_layernorm_bwd(tfin, meanF, rstdF, P["norm.weight"], G["norm.weight"], G["norm.bias"], C=C,
                                         dy_bf16=dyb, dx_f32=R[dt][0], dx_bf16=R[dt][1]))
        l_end = L
        for i in reversed(range(nl)):
            depth = s.depths[i]
            a, b = [k for k in range(3) if k != dt]
            conv_backward(i, R[dt][1], x2b[i], outF=R[a][0], outB=R[a][1], outB_scale=dpm(l_end - 1),
                          tokens_per_sample=tps)
            cur, oth = a, b
            for j in reversed(range(depth)):
                l = l_end - depth + j
                S = blocks[l]
                bwd.append(o.build_gemm_nt(R[cur][1], S.g, dWfc2[l], n1=Cp, n2=Fp, bn=_tile(Fp)))
                bwd.append(o.build_gemm_tn(R[cur][1], Wt["fc2T"][l], N=Fp, K=Cp, bn=_tile(Fp, True), mode=o.EPI_DGELU, aux=S.u,
                                           outB=du))
                bwd.append(o.build_gemm_nt(du, S.xh2, dWfc1[l], n1=Fp, n2=Cp, bn=_tile(Cp)))
                bwd.append(o.build_gemm_tn(du, Wt["fc1T"][l], N=Cp, K=Fp, bn=_tile(Cp, True), mode=o.EPI_LNBWD, aux=S.xh2,
                                           rowstat=S.r2, resid=R[cur][0], outF=R[oth][0], outB=R[oth][1], n_valid=C,
                                           outB_scale=dpa(l), tokens_per_sample=tps))
                cur, oth = oth, cur
                bwd.append(o.build_gemm_nt(R[cur][1], S.att, dWproj[l], n1=Cp, n2=HP, bn=_tile(HP)))
                bwd.append(o.build_gemm_tn(R[cur][1], Wt["projT"][l], N=HP, K=Cp, bn=_tile(HP), outB=datt))
                shift = s.ws // 2 if j % 2 == 1 else 0
                bwd.append(o.build_window_attn_bwd(S.qkv, tables[l], datt, dqkv,
                                                   G[blk_names[l] + "attn.relative_position_bias_table"], batch=B,
                                                   H=H, W=W, num_heads=eng.nH, window=s.ws, shift=shift))
                bwd.append(o.build_gemm_nt(dqkv, S.xh1, dWqkv[l], n1=QP, n2=Cp, bn=_tile(Cp)))
                bwd.append(o.build_gemm_tn(dqkv, Wt["qkvT"][l], N=Cp, K=QP, bn=_tile(Cp, True), mode=o.EPI_LNBWD, aux=S.xh1,
                                           rowstat=S.r1, resid=R[cur][0], outF=R[oth][0], outB=R[oth][1], n_valid=C,
                                           outB_scale=(dpm(l - 1) if j > 0 else None), tokens_per_sample=tps))
                cur, oth = oth, cur
            # RSTB skip: d(t_in) = d(block path) + d(t_out)
            bwd.append(o.build_add_inplace(R[cur][0], R[dt][0], R[cur][1]))
            dt = cur
            l_end -= depth
            # every gradient of RSTB i is now complete in its packed form: map it onto the parameters' own layouts and
            # mark the (contiguous) slice of the flat gradient buffer as ready, so that a data-parallel hook can start
            # its all-reduce while the earlier RSTBs are still in their backward pass (model_base.py:113-119: DDP
            # overlaps per bucket)
            bwd += unpack_rstb(i, l_end, l_end + depth)
            marks.append((len(bwd), ) + rstb_range(i))
        if s.patch_norm:
            dfirst_b, dfirst_f = R[(dt + 1) % 3][1], R[(dt + 1) % 3][0]
            bwd.append(o.build_layernorm_bwd(first, mean0, rstd0, P["patch_embed.norm.weight"],
                                             G["patch_embed.norm.weight"], G["patch_embed.norm.bias"], C=C,
                                             dy_f32=R[dt][0], dx_add=dresf, dx_f32=dfirst_f, dx_bf16=dfirst_b))
        else:
            bwd.append(o.build_add_inplace(R[dt][0], dresf, R[dt][1]))
            dfirst_b, dfirst_f = R[dt][1], R[dt][0]
        # conv_first weight gradient = d(first)^T . im2col(input) on the tensor cores (columns tap*7 + c, ones -> bias)
        xcol, dWcf = bf(T, 64), f32(1, Cp, 64)
        self._thin += (dWcf,)
        bwd.append(o.build_im2col3x3(self.img, xcol, cpad=7, direction=1, mean=mean_t, scale=rng, ones_col=4 * 7 + ci))
        bwd.append(o.build_gemm_nt(dfirst_b, xcol, dWcf[0], n1=Cp, n2=64, bn=64))
        bwd.append(o.build_unpack_conv_grad(dWcf, [G["conv_first.weight"]], [G["conv_first.bias"]], O=C, I=ci, I_pad=7,
                                            ones_col=ci, center_tap=4, dwp_pitch=64))
        bwd += unpack_convs(nl, nl + 1)     # conv_after_body
        self.bwd = bwd + unpack_head
        self.bwd_marks = marks
        self._dx_spec = dict(dhead=dfirst_f, W=cf_w, C=C, pitch=Cp, scale=rng, skip=not up)

    # ------------------------------------------------------------------------------------------ execution
    def set_drop_path(self, rates, generator=None):
        """Draw the per-sample Bernoulli(keep)/keep DropPath scales of this step (network_swinir.py:204, 276-277).
        rates: per-block drop probability (len L); all-zero rates leave the scales at 1."""
        L = self.eng.L
        if not any(r > 0 for r in rates):
            if self.dp_active:
                self.dp.fill_(1.0)
                self.dp_active = False
            return
        keep = torch.tensor([1.0 - r for r in rates] * 2, dtype=torch.float32, device=self.dp.device).unsqueeze(1)
        u = torch.rand(2 * L, self.B, device=self.dp.device, generator=generator)
        self.dp.copy_((u < keep).float() / keep)
        self.dp_active = True

    def _lists(self):
        """the launch lists as C arrays (one FFI call per pass / per segment), built on first use"""
        ls = getattr(self, "_launch_lists", None)
        if ls is None:
            o = self.eng.ops
            segs = []
            if self.bwd is not None:
                start = 0
                for end, lo, hi in getattr(self, "bwd_marks", []):
                    segs.append((o.make_list(self.bwd[start:end]), (lo, hi)))
                    start = end
                segs.append((o.make_list(self.bwd[start:]), None))
            ls = self._launch_lists = (o.make_list(self.fwd), segs)
        return ls

    def run_forward(self, x):
        o = self.eng.ops
        with o.device_guard(self.img.device):   # the launchers use the current device and its current stream
            st = o.stream_handle(self.img.device)
            self.eng.refresh_weights(self.Wt, st)
            self.img.copy_(x)
            self._lists()[0].run(st)
        self.generation += 1
        return self.out.clone()

    def run_backward(self, dout):
        o = self.eng.ops
        with o.device_guard(self.img.device):
            st = o.stream_handle(self.img.device)
            self.dout.copy_(dout)
            self.gflat.zero_()
            self.dwflat.zero_()
            for t in self._thin:
                t.zero_()
            slice_hook = getattr(self.eng, "grad_slice_hook", None)
            for lst, ready in self._lists()[1]:
                lst.run(st)
                if ready is not None and slice_hook is not None:
                    slice_hook(self.gflat, ready[0], ready[1])
            if self.eng.grad_hook is not None:
                self.eng.grad_hook(self.gflat)
        flat = self.gflat.clone()
        grads, off = [], 0
        for p in self.eng.P.values():
            n = p.numel()
            grads.append(flat[off:off + n].view(p.shape))
            off += n
        return grads


def input_grad(plan, dout):
    """d(loss)/d(input image) of a plan whose backward list has just run: the transposed first convolution applied to
    the gradient of its output (kair_conv_out_fwd with the spatially flipped, channel-swapped conv_first weights --
    autograd of network_swinir.py:809-813 / network_dncnn.py:63 / network_unet.py:25), times img_range, plus the global
    skip `x + ...` (:835) where the network has one.  Built on first use: training loops do not need it."""
    sp = plan._dx_spec
    o, W = plan.eng.ops, sp["W"]
    st = getattr(plan, "_dx_state", None)
    if st is None:
        dev = plan.img.device
        ci = W.shape[1]
        Wt = torch.zeros(ci, W.shape[0], 3, 3, dtype=torch.float32, device=dev)
        zc = torch.zeros(ci, dtype=torch.float32, device=dev)
        dimg = torch.zeros_like(plan.img)
        skip = plan.dout if sp["skip"] else None
        launch = o.build_conv_out_fwd(sp["dhead"], Wt, None, dimg, zc, C=sp["C"], pitch=sp["pitch"],
                                      out_scale=sp["scale"], skip=skip, skip_mean=zc if skip is not None else None,
                                      skip_scale=1.0 / sp["scale"])
        st = plan._dx_state = (Wt, dimg, launch)
    Wt, dimg, launch = st
    Wt.copy_(W.detach().flip(2, 3).transpose(0, 1))
    launch.run(o.stream_handle())
    return dimg.clone()


class SwinIRFunction(torch.autograd.Function):
    """forward/backward pair over the whole network (the reference's plugin convention, models/op/fused_act.py:21-72)."""

    @staticmethod
    def forward(ctx, x, plan, *params):
        out = plan.run_forward(x)
        mark_busy(plan)
        ctx.plan = plan
        ctx.generation = plan.generation
        return out

    @staticmethod
    def backward(ctx, dout):
        plan = ctx.plan
        if plan.generation != ctx.generation:
            raise RuntimeError("kair_b200 SwinIR: backward() called after two later forward() calls of the same shape "
                               "overwrote the saved activations (two sets are kept per shape)")
        plan.busy = False
        grads = plan.run_backward(dout.contiguous())
        dx = input_grad(plan, dout) if ctx.needs_input_grad[0] else None
        return (dx, None) + tuple(grads)
