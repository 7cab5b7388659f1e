"""fp32-accurate plans for SwinIR ("SPLIT PRECISION", include/kair_b200.h): north_star's second precision mode, forward
within 1e-4 max-abs and gradients within 1e-3 relative of the reference's fp32 arithmetic
(models/network_swinir.py:114-145, 239-279, 805-839 of the reference).

Same engine, same packing maps, same tcgen05 GEMM kernels as the bf16 plans of swinir_engine.py; what changes:
  * every tensor-core operand is a *split* bf16 operand: an fp32 activation x travels as bf16 components
    [c0 | c1 | c0] (terms = 3) next to weights packed as [C0 | C0 | C1], so ONE kair_gemm_tn launch over a 3x longer K
    accumulates c0*C0 + c1*C0 + c0*C1 in fp32 (TMEM): ~2^-17 relative error per product instead of 2^-9.  terms = 6 adds
    the third component of both sides (~2^-24).  The weight-gradient GEMMs run once per component pair on column slices;
  * GEMM results leave the kernel as fp32 (outF) only; the element-wise stages that the bf16 plans fuse into GEMM
    epilogues (LayerNorm, exact GELU, GELU', LayerNorm backward, DropPath) run as fp32 kernels that write the next split
    operand (kair_layernorm_fwd / kair_split_terms / kair_ln_bwd_split);
  * window attention runs in fp32 on CUDA cores for every window size (kair_window_attn_* with fp32 = 1): ~7 % of the
    FLOPs; the thin first / last convolutions use the fp32 CUDA-core kernels.
Per Swin block: 9 forward launches (5 in bf16 mode) and 27 backward launches (9).  Expect ~1/4 of the bf16 rate.
"""
import math

import torch

from .swinir_engine import _Plan, _round_up, _tile


def _pairs(terms):
    """component pairs (i, j), i + j <= order, of a split product A_i * B_j; component c lives in segment c"""
    order = 1 if terms == 3 else 2
    return [(i, j) for j in range(order + 1) for i in range(order + 1 - j)]


class _BlkF32:
    """Activations of one Swin block kept for the backward pass (fp32-accurate mode)."""

    def __init__(self, T, g, dev, t):
        bf = lambda n: torch.zeros(T, n, dtype=torch.bfloat16, device=dev)
        f32 = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
        self.xh1 = bf(t * g.Cp)
        self.r1 = f32(T)
        self.qkv = f32(T, g.QP)
        self.att = bf(t * g.HP)
        self.xh2 = bf(t * g.Cp)
        self.r2 = f32(T)
        self.u = f32(T, g.Fp)
        self.g = bf(t * g.Fp)


def weight_state_fp32(eng, train, terms):
    """Packed split operands (weight side) + the launches that refresh them from the fp32 master parameters."""
    Wst = eng._weights_fp32
    if Wst is not None and Wst["terms"] == terms and (Wst["train"] or not train):
        return Wst
    o, P, dev, t = eng.ops, eng.P, eng.device, terms
    L, nl, Cp, HP, QP, Fp, C = eng.L, eng.nl, eng.Cp, eng.HP, eng.QP, eng.Fp, eng.C
    s = eng.spec
    if s.resi_connection != "1conv":
        raise NotImplementedError("kair_b200 SwinIR: precision='fp32' covers resi_connection='1conv' (got '%s')"
                                  % s.resi_connection)
    bf = lambda *sh: torch.zeros(sh, dtype=torch.bfloat16, device=dev)
    f32 = lambda *sh: torch.zeros(sh, dtype=torch.float32, device=dev)
    blk = eng._block_names()
    W = {"train": train, "terms": t}
    W["qkv"], W["qkvT"], W["bqkv"] = bf(L, QP, t * Cp), (bf(L, Cp, t * QP) if train else None), f32(L, QP)
    W["proj"], W["projT"], W["bproj"] = bf(L, Cp, t * HP), (bf(L, HP, t * Cp) if train else None), f32(L, Cp)
    W["fc1"], W["fc1T"], W["bfc1"] = bf(L, Fp, t * Cp), (bf(L, Cp, t * Fp) if train else None), f32(L, Fp)
    W["fc2"], W["fc2T"], W["bfc2"] = bf(L, Cp, t * Fp), (bf(L, Fp, t * Cp) if train else None), f32(L, Cp)
    W["conv"], W["convT"], W["bconv"] = bf(nl + 1, Cp, 9 * t * Cp), (bf(nl + 1, Cp, 9 * t * Cp) if train else None), \
        f32(nl + 1, Cp)
    get = lambda suffix: [P[b + suffix] for b in blk]
    qb = get("attn.qkv.bias") if s.qkv_bias else None
    conv_names = ["layers.%d.conv" % i for i in range(nl)] + ["conv_after_body"]
    W["conv_names"] = conv_names
    W["pack"] = [
        o.build_pack_linear(get("attn.qkv.weight"), qb, get("norm1.weight"), get("norm1.bias"), eng.map_qkv, eng.map_c,
                            eng.scale_qkv, eng.const_qkv, W["qkv"], W["qkvT"], W["bqkv"], out_f=3 * C, in_f=C, terms=t),
        o.build_pack_linear(get("attn.proj.weight"), get("attn.proj.bias"), None, None, eng.map_c, eng.map_h, None, None,
                            W["proj"], W["projT"], W["bproj"], out_f=C, in_f=C, terms=t),
        o.build_pack_linear(get("mlp.fc1.weight"), get("mlp.fc1.bias"), get("norm2.weight"), get("norm2.bias"),
                            eng.map_f, eng.map_c, None, None, W["fc1"], W["fc1T"], W["bfc1"], out_f=s.hidden, in_f=C,
                            terms=t),
        o.build_pack_linear(get("mlp.fc2.weight"), get("mlp.fc2.bias"), None, None, eng.map_c, eng.map_f, None, None,
                            W["fc2"], W["fc2T"], W["bfc2"], out_f=C, in_f=s.hidden, terms=t),
        o.build_pack_conv([P[n + ".weight"] for n in conv_names], [P[n + ".bias"] for n in conv_names], W["conv"],
                          W["convT"], W["bconv"], O=C, I=C, I_pad=Cp, terms=t),
    ]
    up = s.upsampler
    if up == "pixelshuffledirect":
        no = s.upscale ** 2 * s.in_chans
        nop = _round_up(no, 64)
        W["psd"], W["bpsd"] = bf(1, nop, 9 * t * Cp), f32(1, nop)
        W["pack"].append(o.build_pack_conv([P["upsample.0.weight"]], [P["upsample.0.bias"]], W["psd"], None, W["bpsd"],
                                           O=no, I=C, I_pad=Cp, terms=t))
    elif up:
        nf = 64
        W["cbu"], W["bcbu"] = bf(1, nf, 9 * t * Cp), f32(1, nf)
        W["pack"].append(o.build_pack_conv([P["conv_before_upsample.0.weight"]], [P["conv_before_upsample.0.bias"]],
                                           W["cbu"], None, W["bcbu"], O=nf, I=C, I_pad=Cp, terms=t))
        if up == "pixelshuffle":
            sc = s.upscale
            r = 3 if sc == 3 else 2
            n = 1 if sc == 3 else int(math.log2(sc))
            names = ["upsample.%d" % (2 * i) for i in range(n)]
            W["ups"], W["bups"], W["ups_r"] = bf(n, r * r * nf, 9 * t * nf), f32(n, r * r * nf), r
        else:
            names = ["conv_up1", "conv_up2", "conv_hr"]
            W["ups"], W["bups"] = bf(3, nf, 9 * t * nf), f32(3, nf)
        W["pack"].append(o.build_pack_conv([P[n_ + ".weight"] for n_ in names], [P[n_ + ".bias"] for n_ in names],
                                           W["ups"], None, W["bups"], O=W["ups"].shape[1], I=nf, I_pad=nf, terms=t))
    W["versions"] = None
    eng._weights_fp32 = W
    eng._plans_fp32 = {}
    return W


class PlanF32(_Plan):
    """Launch lists of the fp32-accurate mode; executed by _Plan.run_forward / run_backward."""

    def __init__(self, eng, Wt, B, H, W, train, terms):
        s, o, dev = eng.spec, eng.ops, eng.device
        if H % s.ws or W % s.ws or min(H, W) <= s.ws:
            raise NotImplementedError("kair_b200 SwinIR: feature map %dx%d must be a multiple of and larger than the "
                                      "window (%d)" % (H, W, s.ws))
        self.eng, self.Wt, self.B, self.H, self.W, self.train = eng, Wt, B, H, W, train
        self.generation = 0
        t = terms
        T = B * H * W
        C, Cp, HP, QP, Fp, L, nl = eng.C, eng.Cp, eng.HP, eng.QP, eng.Fp, eng.L, eng.nl
        P = eng.P
        f32 = lambda *sh: torch.zeros(sh, dtype=torch.float32, device=dev)
        bf = lambda *sh: torch.zeros(sh, dtype=torch.bfloat16, device=dev)
        geo = (B, H, W)
        tps = H * W
        up = s.upsampler
        sc = s.upscale if up else 1
        if train and up:
            raise NotImplementedError("kair_b200 SwinIR: precision='fp32' trains the denoising / restoration models "
                                      "(upsampler=''); the SR heads run forward-only in this mode")
        self.img = f32(B, s.in_chans, H, W)
        self.out = f32(B, s.in_chans, H * sc, W * sc)
        first = f32(T, Cp)
        mean0, rstd0 = f32(T), f32(T)
        xs = [f32(T, Cp), f32(T, Cp)]
        tin = [f32(T, Cp), f32(T, Cp)]
        meanF, rstdF = f32(T), f32(T)
        att_f = f32(T, HP)
        yb_s, res_f = bf(T, t * Cp), f32(T, Cp)
        if train:
            blocks = [_BlkF32(T, eng, dev, t) for _ in range(L)]
            x2s = [bf(T, t * Cp) for _ in range(nl)]
        else:
            shared = _BlkF32(T, eng, dev, t)
            alt = _BlkF32.__new__(_BlkF32)
            alt.__dict__.update(shared.__dict__)
            alt.xh1, alt.r1 = bf(T, t * Cp), f32(T)
            blocks = [shared if l % 2 == 0 else alt for l in range(L)]
            x2s = [bf(T, t * Cp)] * nl
        self.dp = torch.ones(2 * L, B, dtype=torch.float32, device=dev)
        self.dp_active = False
        dpa = lambda l: self.dp[l] if train else None
        dpm = lambda l: self.dp[L + l] if train else None
        cf_w, cf_b = P["conv_first.weight"], P["conv_first.bias"]
        cl_w, cl_b = (P["conv_last.weight"], P["conv_last.bias"]) if up != "pixelshuffledirect" else (None, None)
        tables = [P[b + "attn.relative_position_bias_table"] for b in eng._block_names()]
        mean_t, rng = eng.mean, s.img_range
        split = lambda x, out, n, **kw: o.build_split_terms(x, out, n=n, terms=t, **kw)

        def lin(A, Wm, N, K, **kw):
            launch = o.build_gemm_tn(A, Wm, N=N, K=t * K, bn=_tile(N), **kw)
            try:
                launch.meta = {"terms": t}    # accounting hint for kair_b200.profiling (algorithmic K = K / terms)
            except AttributeError:
                pass
            return launch

        # ================================================================================ forward
        fwd = []
        fwd.append(o.build_conv_in_fwd(self.img, cf_w, cf_b, mean_t, rng, O=C, O_pad=Cp, pitch=Cp, out_f32=first))
        if s.patch_norm:
            fwd.append(o.build_layernorm_fwd(first, P["patch_embed.norm.weight"], P["patch_embed.norm.bias"], C=C,
                                             y_f32=tin[0], mean=mean0, rstd=rstd0))
            t0 = tin[0]
        else:
            t0 = first
        fwd.append(o.build_layernorm_fwd(t0, None, None, C=C, xhat_bf16=blocks[0].xh1, rstd=blocks[0].r1, ones_col=C,
                                         terms=t))
        l = 0
        t_cur = t0
        tfin = None
        for i, depth in enumerate(s.depths):
            xin = t_cur
            for j in range(depth):
                S = blocks[l]
                shift = s.ws // 2 if j % 2 == 1 else 0
                fwd.append(lin(S.xh1, Wt["qkv"][l], QP, Cp, bias=Wt["bqkv"][l], outF=S.qkv))
                fwd.append(o.build_window_attn_fwd(S.qkv, tables[l], att_f, batch=B, H=H, W=W, num_heads=eng.nH,
                                                   window=s.ws, shift=shift))
                fwd.append(split(att_f, S.att, HP))
                fwd.append(lin(S.att, Wt["proj"][l], Cp, HP, bias=Wt["bproj"][l], resid=xin, outF=xs[0],
                               acc_scale=dpa(l), tokens_per_sample=tps))
                fwd.append(o.build_layernorm_fwd(xs[0], None, None, C=C, xhat_bf16=S.xh2, rstd=S.r2, ones_col=C, terms=t))
                fwd.append(lin(S.xh2, Wt["fc1"][l], Fp, Cp, bias=Wt["bfc1"][l], outF=S.u))
                fwd.append(split(S.u, S.g, Fp, op=o.SPLIT_GELU, ones_col=s.hidden))
                fwd.append(lin(S.g, Wt["fc2"][l], Cp, Fp, bias=Wt["bfc2"][l], resid=xs[0], outF=xs[1],
                               acc_scale=dpm(l), tokens_per_sample=tps))
                if j < depth - 1:
                    nxt = blocks[l + 1]
                    fwd.append(o.build_layernorm_fwd(xs[1], None, None, C=C, xhat_bf16=nxt.xh1, rstd=nxt.r1, ones_col=C,
                                                     terms=t))
                    xin = xs[1]
                else:
                    fwd.append(split(xs[1], x2s[i], Cp, ones_col=C))
                l += 1
            t_next = tin[1] if t_cur is tin[0] else tin[0]
            fwd.append(lin(x2s[i], Wt["conv"][i], Cp, Cp, conv=geo, bias=Wt["bconv"][i], resid=t_cur, outF=t_next))
            if i < nl - 1:
                nxt = blocks[l]
                fwd.append(o.build_layernorm_fwd(t_next, None, None, C=C, xhat_bf16=nxt.xh1, rstd=nxt.r1, ones_col=C,
                                                 terms=t))
            else:
                tfin = t_next
            t_cur = t_next
        fwd.append(o.build_layernorm_fwd(tfin, P["norm.weight"], P["norm.bias"], C=C, y_bf16=yb_s, mean=meanF,
                                         rstd=rstdF, ones_col=C, terms=t))
        fwd.append(lin(yb_s, Wt["conv"][nl], Cp, Cp, conv=geo, bias=Wt["bconv"][nl], resid=first, outF=res_f))
        if not up:
            fwd.append(o.build_conv_out_fwd(res_f, cl_w, cl_b, self.out, mean_t, C=C, pitch=Cp, out_scale=1.0 / rng,
                                            skip=self.img, skip_mean=mean_t, skip_scale=rng))
        else:
            res_s = bf(T, t * Cp)
            fwd.append(split(res_f, res_s, Cp))
            if up == "pixelshuffledirect":
                no = Wt["psd"].shape[1]
                u = f32(T, no)
                fwd.append(lin(res_s, Wt["psd"][0], no, Cp, conv=geo, bias=Wt["bpsd"][0], outF=u))
                fwd.append(o.build_pixel_shuffle_out(u, self.out, mean_t, r=s.upscale, scale=1.0 / rng))
            else:
                nf = 64
                f = f32(T, nf)
                fwd.append(lin(res_s, Wt["cbu"][0], nf, Cp, conv=geo, bias=Wt["bcbu"][0], act=o.ACT_LRELU,
                               act_slope=0.01, outF=f))
                h, w = H, W
                if up == "pixelshuffle":
                    r = Wt["ups_r"]
                    for i in range(Wt["ups"].shape[0]):
                        fs, u = bf(B * h * w, t * nf), f32(B * h * w, r * r * nf)
                        fwd.append(split(f, fs, nf))
                        fwd.append(lin(fs, Wt["ups"][i], r * r * nf, nf, conv=(B, h, w), bias=Wt["bups"][i], outF=u))
                        f = f32(B * h * r * w * r, nf)
                        fwd.append(o.build_pixel_shuffle(u, f, batch=B, H=h, W=w, C=nf, r=r))
                        h, w = h * r, w * r
                else:
                    for i in range(2):
                        n_ = f32(B * 4 * h * w, nf)
                        fwd.append(o.build_upsample_nearest2(f, n_, batch=B, H=h, W=w, C=nf))
                        h, w = 2 * h, 2 * w
                        ns = bf(B * h * w, t * nf)
                        fwd.append(split(n_, ns, nf))
                        f = f32(B * h * w, nf)
                        fwd.append(lin(ns, Wt["ups"][i], nf, nf, conv=(B, h, w), bias=Wt["bups"][i], act=o.ACT_LRELU,
                                       act_slope=0.2, outF=f))
                    fs, f2 = bf(B * h * w, t * nf), f32(B * h * w, nf)
                    fwd.append(split(f, fs, nf))
                    fwd.append(lin(fs, Wt["ups"][2], nf, nf, conv=(B, h, w), bias=Wt["bups"][2], act=o.ACT_LRELU,
                                   act_slope=0.2, outF=f2))
                    f = f2
                fwd.append(o.build_conv_out_fwd(f, cl_w, cl_b, self.out, mean_t, C=nf, pitch=nf, out_scale=1.0 / rng))
        self.fwd = fwd
        self.bwd = None
        if not train:
            return

        # ================================================================================ backward
        self.dout = f32(B, s.in_chans, H, W)
        sizes = [p.numel() for p in P.values()]
        self.gflat = f32(sum(sizes))
        G, off = {}, 0
        for (k, p), n in zip(P.items(), sizes):
            G[k] = self.gflat[off:off + n].view(p.shape)
            off += n
        self.G = G
        n_qkv, n_proj, n_fc, n_conv = L * QP * Cp, L * Cp * HP, L * Fp * Cp, (nl + 1) * Cp * 9 * Cp
        self.dwflat = f32(n_qkv + n_proj + 2 * n_fc + n_conv)
        o0 = 0
        dWqkv = self.dwflat[o0:o0 + n_qkv].view(L, QP, Cp); o0 += n_qkv
        dWproj = self.dwflat[o0:o0 + n_proj].view(L, Cp, HP); o0 += n_proj
        dWfc1 = self.dwflat[o0:o0 + n_fc].view(L, Fp, Cp); o0 += n_fc
        dWfc2 = self.dwflat[o0:o0 + n_fc].view(L, Cp, Fp); o0 += n_fc
        dWconv = self.dwflat[o0:o0 + n_conv].view(nl + 1, Cp, 9 * Cp)
        self._thin = ()
        pairs = _pairs(t)
        bwd = []

        def wgrad(gs, n1, xs_, n2, D, conv=None):
            """D += sum over component pairs of gs_i^T . xs_j (column slices of the split operands)"""
            for ia, ib in pairs:
                bwd.append(o.build_gemm_nt(gs[:, ia * n1:(ia + 1) * n1], xs_[:, ib * n2:(ib + 1) * n2], D, n1=n1, n2=n2,
                                           bn=_tile(n2), conv=conv))

        R = [(f32(T, Cp), bf(T, t * Cp)) for _ in range(3)]   # rotating (fp32 gradient, its split operand)
        dres_f, dres_s, dy_f = f32(T, Cp), bf(T, t * Cp), f32(T, Cp)
        dg_f, du_s, acc_f = f32(T, Fp), bf(T, t * Fp), f32(T, Cp)
        datt_f, dqkv_f, dqkv_s = f32(T, HP), f32(T, QP), bf(T, t * QP)
        blk_names = eng._block_names()
        inv = 1.0 / rng
        # conv_last (thin, CUDA cores, fp32): input gradient and weight gradient
        bwd.append(o.build_conv_out_dgrad(self.dout, cl_w, C=C, pitch=Cp, out_scale=inv, dfeat_f32=dres_f))
        bwd.append(o.build_conv_out_wgrad(self.dout, res_f, cl_w, G["conv_last.weight"], G["conv_last.bias"], C=C,
                                          pitch=Cp, out_scale=inv))
        bwd.append(split(dres_f, dres_s, Cp))
        # conv_after_body, then the final LayerNorm
        wgrad(dres_s, Cp, yb_s, Cp, dWconv[nl], conv=geo)
        bwd.append(lin(dres_s, Wt["convT"][nl], Cp, Cp, conv=geo, outF=dy_f))
        dt = 2
        bwd.append(o.build_layernorm_bwd(tfin, meanF, rstdF, P["norm.weight"], G["norm.weight"], G["norm.bias"], C=C,
                                         dy_f32=dy_f, dx_f32=R[dt][0]))
        bwd.append(split(R[dt][0], R[dt][1], Cp))
        l_end = L
        for i in reversed(range(nl)):
            depth = s.depths[i]
            a, b = [k for k in range(3) if k != dt]
            wgrad(R[dt][1], Cp, x2s[i], Cp, dWconv[i], conv=geo)
            bwd.append(lin(R[dt][1], Wt["convT"][i], Cp, Cp, conv=geo, outF=R[a][0]))
            bwd.append(split(R[a][0], R[a][1], Cp, row_scale=dpm(l_end - 1), rows_per_sample=tps))
            cur, oth = a, b
            for j in reversed(range(depth)):
                l = l_end - depth + j
                S = blocks[l]
                wgrad(R[cur][1], Cp, S.g, Fp, dWfc2[l])
                bwd.append(lin(R[cur][1], Wt["fc2T"][l], Fp, Cp, outF=dg_f))
                bwd.append(split(dg_f, du_s, Fp, op=o.SPLIT_DGELU, aux=S.u))
                wgrad(du_s, Fp, S.xh2, Cp, dWfc1[l])
                bwd.append(lin(du_s, Wt["fc1T"][l], Cp, Fp, outF=acc_f))
                bwd.append(o.build_ln_bwd_split(acc_f, S.xh2, S.r2, R[oth][0], C=C, terms=t, resid=R[cur][0],
                                                out=R[oth][1], row_scale=dpa(l), rows_per_sample=tps))
                cur, oth = oth, cur
                wgrad(R[cur][1], Cp, S.att, HP, dWproj[l])
                bwd.append(lin(R[cur][1], Wt["projT"][l], HP, Cp, outF=datt_f))
                shift = s.ws // 2 if j % 2 == 1 else 0
                bwd.append(o.build_window_attn_bwd(S.qkv, tables[l], datt_f, dqkv_f,
                                                   G[blk_names[l] + "attn.relative_position_bias_table"], batch=B,
                                                   H=H, W=W, num_heads=eng.nH, window=s.ws, shift=shift))
                bwd.append(split(dqkv_f, dqkv_s, QP))
                wgrad(dqkv_s, QP, S.xh1, Cp, dWqkv[l])
                bwd.append(lin(dqkv_s, Wt["qkvT"][l], Cp, QP, outF=acc_f))
                bwd.append(o.build_ln_bwd_split(acc_f, S.xh1, S.r1, R[oth][0], C=C, terms=t, resid=R[cur][0],
                                                out=R[oth][1], row_scale=(dpm(l - 1) if j > 0 else None),
                                                rows_per_sample=tps))
                cur, oth = oth, cur
            # RSTB skip: d(t_in) = d(block path) + d(t_out)
            bwd.append(o.build_add_inplace(R[cur][0], R[dt][0]))
            bwd.append(split(R[cur][0], R[cur][1], Cp))
            dt = cur
            l_end -= depth
        if s.patch_norm:
            dfirst = R[(dt + 1) % 3][0]
            bwd.append(o.build_layernorm_bwd(first, mean0, rstd0, P["patch_embed.norm.weight"],
                                             G["patch_embed.norm.weight"], G["patch_embed.norm.bias"], C=C,
                                             dy_f32=R[dt][0], dx_add=dres_f, dx_f32=dfirst))
        else:
            bwd.append(o.build_add_inplace(R[dt][0], dres_f))
            dfirst = R[dt][0]
        self.dfirst = dfirst
        bwd.append(o.build_conv_in_wgrad(self.img, cf_w, mean_t, rng, G["conv_first.weight"], G["conv_first.bias"],
                                         O=C, O_pad=Cp, pitch=Cp, dy_f32=dfirst))
        gl = lambda suffix: [G[bn_ + suffix] for bn_ in blk_names]
        pl = lambda suffix: [P[bn_ + suffix] for bn_ in blk_names]
        bwd.append(o.build_unpack_linear_grad(dWqkv, pl("attn.qkv.weight"), pl("norm1.weight"), pl("norm1.bias"),
                                              gl("attn.qkv.weight"), gl("attn.qkv.bias") if s.qkv_bias else None,
                                              gl("norm1.weight"), gl("norm1.bias"), eng.map_qkv, eng.map_c,
                                              eng.scale_qkv, out_f=3 * C, in_f=C, ones_col=C))
        bwd.append(o.build_unpack_linear_grad(dWproj, pl("attn.proj.weight"), None, None, gl("attn.proj.weight"),
                                              gl("attn.proj.bias"), None, None, eng.map_c, eng.map_h, None,
                                              out_f=C, in_f=C, ones_col=eng.hd))
        bwd.append(o.build_unpack_linear_grad(dWfc1, pl("mlp.fc1.weight"), pl("norm2.weight"), pl("norm2.bias"),
                                              gl("mlp.fc1.weight"), gl("mlp.fc1.bias"), gl("norm2.weight"),
                                              gl("norm2.bias"), eng.map_f, eng.map_c, None, out_f=s.hidden, in_f=C,
                                              ones_col=C))
        bwd.append(o.build_unpack_linear_grad(dWfc2, pl("mlp.fc2.weight"), None, None, gl("mlp.fc2.weight"),
                                              gl("mlp.fc2.bias"), None, None, eng.map_c, eng.map_f, None, out_f=C,
                                              in_f=s.hidden, ones_col=s.hidden))
        cn = Wt["conv_names"]
        bwd.append(o.build_unpack_conv_grad(dWconv, [G[n + ".weight"] for n in cn], [G[n + ".bias"] for n in cn],
                                            O=C, I=C, I_pad=Cp, ones_col=C, center_tap=4))
        self.bwd = bwd
        self._dx_spec = dict(dhead=dfirst, W=cf_w, C=C, pitch=Cp, scale=rng, skip=True)
