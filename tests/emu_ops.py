"""TEST INFRASTRUCTURE: a PyTorch (CPU) emulation of every kair_b200.ops launch builder.

It mirrors the *documented semantics* of the C-ABI kernels (include/kair_b200.h) -- padding, ones columns, bf16
rounding points, accumulate-vs-overwrite -- so that (i) the host-side engines (launch-list construction, weight
packing maps, gradient routing) can be checked on a machine without a GPU against the oracle / the reference, and
(ii) the GPU unit tests have a second, independent statement of what each kernel must produce.

It is never imported by the product package: tests inject it with `Engine(..., ops=emu_ops)`.
"""
import math

import torch
import torch.nn.functional as F

EPI_LINEAR, EPI_GELU, EPI_RESID_LN, EPI_DGELU, EPI_LNBWD = range(5)
ACT_NONE, ACT_LRELU = 0, 1


class EmuLaunch:
    def __init__(self, name, fn):
        self.name, self.fn = name, fn

    def run(self, stream=None):
        self.fn()


def stream_handle(device=None):
    return None


class _EmuList:
    def __init__(self, launches):
        self.launches = list(launches)

    def run(self, stream=None):
        for l in self.launches:
            l.run(stream)

    def __len__(self):
        return len(self.launches)

    def __iter__(self):
        return iter(self.launches)


def make_list(launches):
    return _EmuList(launches)


def device_guard(device):
    import contextlib
    return contextlib.nullcontext()


def zeros(shape, dtype, device):
    return torch.zeros(shape, dtype=dtype, device=device)


def _bf(x):
    return x.to(torch.bfloat16)


SPLIT_COPY, SPLIT_GELU, SPLIT_DGELU, SPLIT_ACT_BWD, SPLIT_LRELU = range(5)


def _components(x):
    """bf16 components of an fp32 tensor (kair_b200.h 'SPLIT PRECISION'): c0 + c1 + c2 ~= x"""
    c0 = x.to(torch.bfloat16).float()
    r1 = x - c0
    c1 = r1.to(torch.bfloat16).float()
    c2 = (r1 - c1).to(torch.bfloat16).float()
    return [c0, c1, c2]


def _seg_act(terms):
    return {1: [0], 3: [0, 1, 0], 6: [0, 1, 2, 0, 1, 0]}[terms]


def _seg_w(terms):
    return {1: [0], 3: [0, 0, 1], 6: [0, 0, 0, 1, 1, 2]}[terms]


def _split_act(x, terms):
    """[T, n] fp32 -> [T, terms*n] bf16 activation-side operand"""
    c = _components(x)
    return torch.cat([c[w] for w in _seg_act(terms)], dim=1).to(torch.bfloat16)


def _gelu_grad(x):
    return 0.5 * (1.0 + torch.erf(x * 0.7071067811865476)) + x * torch.exp(-0.5 * x * x) * 0.3989422804014327


def _conv_acc(A, Bw, conv, K, N, dilation=1):
    """implicit-GEMM 3x3: A NHWC [B,H,W,>=K] bf16, Bw [N, 9*K] with k = tap*K + ci."""
    Bn, H, W = conv
    x = A.reshape(Bn, H, W, -1)[..., :K].float().permute(0, 3, 1, 2)
    w = Bw[:N, :9 * K].float().reshape(N, 9, K).permute(0, 2, 1).reshape(N, K, 3, 3)
    y = F.conv2d(x, w, padding=dilation, dilation=dilation)
    return y.permute(0, 2, 3, 1).reshape(Bn * H * W, N)


def build_gemm_tn(A, B, *, N, K, bn, mode=EPI_LINEAR, act=ACT_NONE, act_slope=0.0, ln_eps=1e-5, bias=None, resid=None,
                  aux=None, rowstat=None, outB=None, outB2=None, outF=None, n_valid=0, ones_col=-1, conv=None,
                  acc_scale=None, outB_scale=None, tokens_per_sample=0, dilation=1):
    assert K % 64 == 0 and N % bn == 0 and bn in (64, 128, 192, 256)
    if mode in (EPI_RESID_LN, EPI_LNBWD):
        assert N == bn

    def run():
        M = conv[0] * conv[1] * conv[2] if conv is not None else A.shape[0]
        if conv is not None:
            acc = _conv_acc(A, B, conv, K, N, dilation)
        else:
            acc = A[:, :K].float() @ B[:N, :K].float().t()
        nv = n_valid if n_valid > 0 else N
        cols = torch.arange(N)

        def flat(t, width):
            return t.reshape(M, -1)[:, :width]

        def per_sample(s):
            return s.repeat_interleave(tokens_per_sample)[:M].unsqueeze(1)

        def put_bf(dst, v, ones=True):
            v = v.clone()
            if ones and 0 <= ones_col < N:
                v[:, ones_col] = 1.0
            flat(dst, N).copy_(_bf(v))

        if mode == EPI_LINEAR:
            v = acc + (bias[:N] if bias is not None else 0.0)
            if acc_scale is not None:
                v = v * per_sample(acc_scale)
            if aux is not None:
                assert act == ACT_NONE
                v = torch.where(flat(aux, N).float() > 0, v, v * act_slope)
            if resid is not None:
                v = v + flat(resid, N)
            if act == ACT_LRELU:
                v = torch.where(v > 0, v, v * act_slope)
            if outF is not None:
                flat(outF, N).copy_(v)
            if outB is not None:
                vb = v * per_sample(outB_scale) if outB_scale is not None else v
                put_bf(outB, vb)
        elif mode == EPI_GELU:
            u = acc + (bias[:N] if bias is not None else 0.0)
            if outB2 is not None:
                flat(outB2, N).copy_(_bf(u))
            put_bf(outB, F.gelu(u))
        elif mode == EPI_RESID_LN:
            v = acc + (bias[:N] if bias is not None else 0.0)
            if acc_scale is not None:
                v = v * per_sample(acc_scale)
            if resid is not None:
                v = v + flat(resid, N)
            if outF is not None:
                flat(outF, N).copy_(v)
            mean = v.sum(1, keepdim=True) / nv
            var = (((v - mean) ** 2).sum(1, keepdim=True) - (N - nv) * mean * mean).clamp_min(0) / nv
            rstd = torch.rsqrt(var + ln_eps)
            if rowstat is not None:
                rowstat[:M].copy_(rstd[:, 0])
            xh = (v - mean) * rstd
            xh[:, nv:] = 0.0
            put_bf(outB, xh)
        elif mode == EPI_DGELU:
            put_bf(outB, acc * _gelu_grad(flat(aux, N).float()), ones=False)
        elif mode == EPI_LNBWD:
            xh = flat(aux, N).float()
            m1 = acc.sum(1, keepdim=True) / nv
            m2 = (acc * xh).sum(1, keepdim=True) / nv
            v = rowstat[:M].unsqueeze(1) * (acc - m1 - xh * m2)
            if resid is not None:
                v = v + flat(resid, N)
            v[:, nv:] = 0.0
            if outF is not None:
                flat(outF, N).copy_(v)
            if outB is not None:
                vb = v * per_sample(outB_scale) if outB_scale is not None else v
                put_bf(outB, vb, ones=False)
        else:
            raise AssertionError(mode)
        del cols

    return EmuLaunch("kair_gemm_tn", run)


def build_mlp_fused(X, W1, W2, *, N, F, bias1=None, bias2=None, resid=None, outF=None, outB=None, outU=None, outG=None,
                    rowstat=None, acc_scale=None, tokens_per_sample=0, ln=True, n_valid=0, ones_col=-1, ones_col_g=-1,
                    ln_eps=1e-5):
    """stated as the unfused pair it replaces (kair_b200.h: bit-identical to it)"""
    assert F == 2 * N and N in (64, 192, 256) and (outF is not None or outB is not None)
    M = X.shape[0]
    g = outG if outG is not None else torch.zeros(M, F, dtype=X.dtype)
    first = build_gemm_tn(X, W1, N=F, K=N, bn=64, mode=EPI_GELU, bias=bias1, outB=g, outB2=outU, ones_col=ones_col_g)
    second = build_gemm_tn(g, W2, N=N, K=F, bn=N, mode=EPI_RESID_LN if ln else EPI_LINEAR, bias=bias2, resid=resid,
                           outF=outF, outB=outB, rowstat=rowstat if ln else None, n_valid=n_valid, ones_col=ones_col,
                           acc_scale=acc_scale, tokens_per_sample=tokens_per_sample, ln_eps=ln_eps)

    def run():
        first.run()
        second.run()

    return EmuLaunch("kair_mlp_fused", run)


def build_gemm_nt(A, B, D, *, n1, n2, bn, splits=0, conv=None, dilation=1, swap=False):
    assert n1 % 64 == 0 and (n1 % bn == 0 and n2 % 64 == 0 and conv is not None if swap else n2 % bn == 0)
    d = dilation

    def run():
        if conv is None:
            D[:n1, :n2] += A[:, :n1].float().t() @ B[:, :n2].float()
            return
        Bn, H, W = conv
        a = A.reshape(Bn, H, W, -1)[..., :n1].float()
        b = B.reshape(Bn, H, W, -1)[..., :n2].float()
        bp = F.pad(b, (0, 0, d, d, d, d))
        for tap in range(9):
            dy, dx = tap // 3, tap % 3
            bs = bp[:, dy * d:dy * d + H, dx * d:dx * d + W, :]
            if swap:   # transposed result: rows tap*n2 + input channel, columns output channel
                D[tap * n2:(tap + 1) * n2, :n1] += torch.einsum("bhwi,bhwj->ji", a, bs)
            else:
                D[:n1, tap * n2:(tap + 1) * n2] += torch.einsum("bhwi,bhwj->ij", a, bs)

    return EmuLaunch("kair_gemm_nt", run)


# ------------------------------------------------------------------------------------------------ window attention

def _win_index(H, W, ws, shift):
    wi, wj = torch.meshgrid(torch.arange(H // ws), torch.arange(W // ws), indexing="ij")
    r, c = torch.meshgrid(torch.arange(ws), torch.arange(ws), indexing="ij")
    h = (wi.reshape(-1, 1) * ws + r.reshape(1, -1) + shift) % H
    w = (wj.reshape(-1, 1) * ws + c.reshape(1, -1) + shift) % W
    return h * W + w


def _rel_index(ws):
    ys, xs = torch.meshgrid(torch.arange(ws), torch.arange(ws), indexing="ij")
    ys, xs = ys.reshape(-1), xs.reshape(-1)
    return (ys[:, None] - ys[None, :] + ws - 1) * (2 * ws - 1) + (xs[:, None] - xs[None, :] + ws - 1)


def _mask(H, W, ws, shift):
    def rid(n):
        p = torch.arange(n)
        return (p >= n - ws).long() + (p >= n - shift).long()
    ids = (3 * rid(H)[:, None] + rid(W)[None, :]).float()
    ids = ids.view(H // ws, ws, W // ws, ws).permute(0, 2, 1, 3).reshape(-1, ws * ws)
    d = ids[:, None, :] - ids[:, :, None]
    return torch.where(d != 0, torch.full_like(d, -100.0), torch.zeros_like(d))


def _attn_core(qkv, table, batch, H, W, nH, ws, shift):
    """qkv fp32 [T, 3*nH*32] -> out [T, nH*32] (autograd-capable)."""
    N = ws * ws
    idx = _win_index(H, W, ws, shift)
    nW = idx.shape[0]
    x = qkv.view(batch, H * W, 3, nH, 32)[:, idx.reshape(-1)]
    x = x.view(batch * nW, N, 3, nH, 32).permute(2, 0, 3, 1, 4)
    q, k, v = x[0], x[1], x[2]
    attn = q @ k.transpose(-2, -1)
    bias = table[_rel_index(ws).reshape(-1)].view(N, N, nH).permute(2, 0, 1)
    attn = attn + bias.unsqueeze(0)
    if shift > 0:
        attn = attn.view(batch, nW, nH, N, N) + _mask(H, W, ws, shift)[None, :, None]
        attn = attn.view(-1, nH, N, N)
    attn = attn.softmax(-1)
    o = (attn @ v).transpose(1, 2).reshape(batch, nW * N, nH * 32)
    out = torch.zeros(batch, H * W, nH * 32)
    out = out.index_copy(1, idx.reshape(-1), o)
    return out.view(batch * H * W, nH * 32)


def build_window_attn_fwd(qkv, table, out, *, batch, H, W, num_heads, window=8, shift=0):
    assert 2 <= window <= 8 and 0 <= shift < window

    def run():
        w = 3 * num_heads * 32
        o = _attn_core(qkv.reshape(batch * H * W, -1)[:, :w].float(), table.float(), batch, H, W, num_heads, window, shift)
        out.reshape(batch * H * W, -1)[:, :num_heads * 32].copy_(o.to(out.dtype))

    return EmuLaunch("kair_window_attn_fwd", run)


def build_window_attn_bwd(qkv, table, dout, dqkv, dtable, *, batch, H, W, num_heads, window=8, shift=0):
    def run():
        T = batch * H * W
        w = 3 * num_heads * 32
        with torch.enable_grad():
            q = qkv.reshape(T, -1)[:, :w].float().detach().requires_grad_(True)
            t = table.float().detach().requires_grad_(True)
            o = _attn_core(q, t, batch, H, W, num_heads, window, shift)
            (o * dout.reshape(T, -1)[:, :num_heads * 32].float()).sum().backward()
        dqkv.reshape(T, -1)[:, :w].copy_(q.grad.to(dqkv.dtype))
        dtable.add_(t.grad)

    return EmuLaunch("kair_window_attn_bwd", run)


# ------------------------------------------------------------------------------------------------ packing

def build_pack_linear(W, b, gamma, beta, row_map, col_map, row_scale, bias_const, Wp, WpT, bp, *, out_f, in_f, terms=1):
    def run():
        L, N_pad = Wp.shape[0], Wp.shape[1]
        K_pad = Wp.shape[2] // terms
        rm, cm = row_map.long(), col_map.long()
        for l in range(L):
            w = W[l].detach()
            rs = row_scale if row_scale is not None else torch.ones(N_pad)
            g = gamma[l].detach() if gamma is not None else torch.ones(in_f)
            full = torch.zeros(N_pad, K_pad)
            rv, cv = rm >= 0, cm >= 0
            sub = w[rm[rv]][:, cm[cv]] * g[cm[cv]][None, :] * rs[rv][:, None]
            full[rv.nonzero()[:, 0][:, None], cv.nonzero()[:, 0][None, :]] = sub
            if terms == 1:
                Wp[l].copy_(_bf(full))
                if WpT is not None:
                    WpT[l].copy_(_bf(full.t()))
            else:
                c = _components(full)
                Wp[l].copy_(_bf(torch.cat([c[w] for w in _seg_w(terms)], dim=1)))
                if WpT is not None:
                    WpT[l].copy_(_bf(torch.cat([c[w].t() for w in _seg_w(terms)], dim=1)))
            if bp is not None:
                bias = torch.zeros(N_pad)
                base = b[l].detach()[rm[rv]] if b is not None else torch.zeros(int(rv.sum()))
                if beta is not None:
                    base = base + w[rm[rv]] @ beta[l].detach()
                bias[rv] = base * rs[rv]
                if bias_const is not None:
                    bias = bias + bias_const
                bp[l].copy_(bias)

    return EmuLaunch("kair_pack_linear", run)


def build_unpack_linear_grad(dWp, W, gamma, beta, dW, db, dgamma, dbeta, row_map, col_map, row_scale, *, out_f, in_f,
                             ones_col):
    def run():
        L, N_pad, K_pad = dWp.shape
        rm, cm = row_map.long(), col_map.long()
        rv, cv = rm >= 0, cm >= 0
        rs = row_scale if row_scale is not None else torch.ones(N_pad)
        for l in range(L):
            g = gamma[l].detach() if gamma is not None else torch.ones(in_f)
            d = dWp[l]
            sub = d[rv][:, cv] * rs[rv][:, None]          # d loss / d (W*gamma) at (src row, src col)
            dWl = sub * g[cm[cv]][None, :]
            if beta is not None:  # W also enters bp = rs * (b + W . beta)
                dWl = dWl + (d[rv, ones_col] * rs[rv])[:, None] * beta[l].detach()[cm[cv]][None, :]
            dW[l][rm[rv][:, None], cm[cv][None, :]] = dWl
            if db is not None:
                db[l][rm[rv]] = d[rv, ones_col] * rs[rv]
            if dgamma is not None:
                w = W[l].detach()[rm[rv]][:, cm[cv]] * rs[rv][:, None]
                dgamma[l][cm[cv]] = (w * d[rv][:, cv]).sum(0)
                dbeta[l][cm[cv]] = (w * d[rv, ones_col][:, None]).sum(0)

    return EmuLaunch("kair_unpack_linear_grad", run)


def build_pack_conv(W, b, Wp, WpT, bp, *, O, I, I_pad, taps=9, oscale=None, oshift=None, terms=1):
    def run():
        L, O_pad = Wp.shape[0], Wp.shape[1]
        for l in range(L):
            w = W[l].detach().reshape(O, I, taps)
            if oscale is not None:
                w = w * oscale[l].view(O, 1, 1)
            full = torch.zeros(O_pad, taps, I_pad)
            full[:O, :, :I] = w.permute(0, 2, 1)
            def segs(m):   # [rows, taps, cols] -> [rows, taps * terms * cols] with k = (tap*terms + seg)*cols + i
                if terms == 1:
                    return _bf(m.reshape(m.shape[0], -1))
                c = _components(m)
                return _bf(torch.stack([c[w] for w in _seg_w(terms)], dim=2).reshape(m.shape[0], -1))
            Wp[l].copy_(segs(full))
            if WpT is not None:
                t = torch.zeros(I_pad, taps, O_pad)
                t[:I, :, :O] = w.flip(2).permute(1, 2, 0)
                WpT[l].copy_(segs(t))
            if bp is not None:
                bias = torch.zeros(O_pad)
                if b is not None:
                    bias[:O] = b[l].detach()
                if oscale is not None:
                    bias[:O] = bias[:O] * oscale[l]
                if oshift is not None:
                    bias[:O] = bias[:O] + oshift[l]
                bp[l].copy_(bias)

    return EmuLaunch("kair_pack_conv", run)


def build_unpack_conv_grad(dWp, dW, db, *, O, I, I_pad, taps=9, ones_col=-1, center_tap=4, dwp_pitch=0, transposed=False):
    def run():
        L, O_pad = dWp.shape[0], dWp.shape[1]
        for l in range(L):
            if transposed:   # [taps*I_pad, O_pad] -> (O, I, taps)
                assert db is None
                d = dWp[l].reshape(taps, I_pad, -1)
                dW[l].copy_(d[:, :I, :O].permute(2, 1, 0).reshape(dW[l].shape))
                continue
            d = dWp[l].reshape(O_pad, -1)[:, :taps * I_pad].reshape(O_pad, taps, I_pad)
            dW[l].copy_(d[:O, :, :I].permute(0, 2, 1).reshape(dW[l].shape))
            if db is not None and ones_col >= 0:
                db[l].copy_(d[:O, center_tap, ones_col])

    return EmuLaunch("kair_unpack_conv_grad", run)


# ------------------------------------------------------------------------------------------------ LayerNorm

def build_layernorm_fwd(x, gamma, beta, *, C, eps=1e-5, y_f32=None, y_bf16=None, xhat_bf16=None, mean=None, rstd=None,
                        ones_col=-1, terms=1):
    def run():
        v = x[:, :C]
        m = v.mean(1, keepdim=True)
        r = torch.rsqrt(((v - m) ** 2).mean(1, keepdim=True) + eps)
        xh = (v - m) * r
        if mean is not None:
            mean.copy_(m[:, 0])
        if rstd is not None:
            rstd.copy_(r[:, 0])
        pitch = x.shape[1]

        def padded(t):
            o = torch.zeros(x.shape[0], pitch)
            o[:, :C] = t
            return o
        if y_f32 is not None or y_bf16 is not None:
            y = padded(xh * gamma.detach() + beta.detach())
            if y_f32 is not None:
                y_f32.copy_(y)
            if y_bf16 is not None:
                if ones_col >= 0:
                    y[:, ones_col] = 1.0
                y_bf16.copy_(_split_act(y, terms) if terms > 1 else _bf(y))
        if xhat_bf16 is not None:
            h = padded(xh)
            if ones_col >= 0:
                h[:, ones_col] = 1.0
            xhat_bf16.copy_(_split_act(h, terms) if terms > 1 else _bf(h))

    return EmuLaunch("kair_layernorm_fwd", run)


def build_layernorm_bwd(x, mean, rstd, gamma, dgamma, dbeta, *, C, dy_f32=None, dy_bf16=None, dx_add=None, dx_f32=None,
                        dx_bf16=None):
    def run():
        dy = torch.zeros(x.shape[0], C)
        if dy_f32 is not None:
            dy = dy + dy_f32[:, :C]
        if dy_bf16 is not None:
            dy = dy + dy_bf16[:, :C].float()
        xh = (x[:, :C] - mean[:, None]) * rstd[:, None]
        g = dy * gamma.detach()
        dx = rstd[:, None] * (g - g.mean(1, keepdim=True) - xh * (g * xh).mean(1, keepdim=True))
        if dx_add is not None:
            dx = dx + dx_add[:, :C]
        full = torch.zeros_like(x)
        full[:, :C] = dx
        if dx_f32 is not None:
            dx_f32.copy_(full)
        if dx_bf16 is not None:
            dx_bf16.copy_(_bf(full))
        dgamma.add_((dy * xh).sum(0))
        dbeta.add_(dy.sum(0))

    return EmuLaunch("kair_layernorm_bwd", run)


# ------------------------------------------------------------------------------------------------ thin convolutions

def _tokens_to_nchw(t, B, H, W, C):
    return t.reshape(B, H, W, -1)[..., :C].permute(0, 3, 1, 2)


def build_conv_in_fwd(img, W, b, mean, rng, *, O, O_pad, pitch, out_f32=None, out_bf16=None, act=ACT_NONE, act_slope=0.0,
                      ones_col=-1):
    def run():
        Bn, ci, H, Wd = img.shape
        xin = (img - mean.view(1, ci, 1, 1)) * rng
        y = F.conv2d(xin, W.detach(), b.detach() if b is not None else None, padding=1)
        if act == ACT_LRELU:
            y = torch.where(y > 0, y, y * act_slope)
        full = torch.zeros(Bn * H * Wd, pitch)
        full[:, :O] = y.permute(0, 2, 3, 1).reshape(-1, O)
        if out_f32 is not None:
            out_f32.copy_(full)
        if out_bf16 is not None:
            if ones_col >= 0:
                full[:, ones_col] = 1.0
            out_bf16.copy_(_bf(full))

    return EmuLaunch("kair_conv_in_fwd", run)


def build_conv_in_wgrad(img, W, mean, rng, dW, db, *, O, O_pad, pitch, dy_f32=None, dy_bf16=None):
    def run():
        Bn, ci, H, Wd = img.shape
        dy = torch.zeros(Bn * H * Wd, O)
        if dy_f32 is not None:
            dy = dy + dy_f32[:, :O]
        if dy_bf16 is not None:
            dy = dy + dy_bf16[:, :O].float()
        dyi = _tokens_to_nchw(dy, Bn, H, Wd, O)
        xin = (img - mean.view(1, ci, 1, 1)) * rng
        dW.add_(torch.nn.grad.conv2d_weight(xin, W.shape, dyi.contiguous(), padding=1))
        if db is not None:
            db.add_(dyi.sum((0, 2, 3)))

    return EmuLaunch("kair_conv_in_wgrad", run)


def build_conv_out_fwd(feat, W, b, out, out_mean, *, C, pitch, out_scale=1.0, skip=None, skip_mean=None, skip_scale=0.0):
    def run():
        Bn, co, H, Wd = out.shape
        x = _tokens_to_nchw(feat.float(), Bn, H, Wd, C)
        y = F.conv2d(x, W.detach(), b.detach() if b is not None else None, padding=1)
        if skip is not None:
            y = y + skip_scale * (skip - skip_mean.view(1, co, 1, 1))
        out.copy_(y * out_scale + out_mean.view(1, co, 1, 1))

    return EmuLaunch("kair_conv_out_fwd", run)


def build_conv_out_dgrad(dout, W, *, C, pitch, out_scale=1.0, dfeat_f32=None, dfeat_bf16=None, mask=None,
                         mask_slope=0.0):
    def run():
        Bn, co, H, Wd = dout.shape
        dx = torch.nn.grad.conv2d_input((Bn, C, H, Wd), W.detach(), dout * out_scale, padding=1)
        full = torch.zeros(Bn * H * Wd, pitch)
        full[:, :C] = dx.permute(0, 2, 3, 1).reshape(-1, C)
        if mask is not None:
            full = torch.where(mask.reshape(-1, pitch).float() > 0, full, full * mask_slope)
        if dfeat_f32 is not None:
            dfeat_f32.copy_(full)
        if dfeat_bf16 is not None:
            dfeat_bf16.copy_(_bf(full))

    return EmuLaunch("kair_conv_out_dgrad", run)


def build_conv_out_wgrad(dout, feat, W, dW, db, *, C, pitch, out_scale=1.0):
    def run():
        Bn, co, H, Wd = dout.shape
        x = _tokens_to_nchw(feat.float(), Bn, H, Wd, C).contiguous()
        dW.add_(torch.nn.grad.conv2d_weight(x, W.shape, (dout * out_scale).contiguous(), padding=1))
        if db is not None:
            db.add_((dout * out_scale).sum((0, 2, 3)))

    return EmuLaunch("kair_conv_out_wgrad", run)


def build_im2col3x3(img, out, *, cpad, direction, mean=None, scale=1.0, ones_col=-1):
    def run():
        B, c, H, W = img.shape
        x = img if mean is None else img - mean.view(1, -1, 1, 1)
        x = F.pad(x * scale, (1, 1, 1, 1))
        res = torch.zeros(B * H * W, out.shape[1])
        for tap in range(9):
            dy, dx = direction * (tap // 3 - 1), direction * (tap % 3 - 1)
            sh = x[:, :, 1 + dy:1 + dy + H, 1 + dx:1 + dx + W]
            res[:, tap * cpad:tap * cpad + c] = sh.permute(0, 2, 3, 1).reshape(-1, c)
        if ones_col >= 0:
            res[:, ones_col] = 1.0
        out.copy_(_bf(res))
    return EmuLaunch("kair_im2col3x3", run)


def build_conv_out_pack(*, mode, co, C, cpad, W=None, Wt=None, D=None, dW=None, db=None, ones_row=-1):
    def run():
        if mode == 0:
            res = torch.zeros(Wt.shape, dtype=torch.float32)
            for tap in range(9):
                res[:C, tap * cpad:tap * cpad + co] = W.reshape(co, C, 9)[:, :, tap].t()
            Wt.copy_(_bf(res))
        else:
            g = torch.stack([D[:C, tap * cpad:tap * cpad + co] for tap in range(9)], dim=2)  # [C, co, 9]
            dW.copy_(g.permute(1, 0, 2).reshape(dW.shape))
            if db is not None:
                db.copy_(D[ones_row, 4 * cpad:4 * cpad + co])
    return EmuLaunch("kair_conv_out_pack", run)


def build_split_terms(x, out, *, n, terms, op=SPLIT_COPY, aux=None, out_f32=None, row_scale=None, rows_per_sample=0,
                      ones_col=-1, slope=0.0):
    def run():
        v = x[:, :n].float().clone()
        if op == SPLIT_GELU:
            v = F.gelu(v)
        elif op == SPLIT_LRELU:
            v = torch.where(v > 0, v, v * slope)
        elif op == SPLIT_DGELU:
            v = v * _gelu_grad(aux[:, :n].float())
        elif op == SPLIT_ACT_BWD:
            v = torch.where(aux[:, :n] > 0, v, v * slope)
        if row_scale is not None:
            v = v * row_scale.repeat_interleave(rows_per_sample)[:, None]
        if ones_col >= 0:
            v[:, ones_col] = 1.0
        if out_f32 is not None:
            out_f32[:, :n].copy_(v)
        if out is not None:
            out[:, :terms * n].copy_(_split_act(v, terms))

    return EmuLaunch("kair_split_terms", run)


def build_ln_bwd_split(g, xhat, rstd, outF, *, C, terms, resid=None, out=None, row_scale=None, rows_per_sample=0):
    def run():
        pitch = g.shape[1]
        nc = {3: 2, 6: 3}[terms]
        xh = sum(xhat[:, k * pitch:k * pitch + C].float() for k in reversed(range(nc)))
        gg = g[:, :C]
        m1 = gg.mean(1, keepdim=True)
        m2 = (gg * xh).mean(1, keepdim=True)
        dx = torch.zeros(g.shape[0], pitch)
        dx[:, :C] = rstd[:, None] * (gg - m1 - xh * m2)
        if resid is not None:
            dx[:, :C] += resid[:, :C]
        outF.copy_(dx)
        if out is not None:
            sc = row_scale.repeat_interleave(rows_per_sample)[:, None] if row_scale is not None else 1.0
            out[:, :terms * pitch].copy_(_split_act(dx * sc, terms))

    return EmuLaunch("kair_ln_bwd_split", run)


def build_add_inplace(y, x, yb=None):
    def run():
        y.add_(x)
        if yb is not None:
            yb.copy_(_bf(y))

    return EmuLaunch("kair_add_inplace", run)


# ------------------------------------------------------------------------------------------------ conv-net helpers

def build_space_to_depth(fine, coarse, *, batch, H, W, C):
    def run():
        x = fine.reshape(batch, H // 2, 2, W // 2, 2, -1)[..., :C]
        coarse.reshape(batch * (H // 2) * (W // 2), -1)[:, :4 * C].copy_(
            x.permute(0, 1, 3, 2, 4, 5).reshape(-1, 4 * C))

    return EmuLaunch("kair_space_to_depth", run)


def build_depth_to_space(coarse, *, batch, H, W, C, out_f32=None, out_bf16=None):
    def run():
        x = coarse.reshape(batch, H // 2, W // 2, -1)[..., :4 * C].float().reshape(batch, H // 2, W // 2, 2, 2, C)
        y = x.permute(0, 1, 3, 2, 4, 5).reshape(batch * H * W, C)
        if out_f32 is not None:
            out_f32.reshape(batch * H * W, -1)[:, :C].copy_(y)
        if out_bf16 is not None:
            out_bf16.reshape(batch * H * W, -1)[:, :C].copy_(_bf(y))

    return EmuLaunch("kair_depth_to_space", run)


def build_colsum(x, out, *, C, scale=1.0):
    def run():
        out[:C].add_(scale * x[:, :C].float().sum(0))

    return EmuLaunch("kair_colsum", run)


def build_bn_stats(z, acc, *, C):
    def run():
        acc[0, :C] += z[:, :C].double().sum(0)
        acc[1, :C] += (z[:, :C].double() ** 2).sum(0)

    return EmuLaunch("kair_bn_stats", run)


def build_bn_finalize(acc, mean, rstd, running_mean, running_var, num_batches_tracked, *, C, T, eps, momentum):
    def run():
        m = acc[0, :C] / T
        var = (acc[1, :C] / T - m * m).clamp_min(0)
        mean[:C].copy_(m.float())
        rstd[:C].copy_((1.0 / torch.sqrt(var + eps)).float())
        with torch.no_grad():
            if running_mean is not None:
                running_mean.mul_(1 - momentum).add_(momentum * m.float())
            if running_var is not None:
                running_var.mul_(1 - momentum).add_(momentum * (var * T / max(T - 1, 1)).float())
            if num_batches_tracked is not None:
                num_batches_tracked.add_(1)
        acc.zero_()

    return EmuLaunch("kair_bn_finalize", run)


def _act_mask(y, act, slope):
    if act != ACT_LRELU:
        return 1.0
    return torch.where(y.float() > 0, torch.ones(()), torch.full((), float(slope)))


def build_bn_apply(z, mean, rstd, gamma, beta, *, C, act=ACT_NONE, act_slope=0.0, out_bf16=None, out_f32=None):
    def run():
        v = (z[:, :C] - mean[:C]) * (gamma.detach() * rstd[:C]) + beta.detach()
        if act == ACT_LRELU:
            v = torch.where(v > 0, v, v * act_slope)
        if out_bf16 is not None:
            out_bf16[:, :C].copy_(_bf(v))
        if out_f32 is not None:
            out_f32[:, :C].copy_(v)

    return EmuLaunch("kair_bn_apply", run)


def build_bn_bwd_reduce(z, y, dy, mean, rstd, dgamma, dbeta, *, C, act=ACT_NONE, act_slope=0.0):
    def run():
        g = dy[:, :C].float() * _act_mask(y[:, :C], act, act_slope)
        xh = (z[:, :C] - mean[:C]) * rstd[:C]
        dbeta.add_(g.sum(0))
        dgamma.add_((g * xh).sum(0))

    return EmuLaunch("kair_bn_bwd_reduce", run)


def build_bn_bwd_apply(z, y, dy, mean, rstd, gamma, dgamma, dbeta, dz, *, C, act=ACT_NONE, act_slope=0.0, dbias=None):
    def run():
        T = z.shape[0]
        g = dy[:, :C].float() * _act_mask(y[:, :C], act, act_slope)
        xh = (z[:, :C] - mean[:C]) * rstd[:C]
        d = gamma.detach() * rstd[:C] * (g - dbeta / T - xh * dgamma / T)
        dz[:, :C].copy_(_bf(d))
        if dbias is not None:
            dbias.add_(d.sum(0))

    return EmuLaunch("kair_bn_bwd_apply", run)


def build_bn_fold(gamma, beta, running_mean, running_var, scale, shift, *, C, eps):
    def run():
        sc = gamma.detach() * torch.rsqrt(running_var + eps)
        scale.copy_(sc)
        shift.copy_(beta.detach() - running_mean * sc)

    return EmuLaunch("kair_bn_fold", run)


def build_pixel_shuffle(x, out, *, batch, H, W, C, r):
    def run():
        v = x.reshape(batch, H, W, -1)[..., :C * r * r].reshape(batch, H, W, C, r, r)
        y = v.permute(0, 1, 4, 2, 5, 3).reshape(batch * H * r * W * r, C)
        out.reshape(batch * H * r * W * r, -1)[:, :C].copy_(y)

    return EmuLaunch("kair_pixel_shuffle", run)


def build_upsample_nearest2(x, out, *, batch, H, W, C):
    def run():
        v = x.reshape(batch, H, 1, W, 1, -1)[..., :C].expand(batch, H, 2, W, 2, C)
        out.reshape(batch * 4 * H * W, -1)[:, :C].copy_(v.reshape(-1, C))

    return EmuLaunch("kair_upsample_nearest2", run)


def build_copy(dst, src):
    return EmuLaunch("copy", lambda: dst.copy_(src))


def build_pixel_shuffle_out(x, out, mean, *, r, scale=1.0):
    def run():
        Bn, co, Ho, Wo = out.shape
        H, W = Ho // r, Wo // r
        v = x.reshape(Bn, H, W, -1)[..., :co * r * r].float().permute(0, 3, 1, 2)
        out.copy_(F.pixel_shuffle(v, r) * scale + mean.view(1, co, 1, 1))

    return EmuLaunch("kair_pixel_shuffle_out", run)


def build_pixel_shuffle_bwd(dy, dx, *, batch, H, W, C, r):
    def run():
        g = dy.reshape(batch, H * r, W * r, -1)[..., :C].float().permute(0, 3, 1, 2)
        dx[:, :C * r * r] = _bf(F.pixel_unshuffle(g, r).permute(0, 2, 3, 1).reshape(batch * H * W, C * r * r))

    return EmuLaunch("kair_pixel_shuffle_bwd", run)


def build_upsample_nearest2_bwd(dy, dx, *, batch, H, W, C, mask=None, mask_slope=0.0):
    def run():
        g = dy.reshape(batch, H, 2, W, 2, -1)[..., :C].float().sum(dim=(2, 4)).reshape(batch * H * W, C)
        if mask is not None:
            g = torch.where(mask[:, :C].float() > 0, g, g * mask_slope)
        dx[:, :C] = _bf(g)

    return EmuLaunch("kair_upsample_nearest2_bwd", run)


def build_pixel_shuffle_out_bwd(dout, dx, *, r, co, scale=1.0):
    def run():
        Bn, _, Ho, Wo = dout.shape
        g = F.pixel_unshuffle(dout * scale, r).permute(0, 2, 3, 1).reshape(-1, co * r * r)
        dx[:, :co * r * r] = _bf(g)

    return EmuLaunch("kair_pixel_shuffle_out_bwd", run)


def build_conv_chain(act0, act1, Wp, bias, slope, sync, *, L, conv):
    """kair_conv_chain = L runs of the conv-mode gemm_tn emulation over the ping-pong buffers (the slope is a device
    value in the real op, so it is read when the launch runs)"""
    acts = (act0, act1)

    def run():
        sl = slope.tolist()
        for l in range(L):
            build_gemm_tn(acts[l & 1], Wp[l], N=64, K=64, bn=64, conv=conv, bias=None if bias is None else bias[l],
                          act=ACT_LRELU, act_slope=sl[l], outB=acts[(l + 1) & 1]).run(None)

    return EmuLaunch("kair_conv_chain", run)


def build_pixel_loss(x, y, dx, loss, scratch, *, eps=1e-9, mode=0, max_pixel=0.0):
    def run():
        import math
        d = x - y
        if mode == 0:
            l = torch.sqrt(d * d + eps)
            g = d / l
        elif mode == 1:
            l = d.abs()
            g = torch.sign(d)
        elif mode == 5:
            q = lambda v: (v.clamp(0, 1) * 255.0).round()
            l = (q(x) - q(y)) ** 2
            g = torch.zeros_like(x)
        else:
            if mode == 2:
                den = lambda v: v * max_pixel
                dlam = torch.full_like(x, max_pixel)
            elif mode == 3:
                lg = math.log(max_pixel + 1.0)
                den = lambda v: torch.exp(v * lg) - 1.0
                dlam = lg * torch.exp(x * lg)
            else:
                s2 = max_pixel + 0.375
                den = lambda v: s2 * v * v - 0.375
                dlam = 2.0 * s2 * x
            raw = den(x)
            lam, k = raw.clamp(min=eps), den(y).clamp(min=0.0)
            l = lam - k * torch.log(lam)
            g = torch.where(raw >= eps, (1.0 - k / lam) * dlam, torch.zeros_like(x))
        loss[0] = l.double().mean().float()
        if dx is not None:
            dx.copy_(g / x.numel())

    return EmuLaunch("kair_pixel_loss", run)


def build_adam_ema(p, g, m, v, e, *, beta1, beta2, eps, weight_decay=0.0, ema_decay=0.0):
    import types
    args = types.SimpleNamespace(step_size=0.0, bc2_sqrt=1.0)

    def run():
        for i in range(len(p)):
            gi = g[i] + weight_decay * p[i] if weight_decay else g[i]
            m[i].lerp_(gi, 1.0 - beta1)
            v[i].mul_(beta2).addcmul_(gi, gi, value=1.0 - beta2)
            denom = (v[i].sqrt() / args.bc2_sqrt).add_(eps)
            p[i].addcdiv_(m[i], denom, value=-args.step_size)
            if e is not None:
                e[i].mul_(ema_decay).add_(p[i], alpha=1.0 - ema_decay)

    l = EmuLaunch("kair_adam_ema", run)
    l.args = args
    l.g = g
    return l


def adam_ema_set_grads(launch, g, tables):
    launch.g[:] = list(g)


# ------------------------------------------------------------------------------------------------ SPECT data path
SPECT_METHODS = {"log": 0, "anscombe": 1, "linear": 2, "none": 3}


def build_spect_patches(frames_L, frames_H, sample, out_L, out_H, *, patch, method, max_pixel):
    from oracle import spect_oracle as spo

    def run():
        for b, (f, top, left, mode) in enumerate(sample.tolist()):
            pl, ph = spo.train_patch(frames_L[f].numpy(), frames_H[f].numpy(), top, left, mode, patch, method, max_pixel)
            out_L[b].copy_(torch.from_numpy(pl))
            out_H[b].copy_(torch.from_numpy(ph))

    return EmuLaunch("kair_spect_patches", run)


def build_spect_normalize(x, y, *, method, max_pixel, inverse=False):
    from oracle import spect_oracle as spo

    def run():
        fn = spo.denormalize_spect if inverse else spo.normalize_spect
        y.copy_(torch.from_numpy(fn(x.numpy(), method, max_pixel).astype("float32")))

    return EmuLaunch("kair_spect_normalize", run)


def build_spect_metrics(E, H, out, *, max_val):
    from oracle import spect_oracle as spo
    import numpy as np

    def run():
        C, h, w = E.shape
        for c in range(C):
            e8, h8 = spo.quantize(E[c].numpy(), np.float32(max_val)), spo.quantize(H[c].numpy(), np.float32(max_val))
            out[c, 0] += float(((e8.astype(np.float64) - h8.astype(np.float64)) ** 2).sum())
            out[c, 1] += spo.ssim_uint8(e8, h8) * (h - 10) * (w - 10)
            out[c, 2] += spo.pll(E[c].numpy(), H[c].numpy()) * h * w

    return EmuLaunch("kair_spect_metrics", run)


# ------------------------------------------------------------------------------------------------ tile drivers
def _dihedral(img, mode):
    """utils_image.augment_img_tensor4 on (..., H, W)"""
    if mode == 0:
        return img
    k, flip = {1: (1, True), 2: (0, True), 3: (3, False), 4: (2, True), 5: (1, False), 6: (2, False), 7: (3, True)}[mode]
    out = img.rot90(k, [-2, -1]) if k else img
    return out.flip([-2]) if flip else out


def build_tiles_gather(img, origins, tiles):
    def run():
        n, C, th, tw = tiles.shape
        for i, (y, x, m) in enumerate(origins.tolist()):
            hs, ws = (tw, th) if m % 2 == 1 else (th, tw)
            tiles[i].copy_(_dihedral(img[:, y:y + hs, x:x + ws], m))

    return EmuLaunch("kair_tiles_gather", run)


def build_tiles_accumulate(tiles, origins, E, Wsum=None, *, weight=1.0):
    def run():
        for i, (y, x, m) in enumerate(origins.tolist()):
            inv = 8 - m if m in (3, 5) else m
            back = _dihedral(tiles[i], inv)
            hb, wb = back.shape[-2:]
            E[:, y:y + hb, x:x + wb] += weight * back
            if Wsum is not None:
                Wsum[y:y + hb, x:x + wb] += weight

    return EmuLaunch("kair_tiles_accumulate", run)


def build_axpby(x, *, n, a=1.0, y=None, b=0.0, out_f32=None, out_bf16=None, mask=None, mask_slope=0.0, n_pad=None):
    def run():
        v = a * x[:, :n].float()
        if y is not None:
            v = v + b * y[:, :n].float()
        if out_f32 is not None:
            out_f32[:, :n].copy_(v)
        if out_bf16 is not None:
            vb = v
            if mask is not None:
                vb = torch.where(mask[:, :n].float() > 0, v, v * mask_slope)
            out_bf16[:, :n].copy_(_bf(vb))
            if n_pad is not None and n_pad > n:
                out_bf16[:, n:n_pad].zero_()

    return EmuLaunch("kair_axpby", run)


def build_ffd_pack(img, sigma, out):
    def run():
        Bn, C, h, w = img.shape
        x = F.pad(img, (0, (w + 1) // 2 * 2 - w, 0, (h + 1) // 2 * 2 - h), mode="replicate")
        u = F.pixel_unshuffle(x, 2)                                         # channel c*4 + dy*2 + dx
        m = sigma.view(Bn, 1, 1, 1).expand(Bn, 1, u.shape[2], u.shape[3])
        t = torch.cat((u, m), 1).permute(0, 2, 3, 1).reshape(-1, 4 * C + 1)
        out.zero_()
        out[:, :4 * C + 1].copy_(_bf(t))

    return EmuLaunch("kair_ffd_pack", run)
