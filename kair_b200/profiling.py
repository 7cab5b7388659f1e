"""Per-launch device timing and algorithmic-work accounting for launch lists (used by bench.py's roofline leg).

Algorithmic FLOPs follow SURVEY.md section 8(d): 2*MAC on the *unpadded* dimensions of the reference arithmetic
(C=180, head_dim=30, hidden=360 ...), so padding work (180 -> 192, 30 -> 32) counts against the achieved number.
"""
import torch

_EPI = ["linear", "gelu", "resid_ln", "dgelu", "lnbwd"]


def describe(launch, real_dims):
    """(tag, algorithmic FLOPs, algorithmic HBM bytes) of one recorded launch.

    real_dims maps padded GEMM extents to the reference's extents, e.g. {192: 180, 576: 540, 384: 360}."""
    a = getattr(launch, "args", None)
    name = launch.name.replace("kair_", "")
    r = lambda n: real_dims.get(int(n), int(n))
    if name == "gemm_tn":
        taps = 9 if a.conv else 1
        terms = (getattr(launch, "meta", None) or {}).get("terms", 1)   # fp32-accurate mode: split operands along K
        tag = "gemm_tn[%s N%d K%d %s]" % ("conv3x3" if a.conv else "dense", a.N, a.K * taps, _EPI[a.mode])
        if terms > 1:
            tag = tag[:-1] + " split%d]" % terms
        if a.mode == 0 and a.aux:
            tag = tag[:-1] + "+actbwd]"
        # operand traffic of this launch on the unpadded extents: A once (the 9 taps re-read it from L2 / smem),
        # fp32 residual in, bf16 aux in, every output once; weights are L2-resident.  Split operands: the algorithmic
        # FLOPs stay those of the reference's fp32 product (K / terms); the bytes are what the split layout moves.
        kr = r(a.K // terms)
        by = 2.0 * a.M * kr * terms
        by += 4.0 * a.M * r(a.N) * (1 if a.resid else 0) + 2.0 * a.M * r(a.N) * (1 if a.aux else 0)
        by += 2.0 * a.M * r(a.N) * ((1 if a.outB else 0) + (1 if a.outB2 else 0)) + 4.0 * a.M * r(a.N) * (1 if a.outF else 0)
        return tag, 2.0 * a.M * r(a.N) * kr * taps, by
    if name == "conv_chain":
        M = float(a.batch) * a.H * a.W
        # per layer: the activation is read and written once (bf16, 64 channels); FLOPs of L 3x3 convolutions 64 -> 64
        return "conv_chain[L%d conv3x3 N64 K576]" % a.L, 2.0 * M * 64 * 576 * a.L, 2.0 * M * 64 * 2 * a.L
    if name == "mlp_fused":
        tag = "mlp_fused[N%d F%d %s%s]" % (a.N, a.F, "resid_ln" if a.ln else "linear", " +u,g" if a.outU else "")
        rn, rf = r(a.N), r(a.F)
        by = 2.0 * a.M * rn + 4.0 * a.M * rn * ((1 if a.resid else 0) + (1 if a.outF else 0))
        by += 2.0 * a.M * rn * (1 if a.outB else 0) + 2.0 * a.M * rf * ((1 if a.outU else 0) + (1 if a.outG else 0))
        return tag, 4.0 * a.M * rn * rf, by
    if name == "gemm_nt":
        taps = 9 if a.conv else 1
        tag = "gemm_nt[%s n1=%d n2=%d%s]" % ("conv3x3" if a.conv else "dense", a.n1, a.n2 * taps, " swap" if a.swap else "")
        return tag, 2.0 * a.T * r(a.n1) * r(a.n2) * taps, 2.0 * a.T * (r(a.n1) + r(a.n2))
    if name in ("window_attn_fwd", "window_attn_bwd"):
        T = a.batch * a.H * a.W
        N = a.window * a.window
        hd = real_dims.get("head_dim", 30)
        esz = 4 if a.fp32 else 2
        sfx = "[fp32 cuda-core]" if a.fp32 else ""
        f = 4.0 * (T // N) * a.num_heads * N * N * hd          # QK^T + PV
        by = 4.0 * T * a.num_heads * hd * esz                  # q, k, v, o in bf16 (15 360 B / window-head)
        if name.endswith("bwd"):
            # autograd's dV, dP, dQ, dK contractions (the kernel's recomputation of S is not algorithmic work);
            # reads q, k, v, dO, writes dq, dk, dv
            return "window_attn_bwd" + sfx, 2.0 * f, 7.0 * T * a.num_heads * hd * esz
        return "window_attn_fwd" + sfx, f, by
    if name == "split_terms":
        return name, 0.0, float(a.T) * a.n * (4.0 + (4.0 if a.aux else 0) + 2.0 * a.terms + (4.0 if a.out_f32 else 0))
    if name == "ln_bwd_split":
        return name, 0.0, float(a.T) * a.C * (4.0 + 2.0 * 2 + (4.0 if a.resid else 0) + 4.0 + (2.0 * a.terms if a.out else 0))
    if name == "layernorm_fwd":
        tm = a.terms if a.terms > 1 else 1
        return name, 0.0, float(a.T) * a.C * (4.0 + (4.0 if a.y_f32 else 0) + 2.0 * tm * ((1 if a.y_bf16 else 0) + (1 if a.xhat_bf16 else 0)))
    T_ = lambda: int(a.T)
    if name in ("bn_stats",):
        return name, 0.0, 4.0 * T_() * a.C
    if name == "bn_apply":
        return name, 0.0, T_() * a.C * (4.0 + (2.0 if a.out_bf16 else 0) + (4.0 if a.out_f32 else 0))
    if name == "bn_bwd_reduce":
        return name, 0.0, T_() * a.C * (4.0 + 2.0 + (4.0 if a.dy_f32 else 2.0))
    if name == "bn_bwd_apply":
        return name, 0.0, T_() * a.C * (4.0 + 2.0 + (4.0 if a.dy_f32 else 2.0) + 2.0)
    if name == "colsum":
        return name, 0.0, T_() * a.C * (2.0 if a.x_bf16 else 4.0)
    if name in ("space_to_depth", "depth_to_space"):
        n = float(a.batch) * a.H * a.W * a.C
        if name == "space_to_depth":
            return name, 0.0, 4.0 * n
        return name, 0.0, n * ((4.0 if a.coarse_in_f32 else 2.0) + (4.0 if a.fine_out_f32 else 0) + (2.0 if a.fine_out_bf16 else 0))
    if name in ("pixel_shuffle", "upsample_nearest2"):
        rr = a.r * a.r if name == "pixel_shuffle" else 4
        n = float(a.batch) * a.H * a.W * a.C * rr
        return name, 0.0, 2.0 * n + (2.0 * n if name == "pixel_shuffle" else 2.0 * n / 4)
    if name in ("conv_in_fwd", "conv_out_fwd", "conv_out_dgrad", "conv_out_wgrad", "conv_in_wgrad"):
        Tn = float(a.batch) * a.H * a.Wd
        if name.startswith("conv_in"):
            ch, img_c = a.O, a.ci
        else:
            ch, img_c = a.C, a.co
        fl = 2.0 * Tn * 9 * ch * img_c
        feat_b = Tn * ch * 2.0
        if name == "conv_in_fwd":
            feat_b = Tn * ch * ((4.0 if a.out_f32 else 0) + (2.0 if a.out_bf16 else 0))
        elif name == "conv_in_wgrad":
            feat_b = Tn * ch * ((4.0 if a.dy_f32 else 0) + (2.0 if a.dy_bf16 else 0))
        elif name == "conv_out_dgrad":
            feat_b = Tn * ch * ((4.0 if a.dfeat_f32 else 0) + (2.0 if a.dfeat_bf16 else 0) + (2.0 if a.mask else 0))
        return name, fl, feat_b + Tn * img_c * 4.0
    if name == "add_inplace":
        return name, 0.0, 0.0
    return name, 0.0, 0.0


def write_bytes(launch, real_dims):
    """Algorithmic HBM bytes WRITTEN by one launch (subset of describe()'s bytes) for the kernels that dominate a step.
    B200 HBM3e sustains ~6.5 TB/s for mixed traffic but only ~3.9 TB/s for pure writes (measured with a fill kernel,
    bench.py `hbm_write_peak`), so write-heavy epilogues (qkv, fc1+GELU: 75-80 % writes) have a lower ceiling."""
    a = getattr(launch, "args", None)
    name = launch.name.replace("kair_", "")
    r = lambda n: real_dims.get(int(n), int(n))
    if name == "mlp_fused":
        return (4.0 * a.M * r(a.N) * (1 if a.outF else 0) + 2.0 * a.M * r(a.N) * (1 if a.outB else 0) +
                2.0 * a.M * r(a.F) * ((1 if a.outU else 0) + (1 if a.outG else 0)))
    if name == "gemm_tn":
        return 2.0 * a.M * r(a.N) * ((1 if a.outB else 0) + (1 if a.outB2 else 0)) + 4.0 * a.M * r(a.N) * (1 if a.outF else 0)
    if name in ("window_attn_fwd", "window_attn_bwd"):
        T = a.batch * a.H * a.W
        hd = real_dims.get("head_dim", 30)
        return (3.0 if name.endswith("bwd") else 1.0) * T * a.num_heads * hd * 2
    return 0.0


def time_launches(launches, stream_handle, reps=1):
    """Run `launches` in order `reps` times on the current stream with a CUDA event between every pair of launches;
    returns per-launch mean milliseconds (list aligned with `launches`)."""
    n = len(launches)
    tot = [0.0] * n
    for _ in range(reps):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
        ev[0].record()
        for i, l in enumerate(launches):
            l.run(stream_handle)
            ev[i + 1].record()
        ev[n].synchronize()
        for i in range(n):
            tot[i] += ev[i].elapsed_time(ev[i + 1])
    return [t / reps for t in tot]


def summarize(launches, ms, real_dims):
    """Group per-launch times by tag -> {tag: dict(count, ms, flops, bytes)} for one pass over the list."""
    out = {}
    for l, t in zip(launches, ms):
        tag, fl, by = describe(l, real_dims)
        d = out.setdefault(tag, dict(count=0, ms=0.0, flops=0.0, bytes=0.0, wbytes=0.0))
        d["count"] += 1
        d["ms"] += t
        d["flops"] += fl
        d["bytes"] += by
        d["wbytes"] += write_bytes(l, real_dims)
    return out


def measure_fill_gbs(device, mib=512, reps=5):
    """Pure-write HBM bandwidth of this GPU (GB/s): best of `reps` fills of a `mib` MiB buffer (calibration only)."""
    x = torch.empty(mib << 20, dtype=torch.uint8, device=device)
    x.zero_()
    torch.cuda.synchronize(device)
    best = float("inf")
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        x.zero_()
        b.record()
        b.synchronize()
        best = min(best, a.elapsed_time(b))
    return (mib << 20) / (best * 1e-3) / 1e9


def roofline_of(group, peak_tflops, peak_gbs):
    """roofline object fields for one kernel group dict(count, ms, flops, bytes): the bound is chosen by arithmetic
    intensity against the ridge peak_tflops / peak_gbs."""
    sec = group["ms"] * 1e-3
    ai = group["flops"] / group["bytes"] if group["bytes"] > 0 else float("inf")
    ridge = peak_tflops * 1e12 / (peak_gbs * 1e9)
    if group["flops"] > 0 and ai >= ridge:
        ach = group["flops"] / sec / 1e12 if sec > 0 else 0.0
        return {"bound": "tensor", "achieved": ach, "peak": peak_tflops, "unit": "TFLOP/s", "frac": ach / peak_tflops}
    ach = group["bytes"] / sec / 1e9 if sec > 0 else 0.0
    return {"bound": "hbm", "achieved": ach, "peak": peak_gbs, "unit": "GB/s", "frac": ach / peak_gbs}
