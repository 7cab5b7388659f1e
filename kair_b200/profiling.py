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
        tag = "gemm_tn[%s N%d K%d %s]" % ("conv3x3" if a.conv else "dense", a.N, a.K * taps, _EPI[a.mode])
        return tag, 2.0 * a.M * r(a.N) * r(a.K) * taps, 2.0 * a.M * (r(a.K) + r(a.N))
    if name == "gemm_nt":
        taps = 9 if a.conv else 1
        tag = "gemm_nt[%s n1=%d n2=%d]" % ("conv3x3" if a.conv else "dense", a.n1, a.n2 * taps)
        return tag, 2.0 * a.T * r(a.n1) * r(a.n2) * taps, 2.0 * a.T * (r(a.n1) + r(a.n2))
    if name in ("window_attn_fwd", "window_attn_bwd"):
        T = a.batch * a.H * a.W
        N = a.window * a.window
        hd = real_dims.get("head_dim", 30)
        f = 4.0 * (T // N) * a.num_heads * N * N * hd          # QK^T + PV
        by = 4.0 * T * a.num_heads * hd * 2                    # q, k, v, o in bf16
        if name.endswith("bwd"):
            return "window_attn_bwd", 2.0 * f, 2.0 * by        # dV, dP, dQ, dK; + dO, dq, dk, dv traffic
        return "window_attn_fwd", f, by
    return name, 0.0, 0.0


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
        d = out.setdefault(tag, dict(count=0, ms=0.0, flops=0.0, bytes=0.0))
        d["count"] += 1
        d["ms"] += t
        d["flops"] += fl
        d["bytes"] += by
    return out
