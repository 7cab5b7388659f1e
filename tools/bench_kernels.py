"""Isolated CUDA-event timings of the hot kernels at cfg2 shapes, with ablations (drop residual / outputs) that show
which part of an epilogue bounds it.  Development aid: python tools/bench_kernels.py [filter]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kair_b200 import ops

T = 131072
dev = "cuda"
bf = lambda *s: (torch.randn(*s, device=dev) * 0.5).bfloat16()
f32 = lambda *s: torch.randn(*s, device=dev)


def timeit(fn, n=20):
    """GPU time per call: the n calls are captured in a CUDA graph so host launch cost is excluded."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if os.environ.get("KB_NO_GRAPH") != "1":
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(n):
                fn()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    flt = sys.argv[1] if len(sys.argv) > 1 else ""
    A192, A384, A576 = bf(T, 192), bf(T, 384), bf(T, 576)
    W = {(n, k): bf(n, k) * 0.1 for n in (192, 384, 576) for k in (192, 384, 576)}
    bias = {n: f32(n) for n in (192, 384, 576)}
    R, OF, OB, OB2, OB384, OB576 = f32(T, 192), f32(T, 192), bf(T, 192), bf(T, 384), bf(T, 384), bf(T, 576)
    AUX192, AUX384 = bf(T, 192), bf(T, 384)
    rs = torch.rand(T, device=dev) + 0.5
    cases = {}
    def add(name, **kw):
        cases[name] = kw
    add("qkv linear N576 K192", A=A192, B=W[(576, 192)], N=576, K=192, bias=bias[576], outB=OB576)
    add("qkv linear N576 K192 no-out", A=A192, B=W[(576, 192)], N=576, K=192, bias=bias[576])
    add("qkv linear N576 K192 no-bias", A=A192, B=W[(576, 192)], N=576, K=192, outB=OB576)
    add("linear N192 K192 no-out", A=A192, B=W[(192, 192)], N=192, K=192)
    add("linear N192 K576 no-out", A=A576, B=W[(192, 576)], N=192, K=576)
    add("resid_ln N192 K192 full", A=A192, B=W[(192, 192)], N=192, K=192, mode=ops.EPI_RESID_LN, bias=bias[192], resid=R, outF=OF, outB=OB, rowstat=rs, n_valid=180, ones_col=180)
    add("resid_ln N192 K192 no-resid", A=A192, B=W[(192, 192)], N=192, K=192, mode=ops.EPI_RESID_LN, bias=bias[192], outF=OF, outB=OB, rowstat=rs, n_valid=180, ones_col=180)
    add("resid_ln N192 K192 no-outF", A=A192, B=W[(192, 192)], N=192, K=192, mode=ops.EPI_RESID_LN, bias=bias[192], resid=R, outB=OB, rowstat=rs, n_valid=180, ones_col=180)
    add("resid_ln N192 K192 no-resid no-outF", A=A192, B=W[(192, 192)], N=192, K=192, mode=ops.EPI_RESID_LN, bias=bias[192], outB=OB, rowstat=rs, n_valid=180, ones_col=180)
    add("resid_ln N192 K384 full", A=A384, B=W[(192, 384)], N=192, K=384, mode=ops.EPI_RESID_LN, bias=bias[192], resid=R, outF=OF, outB=OB, rowstat=rs, n_valid=180, ones_col=180)
    add("gelu N384 K192", A=A192, B=W[(384, 192)], N=384, K=192, mode=ops.EPI_GELU, bias=bias[384], outB=OB384, outB2=OB2, ones_col=360)
    add("gelu N384 K192 no-u", A=A192, B=W[(384, 192)], N=384, K=192, mode=ops.EPI_GELU, bias=bias[384], outB=OB384, ones_col=360)
    add("dgelu N384 K192", A=A192, B=W[(384, 192)], N=384, K=192, mode=ops.EPI_DGELU, aux=AUX384, outB=OB384)
    add("lnbwd N192 K384 full", A=A384, B=W[(192, 384)], N=192, K=384, mode=ops.EPI_LNBWD, aux=AUX192, rowstat=rs, resid=R, outF=OF, outB=OB, n_valid=180)
    add("lnbwd N192 K384 no-resid", A=A384, B=W[(192, 384)], N=192, K=384, mode=ops.EPI_LNBWD, aux=AUX192, rowstat=rs, outF=OF, outB=OB, n_valid=180)
    add("lnbwd N192 K384 no-outF", A=A384, B=W[(192, 384)], N=192, K=384, mode=ops.EPI_LNBWD, aux=AUX192, rowstat=rs, resid=R, outB=OB, n_valid=180)
    add("lnbwd N192 K576 full", A=A576, B=W[(192, 576)], N=192, K=576, mode=ops.EPI_LNBWD, aux=AUX192, rowstat=rs, resid=R, outF=OF, outB=OB, n_valid=180)
    add("linear N192 K192 (projT)", A=A192, B=W[(192, 192)], N=192, K=192, outB=OB)
    for name, kw in cases.items():
        if flt and flt not in name:
            continue
        kw = dict(kw)
        A, B = kw.pop("A"), kw.pop("B")
        l = ops.build_gemm_tn(A, B, bn=192, **kw)
        us = timeit(lambda: l.run(ops.stream_handle()))
        by = A.shape[0] * (kw["K"] * 2 + sum(v.element_size() * v.shape[-1] for k, v in kw.items() if isinstance(v, torch.Tensor) and v.dim() == 2))
        print("%-40s %8.1f us   %6.0f MB  %5.0f GB/s" % (name, us, by / 1e6, by / us / 1e3))
    # attention
    if not flt or "attn" in flt:
        B_, H, Wd, nH = 8, 128, 128, 6
        qkv = bf(T, 576); out = bf(T, 192); table = f32(225, 6) * 0.1
        dout = bf(T, 192); dqkv = bf(T, 576); dtable = torch.zeros(225, 6, device=dev)
        for shift in (0, 4):
            l = ops.build_window_attn_fwd(qkv, table, out, batch=B_, H=H, W=Wd, num_heads=nH, window=8, shift=shift)
            print("attn fwd shift %d  %8.1f us" % (shift, timeit(lambda: l.run(ops.stream_handle()))))
            l2 = ops.build_window_attn_bwd(qkv, table, dout, dqkv, dtable, batch=B_, H=H, W=Wd, num_heads=nH, window=8, shift=shift)
            print("attn bwd shift %d  %8.1f us" % (shift, timeit(lambda: l2.run(ops.stream_handle()))))


if __name__ == "__main__":
    main()
