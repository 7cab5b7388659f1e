"""Micro-benchmark: kair_mlp_fused vs the unfused pair of kair_gemm_tn launches at the cfg2 / cfg3 token count."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kair_b200 import ops  # noqa: E402


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    b.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def main():
    M, N, F, C, hid = 131072, 192, 384, 180, 360
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    X = torch.randn(M, N, device=dev, generator=g).bfloat16()
    W1 = (torch.randn(F, N, device=dev, generator=g) * 0.05).bfloat16()
    W2 = (torch.randn(N, F, device=dev, generator=g) * 0.05).bfloat16()
    b1, b2 = torch.zeros(F, device=dev), torch.zeros(N, device=dev)
    resid = torch.randn(M, N, device=dev, generator=g)
    bf = lambda *s: torch.zeros(*s, dtype=torch.bfloat16, device=dev)
    oF, oB, rs, u, gg = torch.zeros(M, N, device=dev), bf(M, N), torch.zeros(M, device=dev), bf(M, F), bf(M, F)
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device=dev)
    fused_inf = ops.build_mlp_fused(X, W1, W2, N=N, F=F, bias1=b1, bias2=b2, resid=resid, outF=oF, outB=oB, rowstat=rs,
                                    ln=True, n_valid=C, ones_col=C, ones_col_g=hid)
    fused_trn = ops.build_mlp_fused(X, W1, W2, N=N, F=F, bias1=b1, bias2=b2, resid=resid, outF=oF, outB=oB, rowstat=rs,
                                    outU=u, outG=gg, ln=True, n_valid=C, ones_col=C, ones_col_g=hid)
    fc1 = ops.build_gemm_tn(X, W1, N=F, K=N, bn=192, mode=ops.EPI_GELU, bias=b1, outB=gg, outB2=u, ones_col=hid)
    fc2 = ops.build_gemm_tn(gg, W2, N=N, K=F, bn=192, mode=ops.EPI_RESID_LN, bias=b2, resid=resid, outF=oF, outB=oB,
                            rowstat=rs, n_valid=C, ones_col=C)
    st = ops.stream_handle()
    print("fused inference  %.1f us" % timeit(lambda: fused_inf.run(st)))
    print("fused training   %.1f us" % timeit(lambda: fused_trn.run(st)))
    print("unfused pair     %.1f us" % timeit(lambda: (fc1.run(st), fc2.run(st))))


if __name__ == "__main__":
    main()
