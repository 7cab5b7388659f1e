// Shared device-side primitives for the sm_100a kernels: mbarrier, TMA (cp.async.bulk.tensor),
// tcgen05 (alloc / mma / commit / ld / st) and UMMA descriptor builders.
// Everything here is inline PTX; no CUTLASS/CuTe dependency.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

namespace kb {

typedef __nv_bfloat16 bf16;

#ifndef KB_SPIN_TIMEOUT_CYCLES
// ~2.5 s at 1.9 GHz: a barrier that never flips traps the kernel instead of hanging the GPU.
#define KB_SPIN_TIMEOUT_CYCLES 5000000000ll
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
// One elected thread of a CONVERGED warp.  Single-thread issue of TMA / tcgen05 instructions must be guarded by this,
// not by `lane == 0`: their operands live in uniform registers, and only under elect.sync does the compiler know that
// one thread is active -- otherwise every such instruction is wrapped in an ELECT / R2UR / BRA.U.ANY waterfall loop
// (measured: ~95 cycles per tcgen05.mma, 130-300 per TMA operation).  The leader is the same thread on every call
// (same member mask), which the per-thread bulk-group and tcgen05.commit bookkeeping rely on.
__device__ __forceinline__ bool elect_one() {
  __syncwarp();
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ------------------------------------------------------------------ mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > KB_SPIN_TIMEOUT_CYCLES) {
      printf("kair_b200: mbarrier timeout block %d thread %d bar %u parity %u\n", (int)blockIdx.x,
             (int)threadIdx.x, smem_u32(bar), parity);
      __trap();
    }
  }
}

// ------------------------------------------------------------------ proxies / fences
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ------------------------------------------------------------------ TMA
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tm),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tm, const void* smem_src, int c0, int c1, int c2,
                                             int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(tm),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* tm, const void* smem_src, int c0, int c1, int c2,
                                             int c3, int c4) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(tm),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
               : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* tm, const void* smem_src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tm),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ------------------------------------------------------------------ tcgen05
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], kind::f16 (bf16/fp16 inputs, fp32 accumulate)
__device__ __forceinline__ void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// make all previously issued tcgen05.mma of this thread arrive on an mbarrier when they complete
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// 32 lanes x 32 columns of 32-bit: thread i of the warp receives lane (base_lane+i), columns [c, c+32)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,"
      "%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,"
      "%30,%31,%32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]),
      "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
      "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}

// ------------------------------------------------------------------ UMMA descriptors
// Shared-memory matrix descriptor (sm_100 "version 1"):
//   [0,14)  start address >> 4      [16,30) leading byte offset >> 4
//   [32,46) stride byte offset >> 4 [46,48) version = 1   [49,52) base offset   [61,64) layout type
// layout type: 0 none, 1 128B(base32B), 2 SWIZZLE_128B, 4 SWIZZLE_64B, 6 SWIZZLE_32B
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(layout_type & 7) << 61;
  return d;
}
// K-major operand tile [rows][64 bf16] written by TMA with SWIZZLE_128B: 8-row atoms of 1024 B.
__device__ __forceinline__ uint64_t make_desc_kmajor_sw128(uint32_t smem_addr) {
  return make_smem_desc(smem_addr, 0, 1024, 2);
}
// MN-major operand: atoms of [8 k-rows][64 mn elements] (1024 B); atoms along K every 1024 B (SBO),
// 64-wide MN atoms every `k_rows*128` bytes (LBO).
__device__ __forceinline__ uint64_t make_desc_mnmajor_sw128(uint32_t smem_addr, uint32_t k_rows) {
  return make_smem_desc(smem_addr, k_rows * 128, 1024, 2);
}
// Instruction descriptor for kind::f16, bf16 x bf16 -> fp32.
//   [4,6) c_format (1 = f32)  [7,10) a_format (1 = bf16)  [10,13) b_format  [15] a_major  [16] b_major (1 = MN)
//   [17,23) N >> 3            [24,29) M >> 4
__host__ __device__ constexpr uint32_t make_idesc_bf16(uint32_t M, uint32_t N, uint32_t a_mn_major,
                                                       uint32_t b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn_major << 15) | (b_mn_major << 16) | ((N >> 3) << 17) |
         ((M >> 4) << 24);
}

// ------------------------------------------------------------------ misc math
// programmatic dependent launch (see launch_pdl in host.h): wait for the previous kernel of the stream to complete and
// flush, then let the next one start its own prologue as soon as SM resources free up
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
// Exact-form GELU (nn.GELU(approximate='none'), network_swinir.py:21): x * Phi(x), Phi(x) = 0.5 * erfc(-x / sqrt2).
// erfc via Abramowitz-Stegun 7.1.26 (|abs err| <= 1.5e-7, far below the bf16 rounding of the result):
//   h = 0.5 * erfc(|x|/sqrt2) = t * poly(t) * exp(-x^2/2),  t = 1 / (1 + 0.3275911 |x| / sqrt2),  Phi = x < 0 ? h : 1 - h
// with the 0.5 folded into the polynomial and the two transcendental steps as single MUFU instructions
// (ex2.approx / rcp.approx; no range fix-ups are needed: the arguments are in (-inf, 0] and [1, inf)).  16 SASS
// instructions per element instead of 24 -- the GELU / GELU' epilogues were instruction-issue bound (ncu: 68 % issue
// slots at 32 % of DRAM peak).  e = exp(-x^2/2) is shared with the Gaussian density of gelu'.
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Packed fp32x2 arithmetic (sm_100: FFMA2 / FMUL2 / FADD2, one issue slot for two lanes of fp32 math).  The softmax and
// epilogue loops of this library are instruction-issue bound (two epilogue warps per scheduler), so halving the FP
// instruction count is worth more there than any memory trick.  Operands are passed as scalars and packed through
// mov.b64, which ptxas folds away when the pair sits in an aligned register pair (arrays unrolled in steps of two do).
__device__ __forceinline__ uint64_t f2_pack(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& a, float& b) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
}
// (d0, d1) = (a0, a1) * (b0, b1) + (c0, c1)
__device__ __forceinline__ void fma2(float& d0, float& d1, float a0, float a1, float b0, float b1, float c0, float c1) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(f2_pack(a0, a1)), "l"(f2_pack(b0, b1)), "l"(f2_pack(c0, c1)));
  f2_unpack(r, d0, d1);
}
__device__ __forceinline__ void mul2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_pack(a0, a1)), "l"(f2_pack(b0, b1)));
  f2_unpack(r, d0, d1);
}
__device__ __forceinline__ void add2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(f2_pack(a0, a1)), "l"(f2_pack(b0, b1)));
  f2_unpack(r, d0, d1);
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float gauss_cdf_from(float x, float e) {
  const float t = rcp_approx(fmaf(0.23164189f, fabsf(x), 1.0f));
  float pl = fmaf(0.5f * 1.061405429f, t, -0.5f * 1.453152027f);
  pl = fmaf(pl, t, 0.5f * 1.421413741f);
  pl = fmaf(pl, t, -0.5f * 0.284496736f);
  pl = fmaf(pl, t, 0.5f * 0.254829592f);
  const float h = pl * t * e;
  return x < 0.f ? h : 1.0f - h;
}
__device__ __forceinline__ float gelu_erf(float x) {
  const float e = ex2_approx(-0.72134752044448f * x * x);  // exp(-x^2/2)
  return x * gauss_cdf_from(x, e);
}
__device__ __forceinline__ float gelu_erf_grad(float x) {
  // d/dx [x Phi(x)] = Phi(x) + x * exp(-x^2/2) / sqrt(2 pi)
  const float e = ex2_approx(-0.72134752044448f * x * x);
  return fmaf(x * 0.3989422804014327f, e, gauss_cdf_from(x, e));
}
// Two-lane versions for the issue-bound GEMM epilogues (FFMA2 / FMUL2): ~10.5 issue slots per element instead of 16.
// Phi is assembled as 0.5 + copysign(0.5 - h, x), which equals the select above up to one fp32 rounding of 0.5 - h.
__device__ __forceinline__ void gauss_cdf_half2(float x0, float x1, float e0, float e1, float& s0, float& s1) {
  float d0, d1;
  fma2(d0, d1, fabsf(x0), fabsf(x1), 0.23164189f, 0.23164189f, 1.0f, 1.0f);
  const float t0 = rcp_approx(d0), t1 = rcp_approx(d1);
  float a0, a1;
  fma2(a0, a1, t0, t1, 0.5f * 1.061405429f, 0.5f * 1.061405429f, -0.5f * 1.453152027f, -0.5f * 1.453152027f);
  fma2(a0, a1, a0, a1, t0, t1, 0.5f * 1.421413741f, 0.5f * 1.421413741f);
  fma2(a0, a1, a0, a1, t0, t1, -0.5f * 0.284496736f, -0.5f * 0.284496736f);
  fma2(a0, a1, a0, a1, t0, t1, 0.5f * 0.254829592f, 0.5f * 0.254829592f);
  mul2(a0, a1, a0, a1, t0, t1);
  fma2(a0, a1, a0, a1, -e0, -e1, 0.5f, 0.5f);   // 0.5 - h
  s0 = copysignf(a0, x0);                       // Phi(x) - 0.5
  s1 = copysignf(a1, x1);
}
__device__ __forceinline__ void gelu_erf2(float& x0, float& x1) {
  float q0, q1;
  mul2(q0, q1, x0, x1, x0, x1);
  mul2(q0, q1, q0, q1, -0.72134752044448f, -0.72134752044448f);
  const float e0 = ex2_approx(q0), e1 = ex2_approx(q1);
  float s0, s1, h0, h1;
  gauss_cdf_half2(x0, x1, e0, e1, s0, s1);
  mul2(h0, h1, x0, x1, 0.5f, 0.5f);
  fma2(x0, x1, x0, x1, s0, s1, h0, h1);          // x * (0.5 + s)
}
// (g0, g1) = gelu'(x0), gelu'(x1)
__device__ __forceinline__ void gelu_erf_grad2(float x0, float x1, float& g0, float& g1) {
  float q0, q1;
  mul2(q0, q1, x0, x1, x0, x1);
  mul2(q0, q1, q0, q1, -0.72134752044448f, -0.72134752044448f);
  const float e0 = ex2_approx(q0), e1 = ex2_approx(q1);
  float s0, s1, w0, w1;
  gauss_cdf_half2(x0, x1, e0, e1, s0, s1);
  add2(s0, s1, s0, s1, 0.5f, 0.5f);
  mul2(w0, w1, x0, x1, 0.3989422804014327f, 0.3989422804014327f);
  fma2(g0, g1, w0, w1, e0, e1, s0, s1);
}
// SPLIT PRECISION (include/kair_b200.h): bf16 components of an fp32 value and the segment orders of the two operand
// sides.  Component products c_i * C_j with i + j <= 1 (terms 3) / <= 2 (terms 6) are kept.
__device__ __forceinline__ void split_components(float x, float& c0, float& c1, float& c2) {
  c0 = __bfloat162float(__float2bfloat16(x));
  const float r1 = x - c0;
  c1 = __bfloat162float(__float2bfloat16(r1));
  c2 = __bfloat162float(__float2bfloat16(r1 - c1));
}
__host__ __device__ constexpr int split_seg_act(int terms, int s) {
  return terms == 3 ? (s == 1 ? 1 : 0) : (terms == 6 ? (s < 3 ? s : (s < 5 ? s - 3 : 0)) : 0);
}
__host__ __device__ constexpr int split_seg_w(int terms, int s) {
  return terms == 3 ? (s == 2 ? 1 : 0) : (terms == 6 ? (s < 3 ? 0 : (s < 5 ? 1 : 2)) : 0);
}
__device__ __forceinline__ float split_pick(float x, int which) {
  float c0, c1, c2;
  split_components(x, c0, c1, c2);
  return which == 0 ? c0 : (which == 1 ? c1 : c2);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace kb
