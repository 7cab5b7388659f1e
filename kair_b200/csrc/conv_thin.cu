// kair_conv_out_fwd on the tensor cores: the last convolution of every network (network_dncnn.py:69 m_tail,
// network_unet.py:46 m_tail, network_swinir.py:762/:835 conv_last) maps C (60..192) feature channels to 1..4 image
// channels.  As a GEMM it is M = pixels, K = 9*C, N = co <= 4 -- far too thin for the tcgen05 tiles of gemm_tn, and the
// SIMT version (elementwise.cu, kept as the fallback for unusual shapes) spent ~1400 instructions per pixel and re-read
// every feature row nine times through L1/L2 (262 GB/s on the x4 SR heads, 7-8 % of an inference forward).  Here:
//   * a CTA owns a (TH x 32)-pixel tile of one image; the (TH+2) x 34 halo of bf16 feature rows is staged ONCE in shared
//     memory with 16-byte cp.async (zero-filled outside the image), pixel pitch = 2*KC + 16 bytes so that ldmatrix row
//     fetches are bank-conflict free;
//   * one warp per tile row: for each of the 9 taps x KC/16 k-steps, two ldmatrix.x4 fetch the A fragments of the row's
//     two 16-pixel groups straight from the shifted halo rows (im2col by address arithmetic) and feed
//     mma.sync.m16n8k16 (bf16 x bf16 -> fp32);
//   * the N = 8 columns carry the fp32 weights split in two bf16 terms: column c = hi(W[c]), column c + 4 = lo(W[c]),
//     out[c] = acc[c] + acc[c + 4] -- the weights keep ~16 mantissa bits, the features are bf16 as before;
//   * the B fragments (2 registers per lane per k-step) are precomputed per CTA into shared memory in fragment order.
// mma.sync (not tcgen05) on purpose: N = 8 uses 1/32 of a UMMA tile, the kernel is bound by the feature-tile fill.
#include <cuda_bf16.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

int conv_out_fwd_simt_launch(const kair_conv_out_args& a, cudaStream_t st);  // elementwise.cu

namespace {

constexpr int CO_TW = 32;          // tile width in pixels: two m16 groups per warp
constexpr int CO_HW = CO_TW + 2;   // halo width

struct ConvOutTile {
  kair_conv_out_args a;
  int kc;                 // channels per tap, rounded up to 16 (<= pitch)
  int th;                 // tile height = warps per CTA; grid = (ceil(W / 32), ceil(H / th), batch)
};

__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void* src, bool pred) {
  const int sz = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], const uint2 b) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b.x), "r"(b.y));
}

__device__ __forceinline__ float bf16_round(float w) { return __bfloat162float(__float2bfloat16_rn(w)); }

template <int KPT>   // k-steps per tap as a compile-time constant (4: C = 49..64, every SR head / DnCNN / DRUNet); 0: run-time
__global__ void __launch_bounds__(256) conv_out_fwd_mma_kernel(const ConvOutTile p) {
  extern __shared__ __align__(16) uint8_t smem[];
  const kair_conv_out_args& a = p.a;
  const int kpt = KPT ? KPT : (p.kc >> 4);   // k-steps per tap
  const int ksteps = 9 * kpt;
  const uint32_t pitchB = 32u * kpt + 16u;   // bytes per halo pixel (2 * kc + 16)
  uint2* frag = reinterpret_cast<uint2*>(smem);              // [ksteps][32 lanes]
  uint8_t* tile = smem + (size_t)ksteps * 32 * sizeof(uint2);
  const int tid = threadIdx.x, lane = tid & 31, wrow = tid >> 5, nwarp = blockDim.x >> 5;
  const int b = blockIdx.z;
  const int x0 = blockIdx.x * CO_TW, y0 = blockIdx.y * p.th;

  pdl_wait();                 // the features come from the previous kernel of the stream
  pdl_launch_dependents();

  {  // halo tile -> shared memory: a warp per halo row (row base and row validity are warp-uniform), lanes walk the
     // row's 34 * cpp 16-byte chunks; (pixel, chunk) advance incrementally
    const uint8_t* feat = reinterpret_cast<const uint8_t*>(a.feat);
    const int cpp = p.kc >> 3;               // 16-byte chunks per pixel
    const int dq = 32 / cpp, dr = 32 - dq * cpp;
    const uint32_t gpitch = (uint32_t)a.pitch * 2u;
    for (int hy = wrow; hy < p.th + 2; hy += nwarp) {
      const int y = y0 - 1 + hy;
      const bool yin = y >= 0 && y < a.H;
      const uint8_t* grow = feat + ((size_t)b * a.H + (yin ? y : 0)) * a.Wd * gpitch;
      const uint32_t srow = smem_u32(tile) + hy * CO_HW * pitchB;
      int hx = lane / cpp, ch = lane - hx * cpp;
      while (hx < CO_HW) {
        const int x = x0 - 1 + hx;
        const bool in = yin && x >= 0 && x < a.Wd;
        cp_async16_zfill(srow + hx * pitchB + ch * 16, in ? grow + (size_t)x * gpitch + ch * 16 : feat, in);
        hx += dq; ch += dr;
        if (ch >= cpp) { ch -= cpp; ++hx; }
      }
    }
  }
  // B fragments while the copies fly: lane (g = lane/4, t = lane%4) of k-step s holds W^T[k = 2t, 2t+1 | 2t+8, 2t+9][n = g],
  // n = c: hi(W[c]), n = c + 4: lo(W[c]).  Zero everything, then one thread per real (k-step, t, c) writes both terms
  // (disjoint addresses: no ordering needed between the two loops).
  {
    const int nthr = blockDim.x;
    for (int e = tid; e < ksteps * 32; e += nthr)
      if (((e >> 2) & 3) >= a.co) frag[e] = make_uint2(0u, 0u);
    for (int e = tid; e < ksteps * 4 * a.co; e += nthr) {
      const int t = e & 3, sc = e >> 2;
      const int s = sc / a.co, c = sc - s * a.co;
      const int tap = s / kpt, kk = s - tap * kpt;
      const int k0 = kk * 16 + 2 * t;
      const float* wp = a.W + ((size_t)c * a.C + k0) * 9 + tap;
      const float w0 = k0 < a.C ? __ldg(wp) : 0.f, w1 = k0 + 1 < a.C ? __ldg(wp + 9) : 0.f;
      const float w2 = k0 + 8 < a.C ? __ldg(wp + 72) : 0.f, w3 = k0 + 9 < a.C ? __ldg(wp + 81) : 0.f;
      const float h0 = bf16_round(w0), h1 = bf16_round(w1), h2 = bf16_round(w2), h3 = bf16_round(w3);
      frag[s * 32 + c * 4 + t] = make_uint2(pack_bf16x2(h0, h1), pack_bf16x2(h2, h3));
      frag[s * 32 + (c + 4) * 4 + t] = make_uint2(pack_bf16x2(w0 - h0, w1 - h1), pack_bf16x2(w2 - h2, w3 - h3));
    }
  }
  cp_async_wait_all();
  __syncthreads();

  const int y = y0 + wrow;
  if (y >= a.H) return;
  // C fragment: lane (g, t) holds rows g / g+8, columns 2t, 2t+1; hi + lo = columns c and c+4 = lanes t and t+2, so
  // lanes t < 2 finish channels 2t, 2t+1 of pixels g, g+8 of both 16-pixel groups.  Their epilogue operands are
  // fetched before the MMA loop.
  const int g = lane >> 2, t = lane & 3;
  float eb[2] = {0.f, 0.f}, em[2] = {0.f, 0.f}, es[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  if (t < 2) {
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int c = 2 * t + j;
      if (c < a.co) {
        eb[j] = a.b ? __ldg(a.b + c) : 0.f;
        em[j] = __ldg(a.out_mean + c);
        if (a.skip) {
          const float sm = __ldg(a.skip_mean + c);
          const size_t row = (((size_t)b * a.co + c) * a.H + y) * a.Wd;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int x = x0 + q * 8 + g;
            if (x < a.Wd) es[j][q] = a.skip_scale * (__ldg(a.skip + row + x) - sm);
          }
        }
      }
    }
  }
  float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  {
    // ldmatrix row address of this lane: matrix j = lane/8 -> pixel (lane%8) + 8*(j&1), k offset 8*(j>>1)
    const int r = (lane & 7) + ((lane >> 3) & 1) * 8, kofs = (lane >> 4) * 8;
    const uint32_t abase = smem_u32(tile) + (wrow * CO_HW + r) * pitchB + kofs * 2;
    const uint2* fl = frag + lane;
#pragma unroll
    for (int tap = 0; tap < 9; ++tap) {
      const uint32_t toff = abase + ((tap / 3) * CO_HW + tap % 3) * pitchB;
#pragma unroll
      for (int kk = 0; kk < kpt; ++kk) {
        const uint2 bw = fl[(tap * kpt + kk) * 32];
        uint32_t a0[4], a1[4];
        ldmatrix_x4(a0, toff + kk * 32);
        ldmatrix_x4(a1, toff + 16 * pitchB + kk * 32);
        mma_bf16_16816(acc[0], a0, bw);
        mma_bf16_16816(acc[1], a1, bw);
      }
    }
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[h][i] += __shfl_xor_sync(0xffffffffu, acc[h][i], 2);
    if (t < 2) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int j = i & 1, q = 2 * h + (i >> 1);
        const int c = 2 * t + j, x = x0 + q * 8 + g;
        if (c < a.co && x < a.Wd)
          a.out[(((size_t)b * a.co + c) * a.H + y) * a.Wd + x] = (acc[h][i] + eb[j] + es[j][q]) * a.out_scale + em[j];
      }
    }
  }
}

}  // namespace

int conv_out_fwd_launch(const kair_conv_out_args& a, cudaStream_t st) {
  if (a.feat_f32) return conv_out_fwd_simt_launch(a, st);   // fp32-accurate mode: CUDA-core fp32 kernel
  KB_CHECK(a.feat && a.W && a.out && a.out_mean, "kair_conv_out_fwd: null pointer");
  KB_CHECK(a.co >= 1 && a.co <= 4, "kair_conv_out_fwd: co=%d unsupported (max 4)", a.co);
  KB_CHECK(a.pitch % 8 == 0 && a.C <= a.pitch, "kair_conv_out_fwd: pitch must be a multiple of 8 channels, >= C");
  KB_CHECK(!a.skip || a.skip_mean, "kair_conv_out_fwd: skip needs skip_mean");
  ConvOutTile p;
  p.a = a;
  p.kc = (a.C + 15) & ~15;
  if (p.kc > a.pitch || p.kc > 192) return conv_out_fwd_simt_launch(a, st);   // rows too short / too wide for the tile
  p.th = p.kc <= 64 ? 8 : 4;
  const dim3 grid((unsigned)((a.Wd + CO_TW - 1) / CO_TW), (unsigned)((a.H + p.th - 1) / p.th), (unsigned)a.batch);
  KB_CHECK(a.batch > 0 && a.H > 0 && a.Wd > 0 && grid.y < 65536u && grid.z < 65536u,
           "kair_conv_out_fwd: empty or oversized image batch");
  const size_t smem = (size_t)9 * (p.kc >> 4) * 32 * sizeof(uint2) + (size_t)(p.th + 2) * CO_HW * (2 * p.kc + 16);
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(conv_out_fwd_mma_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    KB_CUDA(cudaFuncSetAttribute(conv_out_fwd_mma_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    attr_set[dev & 63] = true;
  }
  KB_CHECK(smem <= 112 * 1024, "kair_conv_out_fwd: tile does not fit in shared memory");
  if (p.kc == 64) KB_CUDA(launch_pdl(conv_out_fwd_mma_kernel<4>, grid, p.th * 32, smem, st, p));
  else KB_CUDA(launch_pdl(conv_out_fwd_mma_kernel<0>, grid, p.th * 32, smem, st, p));
  return 0;
}

}  // namespace kb
