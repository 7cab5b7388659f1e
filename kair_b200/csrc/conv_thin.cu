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
  int th;                 // tile height = warps per CTA
  int tiles_x, tiles_y;
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

// raw W[c][k][tap]; zero outside the real (co, C) extent
__device__ __forceinline__ float w_raw(const kair_conv_out_args& a, int c, int k, int tap) {
  return (c < a.co && k < a.C) ? __ldg(a.W + ((size_t)c * a.C + k) * 9 + tap) : 0.f;
}
// the bf16 term (0: hi, 1: lo) of an fp32 weight
__device__ __forceinline__ float w_term(float w, int lo_term) {
  const float hi = __bfloat162float(__float2bfloat16_rn(w));
  return lo_term ? w - hi : hi;
}

__global__ void __launch_bounds__(256) conv_out_fwd_mma_kernel(const ConvOutTile p) {
  extern __shared__ __align__(16) uint8_t smem[];
  const kair_conv_out_args& a = p.a;
  const int kpt = p.kc >> 4;                 // k-steps per tap
  const int ksteps = 9 * kpt;
  const uint32_t pitchB = 2u * p.kc + 16u;   // bytes per halo pixel
  uint2* frag = reinterpret_cast<uint2*>(smem);              // [ksteps][32 lanes]
  uint8_t* tile = smem + (size_t)ksteps * 32 * sizeof(uint2);
  const int tid = threadIdx.x, lane = tid & 31, wrow = tid >> 5, nthr = blockDim.x;
  const int tx = blockIdx.x % p.tiles_x, ty = (blockIdx.x / p.tiles_x) % p.tiles_y, b = blockIdx.x / (p.tiles_x * p.tiles_y);
  const int x0 = tx * CO_TW, y0 = ty * p.th;

  pdl_wait();                 // the features come from the previous kernel of the stream
  pdl_launch_dependents();

  {  // halo tile -> shared memory; (pixel, chunk) advance incrementally, one thread moves 16 bytes per step
    const bf16* feat = reinterpret_cast<const bf16*>(a.feat);
    const int cpp = p.kc >> 3;               // 16-byte chunks per pixel
    const int npix = (p.th + 2) * CO_HW;
    const int dq = nthr / cpp, dr = nthr - dq * cpp;
    const uint32_t tile_s = smem_u32(tile);
    int pix = tid / cpp, ch = tid - pix * cpp;
    while (pix < npix) {
      const int hy = pix / CO_HW, hx = pix - hy * CO_HW;
      const int y = y0 - 1 + hy, x = x0 - 1 + hx;
      const bool in = y >= 0 && y < a.H && x >= 0 && x < a.Wd;
      const bf16* src = in ? feat + (((size_t)b * a.H + y) * a.Wd + x) * a.pitch + ch * 8 : feat;
      cp_async16_zfill(tile_s + pix * pitchB + ch * 16, src, in);
      pix += dq; ch += dr;
      if (ch >= cpp) { ch -= cpp; ++pix; }
    }
  }
  // B fragments while the copies fly: lane (g = lane/4, t = lane%4) of k-step s holds W^T[k = 2t, 2t+1 | 2t+8, 2t+9][n = g];
  // four entries per thread per round so that the 16 weight gathers of a round are in flight together
  for (int e0 = tid; e0 < ksteps * 32; e0 += 4 * nthr) {
    float w[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = e0 + i * nthr;
      const int s = e >> 5, l = e & 31;
      const int tap = s / kpt, kk = s - tap * kpt;
      const int c = (l >> 2) & 3, k0 = kk * 16 + 2 * (l & 3);
      const bool on = e < ksteps * 32;
      w[i][0] = on ? w_raw(a, c, k0, tap) : 0.f;
      w[i][1] = on ? w_raw(a, c, k0 + 1, tap) : 0.f;
      w[i][2] = on ? w_raw(a, c, k0 + 8, tap) : 0.f;
      w[i][3] = on ? w_raw(a, c, k0 + 9, tap) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = e0 + i * nthr;
      const int lo = (e & 31) >> 4;          // n = lane/4 >= 4: the low-order term
      if (e < ksteps * 32) {
        uint2 v;
        v.x = pack_bf16x2(w_term(w[i][0], lo), w_term(w[i][1], lo));
        v.y = pack_bf16x2(w_term(w[i][2], lo), w_term(w[i][3], lo));
        frag[e] = v;
      }
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
  KB_CHECK(a.feat && a.W && a.out && a.out_mean, "kair_conv_out_fwd: null pointer");
  KB_CHECK(a.co >= 1 && a.co <= 4, "kair_conv_out_fwd: co=%d unsupported (max 4)", a.co);
  KB_CHECK(a.pitch % 8 == 0 && a.C <= a.pitch, "kair_conv_out_fwd: pitch must be a multiple of 8 channels, >= C");
  KB_CHECK(!a.skip || a.skip_mean, "kair_conv_out_fwd: skip needs skip_mean");
  ConvOutTile p;
  p.a = a;
  p.kc = (a.C + 15) & ~15;
  if (p.kc > a.pitch || p.kc > 192) return conv_out_fwd_simt_launch(a, st);   // rows too short / too wide for the tile
  p.th = p.kc <= 64 ? 8 : 4;
  p.tiles_x = (a.Wd + CO_TW - 1) / CO_TW;
  p.tiles_y = (a.H + p.th - 1) / p.th;
  const long long grid = (long long)p.tiles_x * p.tiles_y * a.batch;
  KB_CHECK(grid > 0 && grid < (1ll << 31), "kair_conv_out_fwd: empty or oversized image batch");
  const size_t smem = (size_t)9 * (p.kc >> 4) * 32 * sizeof(uint2) + (size_t)(p.th + 2) * CO_HW * (2 * p.kc + 16);
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(conv_out_fwd_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    attr_set[dev & 63] = true;
  }
  KB_CHECK(smem <= 112 * 1024, "kair_conv_out_fwd: tile does not fit in shared memory");
  KB_CUDA(launch_pdl(conv_out_fwd_mma_kernel, (int)grid, p.th * 32, smem, st, p));
  return 0;
}

}  // namespace kb
