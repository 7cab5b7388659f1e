// Memory-bound kernels of the conv-net paths (DnCNN, DRUNet / UNetRes) and of SwinIR's reconstruction heads:
//   * 2x2 stride-2 pixel regrouping (space-to-depth / depth-to-space) that turns the strided convolution and the
//     transposed convolution into dense tcgen05 GEMMs (models/basicblock.py:471-477, 495-501)
//   * per-channel column sums (bias gradients)
//   * training-mode BatchNorm2d on NHWC token rows: statistics, apply (+ReLU), backward reduce / apply, and the
//     eval-mode fold into the preceding convolution (models/basicblock.py:68-69; utils/utils_bnorm.py:44-58)
//   * PixelShuffle / nearest-neighbour x2 of token features (models/network_swinir.py:572-615, 823-826)
// All of them are HBM-bound: 16-byte vector accesses, consecutive threads on consecutive addresses, grids sized in
// multiples of the SM count, warp/block partial sums before any atomic.
#include <string.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

static inline int grid_for(long long items, int per_block, int max_mult) {
  long long b = (items + per_block - 1) / per_block;
  long long cap = (long long)sm_count() * max_mult;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// ------------------------------------------------------------------------------------------------ regrouping
__global__ void s2d_kernel(const kair_regroup_args a) {
  const int Hc = a.H >> 1, Wc = a.W >> 1, vc = a.C >> 3;  // 8 bf16 per vector
  const long long nvec = (long long)a.batch * Hc * Wc * 4 * vc;
  const uint4* src = reinterpret_cast<const uint4*>(a.fine_bf16);
  uint4* dst = reinterpret_cast<uint4*>(a.coarse_bf16);
  const long long fp = a.fine_pitch >> 3, cp = a.coarse_pitch >> 3;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < nvec; e += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(e % vc);
    const int tap = (int)((e / vc) & 3);
    const long long tc = e / (4 * vc);
    const int w = (int)(tc % Wc), h = (int)((tc / Wc) % Hc);
    const long long b = tc / ((long long)Wc * Hc);
    const long long tf = (b * a.H + 2 * h + (tap >> 1)) * a.W + 2 * w + (tap & 1);
    dst[tc * cp + tap * vc + v] = __ldg(src + tf * fp + v);
  }
}

__global__ void d2s_kernel(const kair_regroup_args a) {
  const int Hc = a.H >> 1, Wc = a.W >> 1, vc = a.C >> 3;
  const long long nvec = (long long)a.batch * a.H * a.W * vc;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < nvec; e += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(e % vc);
    const long long tf = e / vc;
    const int w = (int)(tf % a.W), h = (int)((tf / a.W) % a.H);
    const long long b = tf / ((long long)a.W * a.H);
    const long long tc = (b * Hc + (h >> 1)) * Wc + (w >> 1);
    const int tap = ((h & 1) << 1) | (w & 1);
    const long long ci = tc * a.coarse_pitch + (long long)tap * a.C + v * 8;
    float f[8];
    if (a.coarse_in_f32) {
      const float4 p = __ldg(reinterpret_cast<const float4*>(a.coarse_in_f32 + ci));
      const float4 q = __ldg(reinterpret_cast<const float4*>(a.coarse_in_f32 + ci + 4));
      f[0] = p.x; f[1] = p.y; f[2] = p.z; f[3] = p.w; f[4] = q.x; f[5] = q.y; f[6] = q.z; f[7] = q.w;
    } else {
      const uint4 p = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(a.coarse_in_bf16) + ci));
      const uint32_t w4[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w4[i]));
        f[2 * i] = f2.x;
        f[2 * i + 1] = f2.y;
      }
    }
    const long long fo = tf * a.fine_pitch + v * 8;
    if (a.fine_out_f32) {
      reinterpret_cast<float4*>(a.fine_out_f32 + fo)[0] = make_float4(f[0], f[1], f[2], f[3]);
      reinterpret_cast<float4*>(a.fine_out_f32 + fo)[1] = make_float4(f[4], f[5], f[6], f[7]);
    }
    if (a.fine_out_bf16) {
      uint4 o;
      o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
      o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
      *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(a.fine_out_bf16) + fo) = o;
    }
  }
}

static int regroup_check(const kair_regroup_args& a, const char* who) {
  KB_CHECK(a.batch > 0 && a.H > 0 && a.W > 0 && a.H % 2 == 0 && a.W % 2 == 0, "%s: H=%d, W=%d must be even", who, a.H, a.W);
  KB_CHECK(a.C > 0 && a.C % 8 == 0 && a.fine_pitch % 8 == 0 && a.coarse_pitch % 8 == 0 && a.fine_pitch >= a.C &&
               a.coarse_pitch >= 4 * a.C,
           "%s: C=%d and the pitches must be multiples of 8 elements", who, a.C);
  return 0;
}

int s2d_launch(const kair_regroup_args& a, cudaStream_t st) {
  if (regroup_check(a, "kair_space_to_depth")) return -1;
  KB_CHECK(a.fine_bf16 && a.coarse_bf16, "kair_space_to_depth: null pointer");
  const long long nvec = (long long)a.batch * a.H * a.W * (a.C >> 3);
  s2d_kernel<<<grid_for(nvec, 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int d2s_launch(const kair_regroup_args& a, cudaStream_t st) {
  if (regroup_check(a, "kair_depth_to_space")) return -1;
  KB_CHECK((a.coarse_in_f32 != nullptr) != (a.coarse_in_bf16 != nullptr), "kair_depth_to_space: exactly one input");
  KB_CHECK(a.fine_out_f32 || a.fine_out_bf16, "kair_depth_to_space: no output");
  const long long nvec = (long long)a.batch * a.H * a.W * (a.C >> 3);
  d2s_kernel<<<grid_for(nvec, 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ column sums
// Shared shape of the per-channel reductions below: a block of 256 threads covers `rows` token rows per iteration,
// thread (r, g) owns channel group g (VEC channels) of row r; per-thread fp32 partials, then one shared-memory
// reduction over r and one atomic per channel and block.
struct ColGeom {
  int groups, gp, rows;  // channel groups, groups rounded up to a power of two (<= 256), rows per iteration
};
static inline ColGeom col_geom(int C, int vec) {
  ColGeom g;
  g.groups = C / vec;
  g.gp = 1;
  while (g.gp < g.groups) g.gp <<= 1;
  g.rows = 256 / g.gp;
  return g;
}

template <int NV>
__device__ __forceinline__ void block_reduce_cols(float (&v)[NV], float* sh, int r, int g, int gp, int rows) {
  // sh: [NV][256]
#pragma unroll
  for (int i = 0; i < NV; ++i) sh[i * 256 + r * gp + g] = v[i];
  __syncthreads();
  if (r == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float s = 0.f;
      for (int rr = 0; rr < rows; ++rr) s += sh[i * 256 + rr * gp + g];
      v[i] = s;
    }
  }
}

__global__ void colsum_kernel(const kair_colsum_args a, int groups, int gp, int rows) {
  __shared__ float sh[2 * 256];
  const int g = threadIdx.x % gp, r = threadIdx.x / gp;
  float acc[2] = {0.f, 0.f};
  if (g < groups) {
    for (long long t = (long long)blockIdx.x * rows + r; t < a.T; t += (long long)gridDim.x * rows) {
      if (a.x_bf16) {
        const float2 f = __bfloat1622float2(
            *reinterpret_cast<const __nv_bfloat162*>(reinterpret_cast<const bf16*>(a.x_bf16) + t * a.pitch + 2 * g));
        acc[0] += f.x;
        acc[1] += f.y;
      } else {
        const float2 f = *reinterpret_cast<const float2*>(a.x_f32 + t * a.pitch + 2 * g);
        acc[0] += f.x;
        acc[1] += f.y;
      }
    }
  }
  block_reduce_cols<2>(acc, sh, r, g, gp, rows);
  if (r == 0 && g < groups) {
    atomicAdd(a.out + 2 * g, acc[0] * a.scale);
    atomicAdd(a.out + 2 * g + 1, acc[1] * a.scale);
  }
}

int colsum_launch(const kair_colsum_args& a, cudaStream_t st) {
  KB_CHECK((a.x_bf16 != nullptr) != (a.x_f32 != nullptr) && a.out, "kair_colsum: exactly one input and an output");
  KB_CHECK(a.C > 0 && a.C % 2 == 0 && a.C <= 512 && a.pitch >= a.C && a.pitch % 2 == 0 && a.T > 0, "kair_colsum: bad shape");
  const ColGeom g = col_geom(a.C, 2);
  colsum_kernel<<<grid_for(a.T, g.rows * 16, 4), 256, 0, st>>>(a, g.groups, g.gp, g.rows);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ BatchNorm
__device__ __forceinline__ void ld_bf16x4(const void* base, long long idx, float (&f)[4]) {
  const uint2 p = *reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(base) + idx);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&p.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&p.y));
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
}
__device__ __forceinline__ void st_bf16x4(void* base, long long idx, const float (&f)[4]) {
  uint2 o;
  o.x = pack_bf16x2(f[0], f[1]);
  o.y = pack_bf16x2(f[2], f[3]);
  *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(base) + idx) = o;
}

__device__ __forceinline__ void ld_dy4(const kair_bn_args& a, long long idx, float (&f)[4]) {
  if (a.dy_f32) {
    const float4 v = *reinterpret_cast<const float4*>(a.dy_f32 + idx);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  } else {
    ld_bf16x4(a.dy, idx, f);
  }
}

__global__ void bn_stats_kernel(const kair_bn_args a, int groups, int gp, int rows) {
  __shared__ float sh[8 * 256];
  const int g = threadIdx.x % gp, r = threadIdx.x / gp;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (g < groups) {
    for (long long t = (long long)blockIdx.x * rows + r; t < a.T; t += (long long)gridDim.x * rows) {
      const float4 z = *reinterpret_cast<const float4*>(a.z + t * a.pitch + 4 * g);
      acc[0] += z.x; acc[1] += z.y; acc[2] += z.z; acc[3] += z.w;
      acc[4] += z.x * z.x; acc[5] += z.y * z.y; acc[6] += z.z * z.z; acc[7] += z.w * z.w;
    }
  }
  block_reduce_cols<8>(acc, sh, r, g, gp, rows);
  if (r == 0 && g < groups) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      atomicAdd(a.acc + 4 * g + i, (double)acc[i]);
      atomicAdd(a.acc + a.C + 4 * g + i, (double)acc[4 + i]);
    }
  }
}

__global__ void bn_finalize_kernel(const kair_bn_args a) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < a.C) {
    const double n = (double)a.T;
    const double mean = a.acc[c] / n;
    double var = a.acc[a.C + c] / n - mean * mean;
    if (var < 0.0) var = 0.0;
    a.mean[c] = (float)mean;
    a.rstd[c] = (float)(1.0 / sqrt(var + (double)a.eps));
    if (a.running_mean) a.running_mean[c] = (1.f - a.momentum) * a.running_mean[c] + a.momentum * (float)mean;
    if (a.running_var) {
      const double unbiased = var * n / (n > 1.0 ? n - 1.0 : 1.0);
      a.running_var[c] = (1.f - a.momentum) * a.running_var[c] + a.momentum * (float)unbiased;
    }
    a.acc[c] = 0.0;
    a.acc[a.C + c] = 0.0;
  }
  if (c == 0 && a.num_batches_tracked) *a.num_batches_tracked += 1;
}

__global__ void bn_apply_kernel(const kair_bn_args a) {
  const int groups = a.C >> 2;
  const long long n = a.T * groups;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(e % groups);
    const long long t = e / groups;
    const float4 z = *reinterpret_cast<const float4*>(a.z + t * a.pitch + 4 * g);
    const float zz[4] = {z.x, z.y, z.z, z.w};
    float y[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = 4 * g + i;
      const float sc = __ldg(a.gamma + c) * __ldg(a.rstd + c);
      float v = (zz[i] - __ldg(a.mean + c)) * sc + __ldg(a.beta + c);
      if (a.act == KAIR_ACT_LRELU) v = v > 0.f ? v : v * a.act_slope;
      y[i] = v;
    }
    if (a.out_bf16) st_bf16x4(a.out_bf16, t * a.pitch + 4 * g, y);
    if (a.out_f32) *reinterpret_cast<float4*>(a.out_f32 + t * a.pitch + 4 * g) = make_float4(y[0], y[1], y[2], y[3]);
  }
}

__global__ void bn_bwd_reduce_kernel(const kair_bn_args a, int groups, int gp, int rows) {
  __shared__ float sh[8 * 256];
  const int g = threadIdx.x % gp, r = threadIdx.x / gp;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (g < groups) {
    float mu[4], rs[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      mu[i] = a.mean[4 * g + i];
      rs[i] = a.rstd[4 * g + i];
    }
    for (long long t = (long long)blockIdx.x * rows + r; t < a.T; t += (long long)gridDim.x * rows) {
      const long long idx = t * a.pitch + 4 * g;
      const float4 z = *reinterpret_cast<const float4*>(a.z + idx);
      const float zz[4] = {z.x, z.y, z.z, z.w};
      float dy[4], y[4];
      ld_dy4(a, idx, dy);
      if (a.act == KAIR_ACT_LRELU) ld_bf16x4(a.y, idx, y);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float gq = dy[i];
        if (a.act == KAIR_ACT_LRELU && !(y[i] > 0.f)) gq *= a.act_slope;
        acc[i] += gq;
        acc[4 + i] += gq * (zz[i] - mu[i]) * rs[i];
      }
    }
  }
  block_reduce_cols<8>(acc, sh, r, g, gp, rows);
  if (r == 0 && g < groups) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      atomicAdd(a.dbeta + 4 * g + i, acc[i]);
      atomicAdd(a.dgamma + 4 * g + i, acc[4 + i]);
    }
  }
}

__global__ void bn_bwd_apply_kernel(const kair_bn_args a, int groups, int gp, int rows) {
  __shared__ float sh[4 * 256];
  const int g = threadIdx.x % gp, r = threadIdx.x / gp;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  if (g < groups) {
    const float inv_n = 1.0f / (float)a.T;
    float mu[4], rs[4], k0[4], k1[4], k2[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = 4 * g + i;
      mu[i] = a.mean[c];
      rs[i] = a.rstd[c];
      k0[i] = a.gamma[c] * rs[i];
      k1[i] = a.dbeta[c] * inv_n;
      k2[i] = a.dgamma[c] * inv_n;
    }
    for (long long t = (long long)blockIdx.x * rows + r; t < a.T; t += (long long)gridDim.x * rows) {
      const long long idx = t * a.pitch + 4 * g;
      const float4 z = *reinterpret_cast<const float4*>(a.z + idx);
      const float zz[4] = {z.x, z.y, z.z, z.w};
      float dy[4], y[4], dz[4];
      ld_dy4(a, idx, dy);
      if (a.act == KAIR_ACT_LRELU) ld_bf16x4(a.y, idx, y);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float gq = dy[i];
        if (a.act == KAIR_ACT_LRELU && !(y[i] > 0.f)) gq *= a.act_slope;
        dz[i] = k0[i] * (gq - k1[i] - (zz[i] - mu[i]) * rs[i] * k2[i]);
        acc[i] += dz[i];
      }
      st_bf16x4(a.out_bf16, idx, dz);
    }
  }
  if (a.dbias) {
    block_reduce_cols<4>(acc, sh, r, g, gp, rows);
    if (r == 0 && g < groups) {
#pragma unroll
      for (int i = 0; i < 4; ++i) atomicAdd(a.dbias + 4 * g + i, acc[i]);
    }
  }
}

__global__ void bn_fold_kernel(const kair_bn_args a) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < a.C) {
    const float sc = a.gamma[c] * rsqrtf(a.running_var[c] + a.eps);
    a.scale[c] = sc;
    a.shift[c] = a.beta[c] - a.running_mean[c] * sc;
  }
}

static int bn_check(const kair_bn_args& a, const char* who) {
  KB_CHECK(a.C > 0 && a.C % 4 == 0 && a.C <= 512, "%s: C=%d must be a multiple of 4, <= 512", who, a.C);
  return 0;
}
static int bn_check_rows(const kair_bn_args& a, const char* who) {
  if (bn_check(a, who)) return -1;
  KB_CHECK(a.T > 0 && a.pitch >= a.C && a.pitch % 4 == 0 && a.z, "%s: bad token matrix", who);
  return 0;
}

int bn_stats_launch(const kair_bn_args& a, cudaStream_t st) {
  if (bn_check_rows(a, "kair_bn_stats")) return -1;
  KB_CHECK(a.acc, "kair_bn_stats: null accumulator");
  const ColGeom g = col_geom(a.C, 4);
  bn_stats_kernel<<<grid_for(a.T, g.rows * 8, 4), 256, 0, st>>>(a, g.groups, g.gp, g.rows);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int bn_finalize_launch(const kair_bn_args& a, cudaStream_t st) {
  if (bn_check(a, "kair_bn_finalize")) return -1;
  KB_CHECK(a.acc && a.mean && a.rstd && a.T > 0, "kair_bn_finalize: null pointer");
  bn_finalize_kernel<<<(a.C + 127) / 128, 128, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int bn_apply_launch(const kair_bn_args& a, cudaStream_t st) {
  if (bn_check_rows(a, "kair_bn_apply")) return -1;
  KB_CHECK(a.mean && a.rstd && a.gamma && a.beta && (a.out_bf16 || a.out_f32), "kair_bn_apply: null pointer");
  bn_apply_kernel<<<grid_for(a.T * (a.C >> 2), 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int bn_bwd_reduce_launch(const kair_bn_args& a, cudaStream_t st) {
  if (bn_check_rows(a, "kair_bn_bwd_reduce")) return -1;
  KB_CHECK((a.dy || a.dy_f32) && a.mean && a.rstd && a.dgamma && a.dbeta && (a.act == KAIR_ACT_NONE || a.y),
           "kair_bn_bwd_reduce: null pointer");
  const ColGeom g = col_geom(a.C, 4);
  bn_bwd_reduce_kernel<<<grid_for(a.T, g.rows * 8, 4), 256, 0, st>>>(a, g.groups, g.gp, g.rows);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int bn_bwd_apply_launch(const kair_bn_args& a, cudaStream_t st) {
  if (bn_check_rows(a, "kair_bn_bwd_apply")) return -1;
  KB_CHECK((a.dy || a.dy_f32) && a.mean && a.rstd && a.gamma && a.dgamma && a.dbeta && a.out_bf16 &&
               (a.act == KAIR_ACT_NONE || a.y),
           "kair_bn_bwd_apply: null pointer");
  const ColGeom g = col_geom(a.C, 4);
  bn_bwd_apply_kernel<<<grid_for(a.T, g.rows * 8, 4), 256, 0, st>>>(a, g.groups, g.gp, g.rows);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int bn_fold_launch(const kair_bn_args& a, cudaStream_t st) {
  if (bn_check(a, "kair_bn_fold")) return -1;
  KB_CHECK(a.gamma && a.beta && a.running_mean && a.running_var && a.scale && a.shift, "kair_bn_fold: null pointer");
  bn_fold_kernel<<<(a.C + 127) / 128, 128, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ SR heads
// PixelShuffle(r): out[(b, r*h+dy, r*w+dx)][c] = in[(b,h,w)][c*r*r + dy*r + dx]; one thread per output channel pair.
template <typename E, typename E2>   // (bf16, __nv_bfloat162) or (float, float2)
__global__ void pixel_shuffle_kernel(const kair_shuffle_args a) {
  const int r = a.r, Ho = a.H * r, Wo = a.W * r, rr = r * r;
  const E* in = reinterpret_cast<const E*>(a.in);
  E* out = reinterpret_cast<E*>(a.out);
  const int pc = a.C >> 1;
  const long long n = (long long)a.batch * Ho * Wo * pc;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(e % pc) * 2;
    const long long to = e / pc;
    const int wo = (int)(to % Wo), ho = (int)((to / Wo) % Ho);
    const long long b = to / ((long long)Wo * Ho);
    const long long ti = (b * a.H + ho / r) * a.W + wo / r;
    const int sub = (ho % r) * r + (wo % r);
    E2 v;
    v.x = in[ti * a.in_pitch + (long long)c * rr + sub];
    v.y = in[ti * a.in_pitch + (long long)(c + 1) * rr + sub];
    *reinterpret_cast<E2*>(out + to * a.out_pitch + c) = v;
  }
}

// r = 2 fast path (every x2 / x4 / x8 head of network_swinir.py:614-631 is a chain of PixelShuffle(2)): a thread moves one
// 16-byte chunk = channels (2q, 2q+1) x the 4 sub-pixels of an input pixel and writes four bf16 pairs, so a warp reads
// 512 contiguous bytes and writes 4 x 128 contiguous bytes; 32-bit index chain when the element count allows.
// (The generic kernel above gathers 2-byte elements behind five 64-bit div/mods per 4 output bytes: 1.1 TB/s.)
template <typename idx_t>
__global__ void pixel_shuffle2_kernel(const kair_shuffle_args a) {
  const idx_t cpp = (idx_t)(a.C >> 1);
  const idx_t n = (idx_t)a.batch * (idx_t)a.H * (idx_t)a.W * cpp;
  const idx_t Wi = (idx_t)a.W, Wo = Wi * 2;
  const bf16* in = reinterpret_cast<const bf16*>(a.in);
  uint32_t* out = reinterpret_cast<uint32_t*>(a.out);
  const long long op = a.out_pitch >> 1;
  for (idx_t e = (idx_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (idx_t)gridDim.x * blockDim.x) {
    const idx_t ti = e / cpp, q = e - ti * cpp;
    const idx_t row = ti / Wi, w = ti - row * Wi;   // row = b * H + h
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + (long long)ti * a.in_pitch) + q);
    // v.x = (c0 s0, c0 s1), v.y = (c0 s2, c0 s3), v.z = (c1 s0, c1 s1), v.w = (c1 s2, c1 s3), s = dy * 2 + dx
    const long long o0 = ((long long)(2 * row) * Wo + 2 * w) * op + q, o1 = o0 + (long long)Wo * op;
    out[o0] = __byte_perm(v.x, v.z, 0x5410);
    out[o0 + op] = __byte_perm(v.x, v.z, 0x7632);
    out[o1] = __byte_perm(v.y, v.w, 0x5410);
    out[o1 + op] = __byte_perm(v.y, v.w, 0x7632);
  }
}

__global__ void upsample_nearest2_kernel(const kair_shuffle_args a) {
  const int Ho = a.H * 2, Wo = a.W * 2, vc = a.C >> 3;
  const uint4* in = reinterpret_cast<const uint4*>(a.in);
  uint4* out = reinterpret_cast<uint4*>(a.out);
  const long long ip = a.in_pitch >> 3, op = a.out_pitch >> 3;
  const long long n = (long long)a.batch * Ho * Wo * vc;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(e % vc);
    const long long to = e / vc;
    const int wo = (int)(to % Wo), ho = (int)((to / Wo) % Ho);
    const long long b = to / ((long long)Wo * Ho);
    const long long ti = (b * a.H + (ho >> 1)) * a.W + (wo >> 1);
    out[to * op + v] = __ldg(in + ti * ip + v);
  }
}

// one thread per output pixel pair along x (coalesced fp32 stores; the bf16 row of an input pixel is L1-resident)
__global__ void pixel_shuffle_out_kernel(const kair_shuffle_out_args a) {
  const int r = a.r, Ho = a.H * r, Wo = a.W * r, rr = r * r;
  const bf16* in = reinterpret_cast<const bf16*>(a.in);
  const long long n = (long long)a.batch * a.co * Ho * Wo;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const int wo = (int)(e % Wo), ho = (int)((e / Wo) % Ho);
    const int c = (int)((e / ((long long)Wo * Ho)) % a.co);
    const long long b = e / ((long long)Wo * Ho * a.co);
    const long long ti = (b * a.H + ho / r) * a.W + wo / r;
    const long long ii = ti * a.in_pitch + (long long)c * rr + (ho % r) * r + (wo % r);
    const float v = a.fp32 ? reinterpret_cast<const float*>(a.in)[ii] : __bfloat162float(in[ii]);
    a.out[e] = v * a.scale + a.mean[c];
  }
}

int pixel_shuffle_out_launch(const kair_shuffle_out_args& a, cudaStream_t st) {
  KB_CHECK(a.in && a.out && a.mean && a.batch > 0 && a.H > 0 && a.W > 0, "kair_pixel_shuffle_out: null pointer / empty");
  KB_CHECK(a.r >= 1 && a.r <= 8 && a.co >= 1 && a.in_pitch >= (long long)a.co * a.r * a.r,
           "kair_pixel_shuffle_out: bad geometry (r=%d, co=%d)", a.r, a.co);
  const long long n = (long long)a.batch * a.co * a.H * a.W * a.r * a.r;
  pixel_shuffle_out_kernel<<<grid_for(n, 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// adjoint of pixel_shuffle_kernel: a thread reads one bf16 pair of the high-resolution gradient (coalesced) and
// scatters it to the two channel groups of the low-resolution pixel
__global__ void pixel_shuffle_bwd_kernel(const kair_shuffle_args a) {
  const int r = a.r, Ho = a.H * r, Wo = a.W * r, rr = r * r;
  const bf16* in = reinterpret_cast<const bf16*>(a.in);
  bf16* out = reinterpret_cast<bf16*>(a.out);
  const int pc = a.C >> 1;
  const long long n = (long long)a.batch * Ho * Wo * pc;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(e % pc) * 2;
    const long long to = e / pc;
    const int wo = (int)(to % Wo), ho = (int)((to / Wo) % Ho);
    const long long b = to / ((long long)Wo * Ho);
    const long long ti = (b * a.H + ho / r) * a.W + wo / r;
    const int sub = (ho % r) * r + (wo % r);
    const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(in + to * a.in_pitch + c);
    out[ti * a.out_pitch + (long long)c * rr + sub] = v.x;
    out[ti * a.out_pitch + (long long)(c + 1) * rr + sub] = v.y;
  }
}

// r = 2 fast path, the mirror of pixel_shuffle2_kernel: four coalesced bf16-pair reads (one per sub-pixel), one 16-byte write
template <typename idx_t>
__global__ void pixel_shuffle2_bwd_kernel(const kair_shuffle_args a) {
  const idx_t cpp = (idx_t)(a.C >> 1);
  const idx_t n = (idx_t)a.batch * (idx_t)a.H * (idx_t)a.W * cpp;
  const idx_t Wi = (idx_t)a.W, Wo = Wi * 2;
  const uint32_t* in = reinterpret_cast<const uint32_t*>(a.in);
  bf16* out = reinterpret_cast<bf16*>(a.out);
  const long long ip = a.in_pitch >> 1;
  for (idx_t e = (idx_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (idx_t)gridDim.x * blockDim.x) {
    const idx_t ti = e / cpp, q = e - ti * cpp;
    const idx_t row = ti / Wi, w = ti - row * Wi;   // row = b * H + h
    const long long i0 = ((long long)(2 * row) * Wo + 2 * w) * ip + q, i1 = i0 + (long long)Wo * ip;
    const uint32_t s0 = __ldg(in + i0), s1 = __ldg(in + i0 + ip), s2 = __ldg(in + i1), s3 = __ldg(in + i1 + ip);
    uint4 v;   // (c0 s0, c0 s1), (c0 s2, c0 s3), (c1 s0, c1 s1), (c1 s2, c1 s3)
    v.x = __byte_perm(s0, s1, 0x5410);
    v.y = __byte_perm(s2, s3, 0x5410);
    v.z = __byte_perm(s0, s1, 0x7632);
    v.w = __byte_perm(s2, s3, 0x7632);
    reinterpret_cast<uint4*>(out + (long long)ti * a.out_pitch)[q] = v;
  }
}

// adjoint of nearest x2: low-resolution gradient = sum of the 2x2 footprint (fp32), then the LeakyReLU mask
__global__ void upsample_nearest2_bwd_kernel(const kair_shuffle_args a) {
  const int Wo = a.W * 2, vc = a.C >> 3;
  const uint4* in = reinterpret_cast<const uint4*>(a.in);
  uint4* out = reinterpret_cast<uint4*>(a.out);
  const uint4* mask = reinterpret_cast<const uint4*>(a.mask);
  const long long ip = a.in_pitch >> 3, op = a.out_pitch >> 3, mp = a.mask_pitch >> 3;
  const long long n = (long long)a.batch * a.H * a.W * vc;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(e % vc);
    const long long ti = e / vc;
    const int w = (int)(ti % a.W), h = (int)((ti / a.W) % a.H);
    const long long b = ti / ((long long)a.W * a.H);
    const long long t00 = (b * 2 * a.H + 2 * h) * Wo + 2 * w;
    uint4 q[4];
    q[0] = __ldg(in + t00 * ip + v);
    q[1] = __ldg(in + (t00 + 1) * ip + v);
    q[2] = __ldg(in + (t00 + Wo) * ip + v);
    q[3] = __ldg(in + (t00 + Wo + 1) * ip + v);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q[k]);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __bfloat1622float2(h2[j]);
        acc[2 * j] += f.x;
        acc[2 * j + 1] += f.y;
      }
    }
    if (mask) {
      const uint4 m = __ldg(mask + ti * mp + v);
      const __nv_bfloat162* m2 = reinterpret_cast<const __nv_bfloat162*>(&m);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __bfloat1622float2(m2[j]);
        if (!(f.x > 0.f)) acc[2 * j] *= a.mask_slope;
        if (!(f.y > 0.f)) acc[2 * j + 1] *= a.mask_slope;
      }
    }
    uint4 o;
    o.x = pack_bf16x2(acc[0], acc[1]);
    o.y = pack_bf16x2(acc[2], acc[3]);
    o.z = pack_bf16x2(acc[4], acc[5]);
    o.w = pack_bf16x2(acc[6], acc[7]);
    out[ti * op + v] = o;
  }
}

__global__ void pixel_shuffle_out_bwd_kernel(const kair_shuffle_out_args a) {
  const int r = a.r, Ho = a.H * r, Wo = a.W * r, rr = r * r;
  bf16* in = reinterpret_cast<bf16*>(const_cast<void*>(a.in));
  const long long n = (long long)a.batch * a.co * Ho * Wo;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const int wo = (int)(e % Wo), ho = (int)((e / Wo) % Ho);
    const int c = (int)((e / ((long long)Wo * Ho)) % a.co);
    const long long b = e / ((long long)Wo * Ho * a.co);
    const long long ti = (b * a.H + ho / r) * a.W + wo / r;
    in[ti * a.in_pitch + (long long)c * rr + (ho % r) * r + (wo % r)] = __float2bfloat16(a.out[e] * a.scale);
  }
}

int pixel_shuffle_out_bwd_launch(const kair_shuffle_out_args& a, cudaStream_t st) {
  KB_CHECK(a.in && a.out && a.batch > 0 && a.H > 0 && a.W > 0, "kair_pixel_shuffle_out_bwd: null pointer / empty");
  KB_CHECK(a.r >= 1 && a.r <= 8 && a.co >= 1 && a.in_pitch >= (long long)a.co * a.r * a.r,
           "kair_pixel_shuffle_out_bwd: bad geometry (r=%d, co=%d)", a.r, a.co);
  const long long n = (long long)a.batch * a.co * a.H * a.W * a.r * a.r;
  pixel_shuffle_out_bwd_kernel<<<grid_for(n, 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int pixel_shuffle_bwd_launch(const kair_shuffle_args& a, cudaStream_t st) {
  KB_CHECK(a.in && a.out && a.batch > 0 && a.H > 0 && a.W > 0, "kair_pixel_shuffle_bwd: null pointer / empty");
  KB_CHECK(a.r >= 2 && a.r <= 4 && a.C % 2 == 0 && a.in_pitch % 2 == 0 && a.out_pitch >= (long long)a.C * a.r * a.r &&
               a.in_pitch >= a.C && !a.mask,
           "kair_pixel_shuffle_bwd: bad geometry (r=%d, C=%d) or mask given", a.r, a.C);
  const long long n = (long long)a.batch * a.H * a.W * a.r * a.r * (a.C >> 1);
  if (a.r == 2 && a.out_pitch % 8 == 0 && a.in_pitch % 2 == 0) {
    const long long n2 = n >> 2;   // 16-byte output chunks
    const int grid = grid_for(n2, 256, 16);
    if (n2 + (long long)grid * 256 < (1ll << 32)) pixel_shuffle2_bwd_kernel<unsigned><<<grid, 256, 0, st>>>(a);
    else pixel_shuffle2_bwd_kernel<long long><<<grid, 256, 0, st>>>(a);
    KB_CUDA(cudaGetLastError());
    return 0;
  }
  pixel_shuffle_bwd_kernel<<<grid_for(n, 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int upsample_nearest2_bwd_launch(const kair_shuffle_args& a, cudaStream_t st) {
  KB_CHECK(a.in && a.out && a.batch > 0 && a.H > 0 && a.W > 0, "kair_upsample_nearest2_bwd: null pointer / empty");
  KB_CHECK(a.C % 8 == 0 && a.in_pitch % 8 == 0 && a.out_pitch % 8 == 0 && a.in_pitch >= a.C && a.out_pitch >= a.C &&
               (!a.mask || (a.mask_pitch % 8 == 0 && a.mask_pitch >= a.C)),
           "kair_upsample_nearest2_bwd: C and pitches must be multiples of 8");
  const long long n = (long long)a.batch * a.H * a.W * (a.C >> 3);
  upsample_nearest2_bwd_kernel<<<grid_for(n, 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int pixel_shuffle_launch(const kair_shuffle_args& a, cudaStream_t st) {
  KB_CHECK(a.in && a.out && a.batch > 0 && a.H > 0 && a.W > 0, "kair_pixel_shuffle: null pointer / empty");
  KB_CHECK(a.r >= 2 && a.r <= 4 && a.C % 2 == 0 && a.out_pitch % 2 == 0 && a.in_pitch >= (long long)a.C * a.r * a.r &&
               a.out_pitch >= a.C,
           "kair_pixel_shuffle: bad geometry (r=%d, C=%d)", a.r, a.C);
  const long long n = (long long)a.batch * a.H * a.W * a.r * a.r * (a.C >> 1);
  if (a.fp32) {
    pixel_shuffle_kernel<float, float2><<<grid_for(n, 256, 16), 256, 0, st>>>(a);
    KB_CUDA(cudaGetLastError());
    return 0;
  }
  if (a.r == 2 && a.in_pitch % 8 == 0) {
    const long long n2 = n >> 2;   // 16-byte input chunks
    const int grid = grid_for(n2, 256, 16);
    if (n2 + (long long)grid * 256 < (1ll << 32)) pixel_shuffle2_kernel<unsigned><<<grid, 256, 0, st>>>(a);
    else pixel_shuffle2_kernel<long long><<<grid, 256, 0, st>>>(a);
    KB_CUDA(cudaGetLastError());
    return 0;
  }
  pixel_shuffle_kernel<bf16, __nv_bfloat162><<<grid_for(n, 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int upsample_nearest2_launch(const kair_shuffle_args& a0, cudaStream_t st) {
  kair_shuffle_args a = a0;
  if (a.fp32) {   // a pure 16-byte mover: fp32 rows are twice as many 2-byte elements
    a.C *= 2;
    a.in_pitch *= 2;
    a.out_pitch *= 2;
  }
  KB_CHECK(a.in && a.out && a.batch > 0 && a.H > 0 && a.W > 0, "kair_upsample_nearest2: null pointer / empty");
  KB_CHECK(a.C % 8 == 0 && a.in_pitch % 8 == 0 && a.out_pitch % 8 == 0 && a.in_pitch >= a.C && a.out_pitch >= a.C,
           "kair_upsample_nearest2: C and pitches must be multiples of 8");
  const long long n = (long long)a.batch * a.H * a.W * 4 * (a.C >> 3);
  upsample_nearest2_kernel<<<grid_for(n, 256, 16), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kb
