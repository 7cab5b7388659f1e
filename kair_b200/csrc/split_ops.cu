// SPLIT PRECISION helpers: the memory-bound kernels of the fp32-accurate mode (include/kair_b200.h, "SPLIT PRECISION").
//
// The tensor cores take bf16 operands, so the fp32-accurate mode carries every GEMM operand as a sum of bf16
// components (c0 = bf16(x), c1 = bf16(x - c0), c2 = bf16(x - c0 - c1)) laid out as extra K segments; the GEMM kernels
// themselves (gemm_tn.cu / gemm_nt.cu) are unchanged and accumulate the component products in fp32 (TMEM).
//   kair_split_terms   fp32 rows -> split bf16 operand rows, fused with the element-wise stage that precedes the GEMM
//                      (exact-erf GELU and its derivative: network_swinir.py:21, 26; LeakyReLU masks; DropPath scales)
//   kair_ln_bwd_split  LayerNorm backward + residual gradient on split x-hat (autograd of network_swinir.py:245, 277)
// Both are HBM-bound: 4 B read + 2*terms B written per element; one thread per 8-column group, 16-byte accesses.
#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

namespace {

__device__ __forceinline__ float gelu_exact(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752f));
}
__device__ __forceinline__ float gelu_exact_grad(float x) {
  return 0.5f * (1.0f + erff(x * 0.70710678118654752f)) + x * 0.3989422804014327f * expf(-0.5f * x * x);
}

template <int TERMS>
__global__ void __launch_bounds__(256) split_terms_kernel(const kair_split_args a) {
  const int groups = a.n >> 3;
  const long long total = a.T * groups;
  bf16* out = reinterpret_cast<bf16*>(a.out);
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const long long t = e / groups;
    const int c0 = (int)(e - t * groups) * 8;
    float v[8];
    {
      const float4* xp = reinterpret_cast<const float4*>(a.x + t * a.x_pitch + c0);
      const float4 x0 = xp[0], x1 = xp[1];
      v[0] = x0.x; v[1] = x0.y; v[2] = x0.z; v[3] = x0.w; v[4] = x1.x; v[5] = x1.y; v[6] = x1.z; v[7] = x1.w;
    }
    if (a.op == KAIR_SPLIT_GELU) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = gelu_exact(v[i]);
    } else if (a.op == KAIR_SPLIT_LRELU) {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * a.slope;
    } else if (a.op == KAIR_SPLIT_DGELU || a.op == KAIR_SPLIT_ACT_BWD) {
      const float4* up = reinterpret_cast<const float4*>(a.aux + t * a.aux_pitch + c0);
      const float4 u0 = up[0], u1 = up[1];
      const float u[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
      if (a.op == KAIR_SPLIT_DGELU) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] *= gelu_exact_grad(u[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = u[i] > 0.f ? v[i] : v[i] * a.slope;
      }
    }
    if (a.row_scale) {
      const float s = __ldg(a.row_scale + t / a.rows_per_sample);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] *= s;
    }
    if (a.ones_col >= c0 && a.ones_col < c0 + 8) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (c0 + i == a.ones_col) v[i] = 1.0f;
    }
    if (a.out_f32) {
      float4* op = reinterpret_cast<float4*>(a.out_f32 + t * a.out_f32_pitch + c0);
      op[0] = make_float4(v[0], v[1], v[2], v[3]);
      op[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
    if (out) {
      float comp[3][8];
#pragma unroll
      for (int i = 0; i < 8; ++i) split_components(v[i], comp[0][i], comp[1][i], comp[2][i]);
      bf16* row = out + t * a.out_pitch + c0;
#pragma unroll
      for (int s = 0; s < TERMS; ++s) {
        const int w = split_seg_act(TERMS, s);
        const uint4 pk = make_uint4(pack_bf16x2(comp[w][0], comp[w][1]), pack_bf16x2(comp[w][2], comp[w][3]),
                                    pack_bf16x2(comp[w][4], comp[w][5]), pack_bf16x2(comp[w][6], comp[w][7]));
        *reinterpret_cast<uint4*>(row + (size_t)s * a.n) = pk;
      }
    }
  }
}

// one warp per row, lane owns columns lane + 32*i (pitch <= 256)
template <int TERMS>
__global__ void __launch_bounds__(256) ln_bwd_split_kernel(const kair_ln_bwd_split_args a) {
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  const bf16* xh = reinterpret_cast<const bf16*>(a.xhat);
  bf16* out = reinterpret_cast<bf16*>(a.out);
  const float inv_c = 1.0f / (float)a.C;
  constexpr int NC = TERMS == 6 ? 3 : (TERMS == 3 ? 2 : 1);   // components stored in x-hat (segments 0 .. NC-1)
  for (long long row = (long long)blockIdx.x * warps + (threadIdx.x >> 5); row < a.T;
       row += (long long)gridDim.x * warps) {
    float g[8], x[8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = lane + 32 * i;
      g[i] = 0.f;
      x[i] = 0.f;
      if (c < a.C) {
        g[i] = a.g[row * a.pitch + c];
        float xv = 0.f;
#pragma unroll
        for (int k = NC - 1; k >= 0; --k) xv += __bfloat162float(xh[row * a.xhat_pitch + (size_t)k * a.pitch + c]);
        x[i] = xv;
        s1 += g[i];
        s2 = fmaf(g[i], xv, s2);
      }
    }
    const float m1 = warp_sum(s1) * inv_c, m2 = warp_sum(s2) * inv_c;
    const float rstd = a.rstd[row];
    const float sc = a.row_scale ? __ldg(a.row_scale + row / a.rows_per_sample) : 1.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = lane + 32 * i;
      if (c >= a.pitch) continue;
      float dx = 0.f;
      if (c < a.C) {
        dx = rstd * (g[i] - m1 - x[i] * m2);
        if (a.resid) dx += a.resid[row * a.pitch + c];
      }
      a.outF[row * a.pitch + c] = dx;
      if (out) {
        float comp[3];
        split_components(dx * sc, comp[0], comp[1], comp[2]);
#pragma unroll
        for (int s = 0; s < TERMS; ++s)
          out[row * a.out_pitch + (size_t)s * a.pitch + c] = __float2bfloat16(comp[split_seg_act(TERMS, s)]);
      }
    }
  }
}

__global__ void __launch_bounds__(256) axpby_kernel(const kair_axpby_args a) {
  const int groups = a.n_pad >> 3;
  const long long total = a.T * groups;
  const bf16* mask = reinterpret_cast<const bf16*>(a.mask);
  bf16* ob = reinterpret_cast<bf16*>(a.out_bf16);
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const long long t = e / groups;
    const int c0 = (int)(e - t * groups) * 8;
    float v[8];
    if (c0 < a.n) {
      const float4* xp = reinterpret_cast<const float4*>(a.x + t * a.x_pitch + c0);
      const float4 x0 = xp[0], x1 = xp[1];
      v[0] = a.a * x0.x; v[1] = a.a * x0.y; v[2] = a.a * x0.z; v[3] = a.a * x0.w;
      v[4] = a.a * x1.x; v[5] = a.a * x1.y; v[6] = a.a * x1.z; v[7] = a.a * x1.w;
      if (a.y) {
        const float4* yp = reinterpret_cast<const float4*>(a.y + t * a.y_pitch + c0);
        const float4 y0 = yp[0], y1 = yp[1];
        v[0] = fmaf(a.b, y0.x, v[0]); v[1] = fmaf(a.b, y0.y, v[1]); v[2] = fmaf(a.b, y0.z, v[2]); v[3] = fmaf(a.b, y0.w, v[3]);
        v[4] = fmaf(a.b, y1.x, v[4]); v[5] = fmaf(a.b, y1.y, v[5]); v[6] = fmaf(a.b, y1.z, v[6]); v[7] = fmaf(a.b, y1.w, v[7]);
      }
      if (a.out_f32) {
        float4* op = reinterpret_cast<float4*>(a.out_f32 + t * a.out_f32_pitch + c0);
        op[0] = make_float4(v[0], v[1], v[2], v[3]);
        op[1] = make_float4(v[4], v[5], v[6], v[7]);
      }
      if (mask) {
        const uint4 m = *reinterpret_cast<const uint4*>(mask + t * a.mask_pitch + c0);
        const __nv_bfloat162* m2 = reinterpret_cast<const __nv_bfloat162*>(&m);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = __bfloat1622float2(m2[i]);
          if (!(f.x > 0.f)) v[2 * i] *= a.mask_slope;
          if (!(f.y > 0.f)) v[2 * i + 1] *= a.mask_slope;
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = 0.f;
    }
    if (ob) {
      const uint4 pk = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                  pack_bf16x2(v[6], v[7]));
      *reinterpret_cast<uint4*>(ob + t * a.out_bf16_pitch + c0) = pk;
    }
  }
}

}  // namespace

int axpby_launch(const kair_axpby_args& a, cudaStream_t st) {
  KB_CHECK(a.x && (a.out_f32 || a.out_bf16) && a.T > 0, "kair_axpby: null pointer / empty input");
  KB_CHECK(a.n > 0 && a.n % 8 == 0 && a.n_pad >= a.n && a.n_pad % 8 == 0, "kair_axpby: n=%d n_pad=%d", a.n, a.n_pad);
  KB_CHECK(a.x_pitch % 4 == 0 && (!a.y || a.y_pitch % 4 == 0) && (!a.out_f32 || a.out_f32_pitch % 4 == 0) &&
               (!a.out_bf16 || a.out_bf16_pitch % 8 == 0) && (!a.mask || a.mask_pitch % 8 == 0),
           "kair_axpby: pitches must keep 16-byte alignment");
  KB_CHECK(a.n_pad == a.n || a.out_bf16, "kair_axpby: n_pad > n only pads the bf16 output");
  const long long total = a.T * (a.n_pad >> 3);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  axpby_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int split_terms_launch(const kair_split_args& a, cudaStream_t st) {
  KB_CHECK(a.x && (a.out || a.out_f32) && a.T > 0, "kair_split_terms: null pointer / empty input");
  KB_CHECK(a.n > 0 && a.n % 8 == 0 && a.x_pitch % 4 == 0 && a.x_pitch >= a.n, "kair_split_terms: n=%d must be a "
           "multiple of 8 and fit the row pitch", a.n);
  KB_CHECK(a.terms == 1 || a.terms == 3 || a.terms == 6, "kair_split_terms: terms=%d must be 1, 3 or 6", a.terms);
  KB_CHECK(!a.out || (a.out_pitch % 8 == 0 && a.out_pitch >= (long long)a.terms * a.n),
           "kair_split_terms: out pitch must hold terms*n columns");
  KB_CHECK(a.op >= KAIR_SPLIT_COPY && a.op <= KAIR_SPLIT_LRELU, "kair_split_terms: op=%d", a.op);
  if (a.op == KAIR_SPLIT_DGELU || a.op == KAIR_SPLIT_ACT_BWD)
    KB_CHECK(a.aux && a.aux_pitch % 4 == 0, "kair_split_terms: op %d needs aux", a.op);
  KB_CHECK(!a.row_scale || a.rows_per_sample > 0, "kair_split_terms: row_scale needs rows_per_sample");
  KB_CHECK(!a.out_f32 || a.out_f32_pitch % 4 == 0, "kair_split_terms: fp32 output pitch");
  const long long total = a.T * (a.n >> 3);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  if (a.terms == 6) split_terms_kernel<6><<<(unsigned)blocks, 256, 0, st>>>(a);
  else if (a.terms == 3) split_terms_kernel<3><<<(unsigned)blocks, 256, 0, st>>>(a);
  else split_terms_kernel<1><<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int ln_bwd_split_launch(const kair_ln_bwd_split_args& a, cudaStream_t st) {
  KB_CHECK(a.g && a.xhat && a.rstd && a.outF && a.T > 0, "kair_ln_bwd_split: null pointer / empty input");
  KB_CHECK(a.C > 0 && a.pitch >= a.C && a.pitch <= 256, "kair_ln_bwd_split: C=%d pitch=%d (max 256)", a.C, a.pitch);
  KB_CHECK(a.terms == 3 || a.terms == 6, "kair_ln_bwd_split: terms=%d must be 3 or 6", a.terms);
  KB_CHECK(a.xhat_pitch >= (long long)a.terms * a.pitch, "kair_ln_bwd_split: x-hat pitch must hold terms*pitch columns");
  KB_CHECK(!a.out || a.out_pitch >= (long long)a.terms * a.pitch, "kair_ln_bwd_split: out pitch");
  KB_CHECK(!a.row_scale || a.rows_per_sample > 0, "kair_ln_bwd_split: row_scale needs rows_per_sample");
  long long blocks = (a.T + 7) / 8;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (a.terms == 6) ln_bwd_split_kernel<6><<<(unsigned)blocks, 256, 0, st>>>(a);
  else ln_bwd_split_kernel<3><<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kb
