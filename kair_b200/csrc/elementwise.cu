// Memory-bound helper kernels around the tensor-core path (warp-shuffle reductions, vectorised coalesced access):
//   * weight packing: fp32 master weights -> padded bf16 GEMM operands (LayerNorm gamma/beta and the q scale folded
//     in), and the inverse mapping of the packed weight gradients back onto the parameters' own layouts
//   * standalone LayerNorm forward / backward (patch_embed.norm and the final norm of SwinIR)
//   * the thin first / last convolutions (1..4 image channels <-> C feature channels), forward and backward
//   * residual-stream helpers
// Reference arithmetic: models/network_swinir.py:199,205,683,725 (LayerNorm eps=1e-5), :669 (conv_first),
// :762 (conv_last), :809-810,:837 (mean / img_range), models/network_dncnn.py:63-71, models/network_unet.py:25,57.
#include <string.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

// ------------------------------------------------------------------------------------------------ pack linear
__global__ void pack_linear_kernel(const kair_pack_linear_args a) {
  const int r = blockIdx.x, l = blockIdx.y;
  const float* W = reinterpret_cast<const float*>(a.W[l]);
  const float* b = a.b ? reinterpret_cast<const float*>(a.b[l]) : nullptr;
  const float* gamma = a.gamma ? reinterpret_cast<const float*>(a.gamma[l]) : nullptr;
  const float* beta = a.beta ? reinterpret_cast<const float*>(a.beta[l]) : nullptr;
  const int src = a.row_map[r];
  const float rs = a.row_scale ? a.row_scale[r] : 1.0f;
  const int terms = a.terms > 1 ? a.terms : 1;   // SPLIT PRECISION: weight-side segments along K (Wp) / N (WpT)
  bf16* Wp = reinterpret_cast<bf16*>(a.Wp) + ((size_t)l * a.N_pad + r) * a.K_pad * terms;
  bf16* WpT = a.WpT ? reinterpret_cast<bf16*>(a.WpT) + (size_t)l * a.K_pad * a.N_pad * terms : nullptr;
  float bacc = 0.f;
  for (int c = threadIdx.x; c < a.K_pad; c += blockDim.x) {
    const int k = a.col_map[c];
    float v = 0.f;
    if (src >= 0 && k >= 0) {
      const float w = W[(size_t)src * a.in_f + k];
      v = w * rs * (gamma ? gamma[k] : 1.0f);
      if (beta) bacc += w * beta[k];
    }
    if (terms == 1) {
      Wp[c] = __float2bfloat16(v);
      if (WpT) WpT[(size_t)c * a.N_pad + r] = __float2bfloat16(v);
    } else {
      for (int sg = 0; sg < terms; ++sg) {
        const bf16 comp = __float2bfloat16(split_pick(v, split_seg_w(terms, sg)));
        Wp[(size_t)sg * a.K_pad + c] = comp;
        if (WpT) WpT[((size_t)c * terms + sg) * a.N_pad + r] = comp;
      }
    }
  }
  if (a.bp) {
    __shared__ float red[32];
    bacc = warp_sum(bacc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = bacc;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < (blockDim.x + 31) / 32; ++i) t += red[i];
      float bias = 0.f;
      if (src >= 0) bias = ((b ? b[src] : 0.f) + t) * rs;
      if (a.bias_const) bias += a.bias_const[r];
      a.bp[(size_t)l * a.N_pad + r] = bias;
    }
  }
}

int pack_linear_launch(const kair_pack_linear_args& a, cudaStream_t st) {
  KB_CHECK(a.W && a.row_map && a.col_map && a.Wp, "kair_pack_linear: null pointer");
  KB_CHECK(a.L > 0 && a.N_pad > 0 && a.K_pad > 0, "kair_pack_linear: bad sizes");
  KB_CHECK(a.terms == 0 || a.terms == 1 || a.terms == 3 || a.terms == 6, "kair_pack_linear: terms=%d", a.terms);
  pack_linear_kernel<<<dim3(a.N_pad, a.L), 128, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ unpack linear grads
// dWp = d(loss)/d(Wp) with Wp[r][c] = rs[r] * W[row_map[r]][col_map[c]] * gamma[col_map[c]] and
// bp[r] = rs[r] * (b[src] + sum_k W[src][k] beta[k]); column `ones_col` of dWp is d(loss)/d(bp).
__global__ void unpack_linear_dw_kernel(const kair_unpack_linear_args a) {
  const int r = blockIdx.x, l = blockIdx.y;
  const int src = a.row_map[r];
  if (src < 0) return;
  const float rs = a.row_scale ? a.row_scale[r] : 1.0f;
  const float* dWp = a.dWp + ((size_t)l * a.N_pad + r) * a.K_pad;
  const float* gamma = a.gamma ? reinterpret_cast<const float*>(a.gamma[l]) : nullptr;
  const float* beta = a.beta ? reinterpret_cast<const float*>(a.beta[l]) : nullptr;
  float* dW = reinterpret_cast<float*>(a.dW[l]) + (size_t)src * a.in_f;
  // W enters twice: Wp = rs * W * gamma and bp = rs * (b + W . beta)
  const float dbp = (beta && a.ones_col >= 0) ? dWp[a.ones_col] : 0.f;
  for (int c = threadIdx.x; c < a.K_pad; c += blockDim.x) {
    const int k = a.col_map[c];
    if (k >= 0) dW[k] = rs * (dWp[c] * (gamma ? gamma[k] : 1.0f) + (beta ? dbp * beta[k] : 0.f));
  }
  if (threadIdx.x == 0 && a.db && a.ones_col >= 0) reinterpret_cast<float*>(a.db[l])[src] = dWp[a.ones_col] * rs;
}

__global__ void __launch_bounds__(1024) unpack_linear_dln_kernel(const kair_unpack_linear_args a) {
  // block = 32 columns x 32 row groups; dgamma[k] = sum_r rs*W[src][k] * dWp[r][c], dbeta[k] = sum_r rs*W[src][k] * dbp[r]
  // (8 row groups made this a 72-deep chain of dependent global loads on a 36-block grid: ~20 us per launch)
  __shared__ float sg[32][33], sb[32][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5, l = blockIdx.y;
  const int c = blockIdx.x * 32 + cx;
  const int k = c < a.K_pad ? a.col_map[c] : -1;
  const float* W = reinterpret_cast<const float*>(a.W[l]);
  const float* dWp = a.dWp + (size_t)l * a.N_pad * a.K_pad;
  float dg = 0.f, db = 0.f;
  if (k >= 0) {
    for (int r = ry; r < a.N_pad; r += 32) {
      const int src = a.row_map[r];
      if (src < 0) continue;
      const float rs = a.row_scale ? a.row_scale[r] : 1.0f;
      const float w = W[(size_t)src * a.in_f + k] * rs;
      dg += w * dWp[(size_t)r * a.K_pad + c];
      db += w * dWp[(size_t)r * a.K_pad + a.ones_col];
    }
  }
  sg[ry][cx] = dg;
  sb[ry][cx] = db;
  __syncthreads();
  // warp ry reduces column ry: lane i holds row group i of that column
  float g = sg[cx][ry], b = sb[cx][ry];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    g += __shfl_xor_sync(0xffffffffu, g, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  const int c2 = blockIdx.x * 32 + ry;
  const int k2 = c2 < a.K_pad ? a.col_map[c2] : -1;
  if (cx == 0 && k2 >= 0) {
    reinterpret_cast<float*>(a.dgamma[l])[k2] = g;
    reinterpret_cast<float*>(a.dbeta[l])[k2] = b;
  }
}

int unpack_linear_launch(const kair_unpack_linear_args& a, cudaStream_t st) {
  KB_CHECK(a.dWp && a.dW && a.row_map && a.col_map, "kair_unpack_linear_grad: null pointer");
  unpack_linear_dw_kernel<<<dim3(a.N_pad, a.L), 128, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  if (a.dgamma && a.dbeta) {
    KB_CHECK(a.W && a.ones_col >= 0, "kair_unpack_linear_grad: LayerNorm gradients need W and ones_col");
    unpack_linear_dln_kernel<<<dim3((a.K_pad + 31) / 32, a.L), 1024, 0, st>>>(a);
    KB_CUDA(cudaGetLastError());
  }
  return 0;
}

// ------------------------------------------------------------------------------------------------ pack conv3x3
// W (O, I, kh, kw) fp32 -> Wp [O_pad][taps*I_pad] (k = tap*I_pad + i) and, for the input-gradient convolution,
// WpT [I_pad][taps*O_pad] with WpT[i][tap*O_pad + o] = W[o][i][taps-1-tap] (spatially flipped, channels swapped).
// A block stages a 32(o) x 32(i) x taps tile in shared memory: the fp32 source is read in contiguous runs of 32*taps
// floats, both packed layouts are written in contiguous 64-byte runs (the per-element version read with a 36-byte
// stride and scattered 2-byte writes into WpT: 0.38 ms per step for DRUNet's 32.6 M weights, re-run after every
// optimizer step).
constexpr int PC_T = 32;
constexpr int PC_MAXTAPS = 9;

template <int TAPS>   // 0: run-time tap count (index arithmetic divides by it); 1 / 4 / 9: compile-time
__global__ void __launch_bounds__(256) pack_conv_kernel(const kair_pack_conv_args a) {
  __shared__ float tile[PC_T][PC_T * (TAPS ? TAPS : PC_MAXTAPS) + 1];
  const int l = blockIdx.z;
  const float* W = reinterpret_cast<const float*>(a.W[l]);
  const int taps = TAPS ? TAPS : a.taps;
  const int terms = a.terms > 1 ? a.terms : 1;   // SPLIT PRECISION: weight-side segments inside every tap
  const size_t n = (size_t)a.O_pad * taps * a.I_pad * terms;
  bf16* Wp = reinterpret_cast<bf16*>(a.Wp) + (size_t)l * n;
  bf16* WpT = a.WpT ? reinterpret_cast<bf16*>(a.WpT) + (size_t)l * n : nullptr;
  const float* osc = a.oscale ? reinterpret_cast<const float*>(a.oscale[l]) : nullptr;
  const float* osh = a.oshift ? reinterpret_cast<const float*>(a.oshift[l]) : nullptr;
  const int i0 = blockIdx.x * PC_T, o0 = blockIdx.y * PC_T;
  const int run = PC_T * taps;
  for (int idx = threadIdx.x; idx < PC_T * run; idx += blockDim.x) {
    const int ol = idx / run, r = idx - ol * run;
    const int o = o0 + ol, i = i0 + r / taps;
    float v = 0.f;
    if (o < a.O && i < a.I) v = W[((size_t)o * a.I + i0) * taps + r] * (osc ? osc[o] : 1.f);
    tile[ol][r] = v;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < PC_T * run; idx += blockDim.x) {
    const int il = idx % PC_T, tap = (idx / PC_T) % taps, ol = idx / (PC_T * taps);
    const int o = o0 + ol, i = i0 + il;
    if (o < a.O_pad && i < a.I_pad) {
      const float v = tile[ol][il * taps + tap];
      if (terms == 1) {
        Wp[((size_t)o * taps + tap) * a.I_pad + i] = __float2bfloat16(v);
      } else {
        for (int sg = 0; sg < terms; ++sg)
          Wp[(((size_t)o * taps + tap) * terms + sg) * a.I_pad + i] = __float2bfloat16(split_pick(v, split_seg_w(terms, sg)));
      }
    }
  }
  if (WpT) {
    for (int idx = threadIdx.x; idx < PC_T * run; idx += blockDim.x) {
      const int ol = idx % PC_T, tap = (idx / PC_T) % taps, il = idx / (PC_T * taps);
      const int o = o0 + ol, i = i0 + il;
      if (o < a.O_pad && i < a.I_pad) {
        const float v = tile[ol][il * taps + tap];
        if (terms == 1) {
          WpT[((size_t)i * taps + (taps - 1 - tap)) * a.O_pad + o] = __float2bfloat16(v);
        } else {
          for (int sg = 0; sg < terms; ++sg)
            WpT[(((size_t)i * taps + (taps - 1 - tap)) * terms + sg) * a.O_pad + o] =
                __float2bfloat16(split_pick(v, split_seg_w(terms, sg)));
        }
      }
    }
  }
  if (a.bp && blockIdx.x == 0 && blockIdx.y == 0) {
    const float* b = a.b ? reinterpret_cast<const float*>(a.b[l]) : nullptr;
    for (int o = threadIdx.x; o < a.O_pad; o += blockDim.x) {
      float v = (b && o < a.O) ? b[o] : 0.f;
      if (o < a.O) v = v * (osc ? osc[o] : 1.f) + (osh ? osh[o] : 0.f);
      a.bp[(size_t)l * a.O_pad + o] = v;
    }
  }
}

int pack_conv_launch(const kair_pack_conv_args& a, cudaStream_t st) {
  KB_CHECK(a.W && a.Wp, "kair_pack_conv: null pointer");
  KB_CHECK(a.taps >= 1 && a.taps <= PC_MAXTAPS, "kair_pack_conv: taps=%d must be in 1..%d", a.taps, PC_MAXTAPS);
  KB_CHECK(a.L >= 1 && a.L <= 65535, "kair_pack_conv: L=%d layers per launch", a.L);
  KB_CHECK(a.terms == 0 || a.terms == 1 || a.terms == 3 || a.terms == 6, "kair_pack_conv: terms=%d", a.terms);
  dim3 grid((a.I_pad + PC_T - 1) / PC_T, (a.O_pad + PC_T - 1) / PC_T, a.L);
  if (a.taps == 9) pack_conv_kernel<9><<<grid, 256, 0, st>>>(a);
  else if (a.taps == 4) pack_conv_kernel<4><<<grid, 256, 0, st>>>(a);
  else if (a.taps == 1) pack_conv_kernel<1><<<grid, 256, 0, st>>>(a);
  else pack_conv_kernel<0><<<grid, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

template <int TAPS>   // 0: run-time tap count; 1 / 4 / 9: compile-time (the index chain divides by it)
__global__ void unpack_conv_kernel(const kair_unpack_conv_args a) {
  const int l = blockIdx.y;
  const unsigned taps = TAPS ? TAPS : a.taps;
  const int rp = a.dwp_pitch > 0 ? a.dwp_pitch : (a.transposed ? a.O_pad : (int)taps * a.I_pad);  // row pitch
  const float* dWp = a.dWp + (size_t)l * (a.transposed ? (size_t)taps * a.I_pad : (size_t)a.O_pad) * rp;
  float* dW = reinterpret_cast<float*>(a.dW[l]);
  // one layer's (O, I, taps) count fits 32 bits (launcher check): 32-bit div/mod chain, coalesced writes, the strided
  // reads of a warp (taps rows x ~4 consecutive inputs) are served by L1
  const unsigned n = (unsigned)a.O * (unsigned)a.I * taps, I = (unsigned)a.I;
  for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
    const unsigned oi = e / taps, tap = e - oi * taps;
    const unsigned o = oi / I, i = oi - o * I;
    dW[e] = a.transposed ? __ldg(dWp + (size_t)(tap * a.I_pad + i) * rp + o) : __ldg(dWp + (size_t)o * rp + tap * a.I_pad + i);
  }
  if (a.db && a.ones_col >= 0 && blockIdx.x == 0) {
    float* db = reinterpret_cast<float*>(a.db[l]);
    for (int o = threadIdx.x; o < a.O; o += blockDim.x)
      db[o] = dWp[(size_t)o * rp + a.center_tap * a.I_pad + a.ones_col];
  }
}

int unpack_conv_launch(const kair_unpack_conv_args& a, cudaStream_t st) {
  KB_CHECK(a.dWp && a.dW, "kair_unpack_conv_grad: null pointer");
  KB_CHECK(!a.transposed || !a.db, "kair_unpack_conv_grad: the transposed layout carries no bias column");
  size_t n = (size_t)a.O * a.I * a.taps;
  KB_CHECK(a.taps >= 1 && n < (1ull << 31), "kair_unpack_conv_grad: a layer's weight count must fit 31 bits");
  int blocks = (int)((n + 255) / 256);
  if (blocks > 1024) blocks = 1024;
  const dim3 grid(blocks, a.L);
  if (a.taps == 9) unpack_conv_kernel<9><<<grid, 256, 0, st>>>(a);
  else if (a.taps == 4) unpack_conv_kernel<4><<<grid, 256, 0, st>>>(a);
  else if (a.taps == 1) unpack_conv_kernel<1><<<grid, 256, 0, st>>>(a);
  else unpack_conv_kernel<0><<<grid, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ LayerNorm
// One warp per token row.  x fp32 [T][pitch]; C real channels, padded channels of every output are written as 0
// (xhat gets 1.0 in `ones_col`).
__global__ void layernorm_fwd_kernel(const kair_layernorm_args a) {
  const int warps = blockDim.x >> 5;
  const long long row = (long long)blockIdx.x * warps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= a.T) return;
  const float* x = a.x + row * a.pitch;
  float v[8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane + 32 * i;
    v[i] = (c < a.C) ? x[c] : 0.f;
    s += v[i];
  }
  const float mean = warp_sum(s) / a.C;
  float s2 = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane + 32 * i;
    const float d = (c < a.C) ? v[i] - mean : 0.f;
    s2 += d * d;
  }
  const float rstd = rsqrtf(warp_sum(s2) / a.C + a.eps);
  if (lane == 0) {
    if (a.mean) a.mean[row] = mean;
    if (a.rstd) a.rstd[row] = rstd;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane + 32 * i;
    if (c >= a.pitch) continue;
    const float xh = (c < a.C) ? (v[i] - mean) * rstd : 0.f;
    const float y = (c < a.C && a.gamma) ? xh * a.gamma[c] + a.beta[c] : 0.f;
    if (a.y_f32) a.y_f32[row * a.pitch + c] = y;
    if (a.terms > 1) {   // SPLIT PRECISION: bf16 outputs are split operands [T][terms*pitch]
      const size_t base = (size_t)row * a.pitch * a.terms + c;
      for (int sg = 0; sg < a.terms; ++sg) {
        const int w = split_seg_act(a.terms, sg);
        if (a.y_bf16)
          reinterpret_cast<bf16*>(a.y_bf16)[base + (size_t)sg * a.pitch] =
              __float2bfloat16(split_pick(c == a.ones_col ? 1.0f : y, w));
        if (a.xhat_bf16)
          reinterpret_cast<bf16*>(a.xhat_bf16)[base + (size_t)sg * a.pitch] =
              __float2bfloat16(split_pick(c == a.ones_col ? 1.0f : xh, w));
      }
      continue;
    }
    if (a.y_bf16) reinterpret_cast<bf16*>(a.y_bf16)[row * a.pitch + c] = __float2bfloat16(c == a.ones_col ? 1.0f : y);
    if (a.xhat_bf16)
      reinterpret_cast<bf16*>(a.xhat_bf16)[row * a.pitch + c] = __float2bfloat16(c == a.ones_col ? 1.0f : xh);
  }
}

// Fast path (terms <= 1, even pitch): a lane owns column PAIRS c = 2 * lane + 64 * i, so every warp access is one
// 256-byte (fp32) or 128-byte (bf16) contiguous run instead of 128 / 64 bytes, and a warp keeps two rows in flight.
// Same arithmetic and summation order per row as layernorm_fwd_kernel up to the pairing of the partial sums.
__global__ void __launch_bounds__(256) layernorm_fwd_pairs_kernel(const kair_layernorm_args a) {
  const int warps = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nv = (a.pitch + 63) >> 6;   // float2 per lane and row (<= 4)
  const long long row0 = ((long long)blockIdx.x * warps + (threadIdx.x >> 5)) * 2;
  float2 v[2][4];
  float s[2] = {0.f, 0.f};
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const long long row = row0 + r;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = 2 * lane + 64 * i;
      v[r][i] = make_float2(0.f, 0.f);
      if (i < nv && row < a.T && c < a.C) {
        const float2 t = *reinterpret_cast<const float2*>(a.x + row * a.pitch + c);
        v[r][i].x = t.x;
        v[r][i].y = (c + 1 < a.C) ? t.y : 0.f;
      }
      s[r] += v[r][i].x + v[r][i].y;
    }
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const long long row = row0 + r;
    if (row >= a.T) return;   // rows come in order: nothing follows an absent row (warp-uniform)
    const float mean = warp_sum(s[r]) / a.C;
    float s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = 2 * lane + 64 * i;
      const float d0 = (i < nv && c < a.C) ? v[r][i].x - mean : 0.f;
      const float d1 = (i < nv && c + 1 < a.C) ? v[r][i].y - mean : 0.f;
      s2 += d0 * d0 + d1 * d1;
    }
    const float rstd = rsqrtf(warp_sum(s2) / a.C + a.eps);
    if (lane == 0) {
      if (a.mean) a.mean[row] = mean;
      if (a.rstd) a.rstd[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = 2 * lane + 64 * i;
      if (i >= nv || c >= a.pitch) continue;
      float xh[2], y[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int ce = c + e;
        const float xv = e ? v[r][i].y : v[r][i].x;
        xh[e] = (ce < a.C) ? (xv - mean) * rstd : 0.f;
        y[e] = (ce < a.C && a.gamma) ? xh[e] * a.gamma[ce] + a.beta[ce] : 0.f;
      }
      if (a.y_f32) *reinterpret_cast<float2*>(a.y_f32 + row * a.pitch + c) = make_float2(y[0], y[1]);
      if (a.y_bf16)
        *reinterpret_cast<uint32_t*>(reinterpret_cast<bf16*>(a.y_bf16) + row * a.pitch + c) =
            pack_bf16x2(c == a.ones_col ? 1.0f : y[0], c + 1 == a.ones_col ? 1.0f : y[1]);
      if (a.xhat_bf16)
        *reinterpret_cast<uint32_t*>(reinterpret_cast<bf16*>(a.xhat_bf16) + row * a.pitch + c) =
            pack_bf16x2(c == a.ones_col ? 1.0f : xh[0], c + 1 == a.ones_col ? 1.0f : xh[1]);
    }
  }
}

int layernorm_fwd_launch(const kair_layernorm_args& a, cudaStream_t st) {
  KB_CHECK(a.x && a.T > 0 && a.C > 0 && a.pitch >= a.C && a.pitch <= 256, "kair_layernorm_fwd: bad arguments");
  KB_CHECK((a.y_f32 == nullptr && a.y_bf16 == nullptr) || (a.gamma && a.beta), "kair_layernorm_fwd: affine output needs gamma/beta");
  KB_CHECK(a.terms == 0 || a.terms == 1 || a.terms == 3 || a.terms == 6, "kair_layernorm_fwd: terms=%d", a.terms);
  const int warps = 8;
  const bool aligned = (reinterpret_cast<uintptr_t>(a.x) % 8 == 0) && (!a.y_f32 || reinterpret_cast<uintptr_t>(a.y_f32) % 8 == 0) &&
                       (!a.y_bf16 || reinterpret_cast<uintptr_t>(a.y_bf16) % 4 == 0) &&
                       (!a.xhat_bf16 || reinterpret_cast<uintptr_t>(a.xhat_bf16) % 4 == 0);
  if (a.terms <= 1 && a.pitch % 2 == 0 && aligned) {
    layernorm_fwd_pairs_kernel<<<(unsigned)((a.T + 2 * warps - 1) / (2 * warps)), warps * 32, 0, st>>>(a);
    KB_CUDA(cudaGetLastError());
    return 0;
  }
  layernorm_fwd_kernel<<<(unsigned)((a.T + warps - 1) / warps), warps * 32, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)),  g = dy * gamma;  dgamma += dy * xhat;  dbeta += dy
// dy is given as fp32 and/or bf16 (summed when both are present); dx_out = dx (+ dx_add).
__global__ void layernorm_bwd_kernel(const kair_layernorm_bwd_args a) {
  // one warp per row; lane owns column pairs c = 2*lane + 64*i (8-byte coalesced accesses), pitch <= 256, pitch % 64 == 0
  __shared__ float sg[256], sb[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    sg[i] = 0.f;
    sb[i] = 0.f;
  }
  __syncthreads();
  const int warps = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nv = a.pitch >> 6;  // float2 per lane
  const float inv_c = 1.0f / (float)a.C;
  float2 pg[4], pb[4], gam[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    pg[i] = pb[i] = make_float2(0.f, 0.f);
    const int c = 2 * lane + 64 * i;
    gam[i].x = (i < nv && c < a.C) ? a.gamma[c] : 0.f;
    gam[i].y = (i < nv && c + 1 < a.C) ? a.gamma[c + 1] : 0.f;
  }
  const bf16* dyb = reinterpret_cast<const bf16*>(a.dy_bf16);
  bf16* dxb = reinterpret_cast<bf16*>(a.dx_bf16);
  for (long long row = (long long)blockIdx.x * warps + (threadIdx.x >> 5); row < a.T;
       row += (long long)gridDim.x * warps) {
    const size_t base = (size_t)row * a.pitch;
    float2 dy[4], xv[4], ad[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (i < nv) {
        const int c = 2 * lane + 64 * i;
        xv[i] = *reinterpret_cast<const float2*>(a.x + base + c);
        dy[i] = a.dy_f32 ? *reinterpret_cast<const float2*>(a.dy_f32 + base + c) : make_float2(0.f, 0.f);
        if (dyb) {
          float2 t = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(dyb + base + c));
          dy[i].x += t.x;
          dy[i].y += t.y;
        }
        ad[i] = a.dx_add ? *reinterpret_cast<const float2*>(a.dx_add + base + c) : make_float2(0.f, 0.f);
      }
    }
    const float mean = a.mean[row], rstd = a.rstd[row];
    float s1 = 0.f, s2 = 0.f;
    float2 xh[4], g[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      xh[i] = g[i] = make_float2(0.f, 0.f);
      if (i < nv) {
        const int c = 2 * lane + 64 * i;
        if (c < a.C) {
          xh[i].x = (xv[i].x - mean) * rstd;
          pg[i].x += dy[i].x * xh[i].x;
          pb[i].x += dy[i].x;
          g[i].x = dy[i].x * gam[i].x;
        }
        if (c + 1 < a.C) {
          xh[i].y = (xv[i].y - mean) * rstd;
          pg[i].y += dy[i].y * xh[i].y;
          pb[i].y += dy[i].y;
          g[i].y = dy[i].y * gam[i].y;
        }
        s1 += g[i].x + g[i].y;
        s2 += g[i].x * xh[i].x + g[i].y * xh[i].y;
      }
    }
    const float m1 = warp_sum(s1) * inv_c, m2 = warp_sum(s2) * inv_c;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (i < nv) {
        const int c = 2 * lane + 64 * i;
        float2 dx;
        dx.x = (c < a.C) ? rstd * (g[i].x - m1 - xh[i].x * m2) + ad[i].x : 0.f;
        dx.y = (c + 1 < a.C) ? rstd * (g[i].y - m1 - xh[i].y * m2) + ad[i].y : 0.f;
        if (a.dx_f32) *reinterpret_cast<float2*>(a.dx_f32 + base + c) = dx;
        if (dxb) *reinterpret_cast<__nv_bfloat162*>(dxb + base + c) = __floats2bfloat162_rn(dx.x, dx.y);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = 2 * lane + 64 * i;
    if (i < nv && c < a.C) {
      atomicAdd(&sg[c], pg[i].x);
      atomicAdd(&sb[c], pb[i].x);
      if (c + 1 < a.C) {
        atomicAdd(&sg[c + 1], pg[i].y);
        atomicAdd(&sb[c + 1], pb[i].y);
      }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < a.C; c += blockDim.x) {
    atomicAdd(&a.dgamma[c], sg[c]);
    atomicAdd(&a.dbeta[c], sb[c]);
  }
}

int layernorm_bwd_launch(const kair_layernorm_bwd_args& a, cudaStream_t st) {
  KB_CHECK(a.x && a.mean && a.rstd && a.gamma && a.dgamma && a.dbeta && (a.dy_f32 || a.dy_bf16),
           "kair_layernorm_bwd: null pointer");
  KB_CHECK(a.pitch <= 256 && a.pitch >= a.C && a.pitch % 64 == 0, "kair_layernorm_bwd: pitch must be a multiple of 64, <= 256");
  int blocks = sm_count() * 8;
  long long need = (a.T + 7) / 8;
  if (blocks > need) blocks = (int)need;
  layernorm_bwd_kernel<<<blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ thin convolutions
constexpr int CI_MAX = 4;  // image-side channel count handled by the SIMT convolutions
constexpr int CI_PIX = 4;  // conv_in: pixels (along a row) per thread

// conv_in: NCHW fp32 image (B, ci, H, W) -> token features [T][pitch]; one thread per (pixel, 8 consecutive outputs):
// every thread is independent (no reductions), its 9*ci image loads are issued together, the weights are read from
// shared memory as float4 pairs and the result leaves as one 16-byte bf16 / two 16-byte fp32 stores.
//   feat[t][o] = act( b[o] + sum_{c,tap} W[o][c][tap] * ((img[c][p+tap] - mean[c]) * range) )
// CI: image channels as a compile-time constant (the 9*CI tap loads / index chains are the issue budget);
// idx_t: 32-bit element index whenever T * pitch/8 fits (the 64-bit div/mod chain cost more than the 72 FMAs)
template <int CI, typename idx_t>
__global__ void conv_in_fwd_kernel(const kair_conv_in_args a) {
  extern __shared__ float sw[];  // [ci*9][P8] weights, then [P8] bias; P8 = pitch rounded to 8
  const int K = CI * 9;
  const int P8 = a.pitch;        // launcher guarantees pitch % 8 == 0
  for (int e = threadIdx.x; e < K * P8; e += blockDim.x) {
    const int o = e % P8, k = e / P8;
    sw[e] = o < a.O ? a.W[(size_t)o * K + k] : 0.f;
  }
  float* sbias = sw + K * P8;
  for (int o = threadIdx.x; o < P8; o += blockDim.x) sbias[o] = (o < a.O && a.b) ? a.b[o] : 0.f;
  __syncthreads();
  // one thread = 8 consecutive outputs of CI_PIX consecutive pixels of a row: the 18 float4 weight reads per tap set are
  // amortised over four pixels (one pixel per thread was bound by the shared-memory bandwidth of those reads: every
  // warp-wide LDS.128 is 8 wavefronts), and the 3 x (CI_PIX + 2) image window is loaded once
  const idx_t og = (idx_t)(P8 >> 3);  // output groups per pixel
  const idx_t Wq = (idx_t)((a.Wd + CI_PIX - 1) / CI_PIX);
  const idx_t rows = (idx_t)a.batch * (idx_t)a.H;
  const idx_t n = rows * Wq * og;
  for (idx_t e = (idx_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (idx_t)gridDim.x * blockDim.x) {
    const idx_t tq = e / og;
    const int g = (int)(e - tq * og);
    const idx_t row = tq / Wq;
    const int w0 = (int)(tq - row * Wq) * CI_PIX;
    const int b = (int)(row / (idx_t)a.H), h = (int)(row - (idx_t)b * (idx_t)a.H);
    float in[CI][3][CI_PIX + 2];
#pragma unroll
    for (int c = 0; c < CI; ++c) {
#pragma unroll
      for (int dy = 0; dy < 3; ++dy) {
        const int hh = h + dy - 1;
#pragma unroll
        for (int x = 0; x < CI_PIX + 2; ++x) {
          const int ww = w0 + x - 1;
          float v = 0.f;
          if (hh >= 0 && hh < a.H && ww >= 0 && ww < a.Wd) {
            v = __ldg(a.img + (((size_t)b * CI + c) * a.H + hh) * a.Wd + ww);
            v = (v - a.mean[c]) * a.range;
          }
          in[c][dy][x] = v;
        }
      }
    }
    float acc[CI_PIX][8];
    {
      const float4 b0 = *reinterpret_cast<const float4*>(sbias + 8 * g);
      const float4 b1 = *reinterpret_cast<const float4*>(sbias + 8 * g + 4);
#pragma unroll
      for (int j = 0; j < CI_PIX; ++j) {
        acc[j][0] = b0.x; acc[j][1] = b0.y; acc[j][2] = b0.z; acc[j][3] = b0.w;
        acc[j][4] = b1.x; acc[j][5] = b1.y; acc[j][6] = b1.z; acc[j][7] = b1.w;
      }
    }
    // same accumulation order per output as before: channels outer, taps in (dy, dx) order
#pragma unroll
    for (int c = 0; c < CI; ++c) {
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const float* wr = sw + (c * 9 + tap) * P8 + 8 * g;
        const float4 w0v = *reinterpret_cast<const float4*>(wr);
        const float4 w1v = *reinterpret_cast<const float4*>(wr + 4);
#pragma unroll
        for (int j = 0; j < CI_PIX; ++j) {
          const float x = in[c][tap / 3][j + tap % 3];
          acc[j][0] += w0v.x * x; acc[j][1] += w0v.y * x; acc[j][2] += w0v.z * x; acc[j][3] += w0v.w * x;
          acc[j][4] += w1v.x * x; acc[j][5] += w1v.y * x; acc[j][6] += w1v.z * x; acc[j][7] += w1v.w * x;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < CI_PIX; ++j) {
      if (w0 + j >= a.Wd) break;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int o = 8 * g + i;
        if (o >= a.O) acc[j][i] = 0.f;
        else if (a.act == KAIR_ACT_LRELU) acc[j][i] = acc[j][i] > 0.f ? acc[j][i] : acc[j][i] * a.act_slope;
      }
      const long long off = ((long long)row * a.Wd + w0 + j) * a.pitch + 8 * g;
      if (a.out_f32) {
        reinterpret_cast<float4*>(a.out_f32 + off)[0] = make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
        reinterpret_cast<float4*>(a.out_f32 + off)[1] = make_float4(acc[j][4], acc[j][5], acc[j][6], acc[j][7]);
      }
      if (a.out_bf16) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (8 * g + i == a.ones_col) acc[j][i] = 1.0f;
        uint4 o4;
        o4.x = pack_bf16x2(acc[j][0], acc[j][1]); o4.y = pack_bf16x2(acc[j][2], acc[j][3]);
        o4.z = pack_bf16x2(acc[j][4], acc[j][5]); o4.w = pack_bf16x2(acc[j][6], acc[j][7]);
        *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(a.out_bf16) + off) = o4;
      }
    }
  }
}

int conv_in_fwd_launch(const kair_conv_in_args& a, cudaStream_t st) {
  KB_CHECK(a.img && a.W && a.mean, "kair_conv_in_fwd: null pointer");
  KB_CHECK(a.ci >= 1 && a.ci <= CI_MAX, "kair_conv_in_fwd: ci=%d unsupported (max %d)", a.ci, CI_MAX);
  KB_CHECK(a.O_pad >= a.O && a.pitch >= a.O && a.pitch % 8 == 0, "kair_conv_in_fwd: pitch must be a multiple of 8, >= O");
  size_t smem = ((size_t)a.ci * 9 * a.pitch + a.pitch) * 4;
  KB_CHECK(smem <= 48 * 1024, "kair_conv_in_fwd: weights do not fit in shared memory");
  const long long n = (long long)a.batch * a.H * ((a.Wd + CI_PIX - 1) / CI_PIX) * (a.pitch >> 3);
  long long blocks = (n + 255) / 256;
  if (blocks > (long long)sm_count() * 8) blocks = (long long)sm_count() * 8;
  const bool small = n + (long long)blocks * 256 < (1ll << 32);   // the grid-stride index stays inside 32 bits
#define KB_CONV_IN(CI_)                                                                       \
  if (small) conv_in_fwd_kernel<CI_, unsigned><<<(int)blocks, 256, smem, st>>>(a);            \
  else conv_in_fwd_kernel<CI_, long long><<<(int)blocks, 256, smem, st>>>(a)
  switch (a.ci) {
    case 1: KB_CONV_IN(1); break;
    case 2: KB_CONV_IN(2); break;
    case 3: KB_CONV_IN(3); break;
    default: KB_CONV_IN(4); break;
  }
#undef KB_CONV_IN
  KB_CUDA(cudaGetLastError());
  return 0;
}

// conv_in weight gradient: dW[o][c][tap] = sum_t dy[t][o] * in(t, c, tap), db[o] = sum_t dy[t][o]
__global__ void conv_in_wgrad_kernel(const kair_conv_in_args a) {
  const int K = a.ci * 9;
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  const long long T = (long long)a.batch * a.H * a.Wd;
  // each lane owns outputs o = lane + 32*j ; accumulators acc[j][k]
  constexpr int JMAX = 8;
  const int nj = (a.O + 31) / 32;
  for (int j0 = 0; j0 < nj; ++j0) {
    const int o = lane + 32 * j0;
    float acc[CI_MAX * 9 + 1];
#pragma unroll
    for (int k = 0; k <= CI_MAX * 9; ++k) acc[k] = 0.f;
    for (long long t = (long long)blockIdx.x * warps + (threadIdx.x >> 5); t < T; t += (long long)gridDim.x * warps) {
      const int w = (int)(t % a.Wd), h = (int)((t / a.Wd) % a.H), b = (int)(t / ((long long)a.Wd * a.H));
      float dy = 0.f;
      if (o < a.O) {
        if (a.dy_f32) dy += a.dy_f32[t * a.pitch + o];
        if (a.dy_bf16) dy += __bfloat162float(reinterpret_cast<const bf16*>(a.dy_bf16)[t * a.pitch + o]);
      }
#pragma unroll
      for (int c = 0; c < CI_MAX; ++c) {
        if (c >= a.ci) break;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const int hh = h + tap / 3 - 1, ww = w + tap % 3 - 1;
          float v = 0.f;
          if (hh >= 0 && hh < a.H && ww >= 0 && ww < a.Wd) {
            v = a.img[(((size_t)b * a.ci + c) * a.H + hh) * a.Wd + ww];
            v = (v - a.mean[c]) * a.range;
          }
          acc[c * 9 + tap] += dy * v;
        }
      }
      acc[CI_MAX * 9] += dy;
    }
    if (o < a.O) {
      for (int k = 0; k < K; ++k) atomicAdd(&a.dW[(size_t)o * K + k], acc[k]);
      if (a.db) atomicAdd(&a.db[o], acc[CI_MAX * 9]);
    }
  }
  (void)JMAX;
}

int conv_in_wgrad_launch(const kair_conv_in_args& a, cudaStream_t st) {
  KB_CHECK(a.img && a.dW && a.mean && (a.dy_f32 || a.dy_bf16), "kair_conv_in_wgrad: null pointer");
  KB_CHECK(a.ci >= 1 && a.ci <= CI_MAX, "kair_conv_in_wgrad: ci=%d unsupported", a.ci);
  conv_in_wgrad_kernel<<<sm_count() * 2, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// conv_out: token features bf16 [T][pitch] (C channels) -> NCHW fp32 image (B, co, H, W).
// 8 lanes per pixel (a warp covers 4 consecutive pixels): lane j of a pixel owns the 16-byte channel chunks
// j, j+8, ...; the 9 tap rows of a pixel are fetched with independent 16-byte loads issued before any arithmetic
// (memory-level parallelism instead of one dependent load chain per warp), partial sums meet in 3 shuffle steps.
//   out[c][p] = (b[c] + sum_{tap,k} W[c][k][tap] * feat[p+tap][k] + skip_scale * skip[c][p]) * out_scale + out_mean[c]
template <bool kF32>   // kF32: fp32 features (a.feat_f32), 8 channels = two 16-byte loads
__global__ void conv_out_fwd_kernel(const kair_conv_out_args a) {
  extern __shared__ float sw[];  // [co][9][C_pad]
  const int Cp = a.pitch;
  for (int e = threadIdx.x; e < a.co * 9 * Cp; e += blockDim.x) {
    const int k = e % Cp, tap = (e / Cp) % 9, c = e / (9 * Cp);
    sw[e] = k < a.C ? a.W[((size_t)c * a.C + k) * 9 + tap] : 0.f;
  }
  __syncthreads();
  const int sub = threadIdx.x & 7;
  const long long T = (long long)a.batch * a.H * a.Wd;
  const long long groups = (T + 3) / 4;  // 4 pixels per warp iteration
  const int warps = blockDim.x >> 5;
  const bf16* feat = reinterpret_cast<const bf16*>(a.feat);
  const int chunks = (a.C + 7) >> 3;     // 16-byte chunks per row that hold real channels
  for (long long gidx = (long long)blockIdx.x * warps + (threadIdx.x >> 5); gidx < groups; gidx += (long long)gridDim.x * warps) {
    const long long t = gidx * 4 + ((threadIdx.x & 31) >> 3);
    const bool live = t < T;
    const long long tt = live ? t : T - 1;
    const int w = (int)(tt % a.Wd), h = (int)((tt / a.Wd) % a.H), b = (int)(tt / ((long long)a.Wd * a.H));
    float acc[CI_MAX];
#pragma unroll
    for (int c = 0; c < CI_MAX; ++c) acc[c] = 0.f;
    for (int ch = sub; ch < chunks; ch += 8) {
      uint4 v[9], v2[9];
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int hh = h + tap / 3 - 1, ww = w + tap % 3 - 1;
        v[tap] = make_uint4(0u, 0u, 0u, 0u);
        v2[tap] = make_uint4(0u, 0u, 0u, 0u);
        if (hh >= 0 && hh < a.H && ww >= 0 && ww < a.Wd) {
          const size_t off = (((size_t)b * a.H + hh) * a.Wd + ww) * Cp + ch * 8;
          if (kF32) {
            v[tap] = __ldg(reinterpret_cast<const uint4*>(a.feat_f32 + off));
            v2[tap] = __ldg(reinterpret_cast<const uint4*>(a.feat_f32 + off + 4));
          } else {
            v[tap] = __ldg(reinterpret_cast<const uint4*>(feat + off));
          }
        }
      }
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const uint32_t wd[4] = {v[tap].x, v[tap].y, v[tap].z, v[tap].w};
        float f[8];
        if (kF32) {
          f[0] = __uint_as_float(v[tap].x); f[1] = __uint_as_float(v[tap].y);
          f[2] = __uint_as_float(v[tap].z); f[3] = __uint_as_float(v[tap].w);
          f[4] = __uint_as_float(v2[tap].x); f[5] = __uint_as_float(v2[tap].y);
          f[6] = __uint_as_float(v2[tap].z); f[7] = __uint_as_float(v2[tap].w);
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float2 f2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&wd[i]));
            f[2 * i] = f2.x;
            f[2 * i + 1] = f2.y;
          }
        }
#pragma unroll
        for (int c = 0; c < CI_MAX; ++c) {
          if (c >= a.co) break;
          const float* wr = sw + (c * 9 + tap) * Cp + ch * 8;
          const float4 w0 = *reinterpret_cast<const float4*>(wr);
          const float4 w1 = *reinterpret_cast<const float4*>(wr + 4);
          acc[c] += w0.x * f[0] + w0.y * f[1] + w0.z * f[2] + w0.w * f[3] + w1.x * f[4] + w1.y * f[5] + w1.z * f[6] +
                    w1.w * f[7];
        }
      }
    }
#pragma unroll
    for (int c = 0; c < CI_MAX; ++c) {
      if (c >= a.co) break;
      float v = acc[c];
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      if (sub == 0 && live) {
        v += a.b ? a.b[c] : 0.f;
        const size_t oi = (((size_t)b * a.co + c) * a.H + h) * a.Wd + w;
        if (a.skip) v += a.skip_scale * (a.skip[oi] - a.skip_mean[c]);
        a.out[oi] = v * a.out_scale + a.out_mean[c];
      }
    }
  }
}

// fallback of conv_thin.cu's tensor-core kernel (C > 192 or feature rows shorter than the k-padding)
int conv_out_fwd_simt_launch(const kair_conv_out_args& a, cudaStream_t st) {
  KB_CHECK((a.feat || a.feat_f32) && a.W && a.out && a.out_mean, "kair_conv_out_fwd: null pointer");
  KB_CHECK(a.co >= 1 && a.co <= CI_MAX, "kair_conv_out_fwd: co=%d unsupported (max %d)", a.co, CI_MAX);
  KB_CHECK(a.pitch % 8 == 0 && a.C <= a.pitch, "kair_conv_out_fwd: pitch must be a multiple of 8 channels, >= C");
  size_t smem = (size_t)a.co * 9 * a.pitch * 4;
  KB_CHECK(smem <= 96 * 1024, "kair_conv_out_fwd: weights do not fit in shared memory");
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(conv_out_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    KB_CUDA(cudaFuncSetAttribute(conv_out_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_set[dev & 63] = true;
  }
  const long long groups = ((long long)a.batch * a.H * a.Wd + 3) / 4;
  long long blocks = (groups + 7) / 8;
  const long long cap = (long long)sm_count() * (smem > 32 * 1024 ? 2 : 6);
  if (blocks > cap) blocks = cap;
  if (a.feat_f32) conv_out_fwd_kernel<true><<<(int)blocks, 256, smem, st>>>(a);
  else conv_out_fwd_kernel<false><<<(int)blocks, 256, smem, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// conv_out input gradient: dfeat[t][k] = sum_{c,tap} dout[c][p - off(tap)] * W[c][k][tap]   (dout already scaled)
__global__ void conv_out_dgrad_kernel(const kair_conv_out_args a) {
  extern __shared__ float sw[];  // [co][9][C_pad]
  const int Cp = a.pitch;
  for (int e = threadIdx.x; e < a.co * 9 * Cp; e += blockDim.x) {
    const int k = e % Cp, tap = (e / Cp) % 9, c = e / (9 * Cp);
    sw[e] = k < a.C ? a.W[((size_t)c * a.C + k) * 9 + tap] : 0.f;
  }
  __syncthreads();
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  const long long T = (long long)a.batch * a.H * a.Wd;
  for (long long t = (long long)blockIdx.x * warps + (threadIdx.x >> 5); t < T; t += (long long)gridDim.x * warps) {
    const int w = (int)(t % a.Wd), h = (int)((t / a.Wd) % a.H), b = (int)(t / ((long long)a.Wd * a.H));
    float g[CI_MAX * 9];
#pragma unroll
    for (int c = 0; c < CI_MAX; ++c) {
      if (c >= a.co) break;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int hh = h - (tap / 3 - 1), ww = w - (tap % 3 - 1);
        g[c * 9 + tap] = (hh >= 0 && hh < a.H && ww >= 0 && ww < a.Wd)
                             ? a.dout[(((size_t)b * a.co + c) * a.H + hh) * a.Wd + ww] * a.out_scale
                             : 0.f;
      }
    }
    for (int k = lane; k < Cp; k += 32) {
      float acc = 0.f;
      if (k < a.C) {
#pragma unroll
        for (int c = 0; c < CI_MAX; ++c) {
          if (c >= a.co) break;
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) acc += g[c * 9 + tap] * sw[(c * 9 + tap) * Cp + k];
        }
        if (a.mask && !(__bfloat162float(reinterpret_cast<const bf16*>(a.mask)[t * Cp + k]) > 0.f)) acc *= a.mask_slope;
      }
      if (a.dfeat_f32) a.dfeat_f32[t * Cp + k] = acc;
      if (a.dfeat_bf16) reinterpret_cast<bf16*>(a.dfeat_bf16)[t * Cp + k] = __float2bfloat16(acc);
    }
  }
}

int conv_out_dgrad_launch(const kair_conv_out_args& a, cudaStream_t st) {
  KB_CHECK(a.dout && a.W && (a.dfeat_f32 || a.dfeat_bf16), "kair_conv_out_dgrad: null pointer");
  KB_CHECK(a.co >= 1 && a.co <= CI_MAX, "kair_conv_out_dgrad: co=%d unsupported", a.co);
  size_t smem = (size_t)a.co * 9 * a.pitch * 4;
  KB_CHECK(smem <= 96 * 1024, "kair_conv_out_dgrad: weights do not fit in shared memory");
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(conv_out_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    attr_set[dev & 63] = true;
  }
  conv_out_dgrad_kernel<<<sm_count() * 4, 256, smem, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// conv_out weight gradient: dW[c][k][tap] = sum_t dout[c][t] * feat[t + off(tap)][k],  db[c] = sum_t dout[c][t]
__global__ void conv_out_wgrad_kernel(const kair_conv_out_args a) {
  const int Cp = a.pitch;
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  const long long T = (long long)a.batch * a.H * a.Wd;
  const bf16* feat = reinterpret_cast<const bf16*>(a.feat);
  for (int c = 0; c < a.co; ++c) {
    float acc[9][8];  // lane owns channels k = lane*2 + 64*j (+1)
#pragma unroll
    for (int tap = 0; tap < 9; ++tap)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[tap][j] = 0.f;
    float accb = 0.f;
    for (long long t = (long long)blockIdx.x * warps + (threadIdx.x >> 5); t < T; t += (long long)gridDim.x * warps) {
      const int w = (int)(t % a.Wd), h = (int)((t / a.Wd) % a.H), b = (int)(t / ((long long)a.Wd * a.H));
      const float d = a.dout[(((size_t)b * a.co + c) * a.H + h) * a.Wd + w] * a.out_scale;
      accb += d;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const int hh = h + tap / 3 - 1, ww = w + tap % 3 - 1;
        if (hh < 0 || hh >= a.H || ww < 0 || ww >= a.Wd) continue;
        const size_t roff = (((size_t)b * a.H + hh) * a.Wd + ww) * Cp;
        const bf16* row = feat + roff;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int k = lane * 2 + 64 * j;
          if (k < a.C) {
            const float2 f = a.feat_f32
                                 ? *reinterpret_cast<const float2*>(a.feat_f32 + roff + k)
                                 : __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(row + k));
            acc[tap][2 * j] += d * f.x;
            acc[tap][2 * j + 1] += d * f.y;
          }
        }
      }
    }
#pragma unroll
    for (int tap = 0; tap < 9; ++tap)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = lane * 2 + 64 * j;
        if (k < a.C) {
          atomicAdd(&a.dW[((size_t)c * a.C + k) * 9 + tap], acc[tap][2 * j]);
          if (k + 1 < a.C) atomicAdd(&a.dW[((size_t)c * a.C + k + 1) * 9 + tap], acc[tap][2 * j + 1]);
        }
      }
    if (a.db) {
      accb = (lane == 0) ? accb : 0.f;
      if (lane == 0) atomicAdd(&a.db[c], accb);
    }
  }
}

int conv_out_wgrad_launch(const kair_conv_out_args& a, cudaStream_t st) {
  KB_CHECK(a.dout && (a.feat || a.feat_f32) && a.dW, "kair_conv_out_wgrad: null pointer");
  KB_CHECK(a.C <= 256, "kair_conv_out_wgrad: C=%d unsupported (max 256)", a.C);
  conv_out_wgrad_kernel<<<sm_count(), 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------ residual helpers
// ------------------------------------------------------------------------------------------------ thin convs as GEMMs
// im2col for <= 4-channel images: one thread per (token, 8-column group); see include/kair_b200.h
__global__ void im2col3x3_kernel(const kair_im2col_args a) {
  const long long T = (long long)a.batch * a.H * a.Wd;
  const int groups = a.pitch >> 3;
  bf16* out = reinterpret_cast<bf16*>(a.out);
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < T * groups;
       e += (long long)gridDim.x * blockDim.x) {
    const long long t = e / groups;
    const int g = (int)(e - t * groups);
    const int w = (int)(t % a.Wd), h = (int)((t / a.Wd) % a.H), b = (int)(t / ((long long)a.Wd * a.H));
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int col = g * 8 + i;
      const int tap = col / a.cpad, ch = col - tap * a.cpad;
      float x = 0.f;
      if (tap < 9 && ch < a.c) {
        const int hh = h + a.dir * (tap / 3 - 1), ww = w + a.dir * (tap % 3 - 1);
        if (hh >= 0 && hh < a.H && ww >= 0 && ww < a.Wd) {
          x = a.img[(((size_t)b * a.c + ch) * a.H + hh) * a.Wd + ww];
          x = (x - (a.mean ? a.mean[ch] : 0.f)) * a.scale;
        }
      }
      if (col == a.ones_col) x = 1.0f;
      v[i] = x;
    }
    uint4 pk = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                          pack_bf16x2(v[6], v[7]));
    *reinterpret_cast<uint4*>(out + t * a.pitch + g * 8) = pk;
  }
}

int im2col3x3_launch(const kair_im2col_args& a, cudaStream_t st) {
  KB_CHECK(a.img && a.out, "kair_im2col3x3: null pointer");
  KB_CHECK(a.c >= 1 && a.cpad >= a.c && 9 * a.cpad <= a.pitch && a.pitch % 8 == 0 && (a.dir == 1 || a.dir == -1),
           "kair_im2col3x3: bad geometry c=%d cpad=%d pitch=%d dir=%d", a.c, a.cpad, a.pitch, a.dir);
  im2col3x3_kernel<<<sm_count() * 8, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

__global__ void conv_out_pack_kernel(const kair_conv_out_pack_args a) {
  const int n = a.C_pad * a.pitch;
  if (a.mode == 0) {
    bf16* Wt = reinterpret_cast<bf16*>(a.Wt);
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
      const int k = e / a.pitch, col = e - k * a.pitch;
      const int tap = col / a.cpad, c = col - tap * a.cpad;
      float v = 0.f;
      if (k < a.C && tap < 9 && c < a.co) v = a.W[((size_t)c * a.C + k) * 9 + tap];
      Wt[e] = __float2bfloat16(v);
    }
  } else {
    const int m = a.co * a.C * 9;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < m; e += gridDim.x * blockDim.x) {
      const int tap = e % 9, k = (e / 9) % a.C, c = e / (9 * a.C);
      a.dW[e] = a.D[(size_t)k * a.pitch + tap * a.cpad + c];
    }
    if (a.db && blockIdx.x == 0 && threadIdx.x < a.co)
      a.db[threadIdx.x] = a.D[(size_t)a.ones_row * a.pitch + 4 * a.cpad + threadIdx.x];
  }
}

int conv_out_pack_launch(const kair_conv_out_pack_args& a, cudaStream_t st) {
  KB_CHECK(a.mode == 0 ? (a.W && a.Wt) : (a.D && a.dW), "kair_conv_out_pack: null pointer");
  KB_CHECK(a.co >= 1 && a.cpad >= a.co && 9 * a.cpad <= a.pitch && a.C <= a.C_pad, "kair_conv_out_pack: bad geometry");
  conv_out_pack_kernel<<<64, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

// y += x (fp32, in place) and refresh the bf16 shadow copy of y
__global__ void add_inplace_kernel(float* y, const float* x, bf16* yb, size_t n4) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 a = reinterpret_cast<float4*>(y)[i];
    const float4 b = reinterpret_cast<const float4*>(x)[i];
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    reinterpret_cast<float4*>(y)[i] = a;
    if (yb) {
      uint2 o;
      o.x = pack_bf16x2(a.x, a.y);
      o.y = pack_bf16x2(a.z, a.w);
      reinterpret_cast<uint2*>(yb)[i] = o;
    }
  }
}

int add_inplace_launch(float* y, const float* x, void* yb, long long n, cudaStream_t st) {
  KB_CHECK(y && x && n > 0 && n % 4 == 0, "kair_add_inplace: bad arguments");
  add_inplace_kernel<<<sm_count() * 8, 256, 0, st>>>(y, x, reinterpret_cast<bf16*>(yb), (size_t)n / 4);
  KB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kb
