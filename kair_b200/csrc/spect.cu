// SPECT input pipeline and full-frame validation metrics on the device (the fork trains and
// validates SwinIR on 2-view planar SPECT frames, BASELINE.json configs[1] "SPECT-shaped").
//
//   kair_spect_patches    data/dataset_spect.py:57-118 of the reference for a whole batch in one launch: the .dat
//                         frame pair (anterior, posterior) -> np.fliplr of the posterior view (:71-79) -> random crop
//                         (:92-96) -> one of the 8 flip / rotation modes of utils_image.augment_img (:101-103,
//                         utils/utils_image.py:387-405) -> utils_spect.normalize_spect (:108-109) -> (B, 2, P, P).
//                         Pure index arithmetic + one transcendental per element; frames stay resident in HBM
//                         (one 2 x 1024 x 256 fp32 frame is 2 MB: a 10 000-frame study is 20 GB of the 180 GB).
//   kair_spect_normalize  utils/utils_spect.py:3-35, both directions (log / anscombe / linear).
//   kair_spect_metrics    models/model_base.py:355-377 (_compute_channel_metrics: clip to [0, max], /max*255 -> uint8,
//                         PSNR and SSIM of the quantised channel; utils/utils_image.py:629-700) and utils_image.pll
//                         (:833-865, Poisson log-likelihood per pixel) for one image, all channels, as sums in double
//                         precision; the host finishes the scalars (20 log10(255 / sqrt(mse)), means).
// HBM-bound, one pass each; the SSIM windows (11 x 11 Gaussian, "valid" region) are evaluated from a shared-memory halo
// tile in fp64 like the reference's cv2.filter2D on float64 images.
#include <math.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

namespace {

__device__ __forceinline__ float spect_norm(float x, int method, float max_pixel) {
  if (method == 0) return logf(x + 1.0f) / logf(max_pixel + 1.0f);
  if (method == 1) return sqrtf(x + 0.375f) / sqrtf(max_pixel + 0.375f);
  if (method == 2) return x / max_pixel;
  return x;
}
__device__ __forceinline__ float spect_denorm(float y, int method, float max_pixel) {
  if (method == 0) return expf(y * logf(max_pixel + 1.0f)) - 1.0f;
  if (method == 1) {
    const float t = sqrtf(max_pixel + 0.375f) * y;
    return t * t - 0.375f;
  }
  if (method == 2) return y * max_pixel;
  return y;
}

// source position (row, col) inside the P x P crop of output position (i, j) under augment mode m
// (np.rot90 / np.flipud compositions of utils_image.augment_img written out)
__device__ __forceinline__ void augment_src(int m, int i, int j, int P, int& r, int& c) {
  const int n = P - 1;
  switch (m) {
    case 0: r = i; c = j; break;
    case 1: r = j; c = i; break;              // flipud(rot90)       = transpose
    case 2: r = n - i; c = j; break;          // flipud
    case 3: r = n - j; c = i; break;          // rot90, k = 3
    case 4: r = i; c = n - j; break;          // flipud(rot90, k=2)  = fliplr
    case 5: r = j; c = n - i; break;          // rot90
    case 6: r = n - i; c = n - j; break;      // rot90, k = 2
    default: r = n - j; c = n - i; break;     // flipud(rot90, k=3)  = anti-transpose
  }
}

__global__ void spect_patches_kernel(const kair_spect_patch_args a) {
  const long long per = 2ll * a.P * a.P;
  const long long total = (long long)a.B * per;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(e / per);
    const int rem = (int)(e - (long long)b * per);
    const int ch = rem / (a.P * a.P);
    const int ij = rem - ch * a.P * a.P;
    const int i = ij / a.P, j = ij - i * a.P;
    const int* sp = a.sample + 4 * b;
    const int file = sp[0], top = sp[1], left = sp[2], mode = sp[3];
    int r, c;
    augment_src(mode, i, j, a.P, r, c);
    const int y = top + r;
    int x = left + c;
    if (ch == 1) x = a.FW - 1 - x;            // posterior view is mirrored left-right before stacking
    const size_t src = (((size_t)file * 2 + ch) * a.FH + y) * a.FW + x;
    if (a.out_L) a.out_L[e] = spect_norm(__ldg(a.frames_L + src), a.method, a.max_pixel);
    if (a.out_H) a.out_H[e] = spect_norm(__ldg(a.frames_H + src), a.method, a.max_pixel);
  }
}

__global__ void spect_normalize_kernel(const kair_spect_norm_args a) {
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < a.n; e += (long long)gridDim.x * blockDim.x)
    a.y[e] = a.inverse ? spect_denorm(a.x[e], a.method, a.max_pixel) : spect_norm(a.x[e], a.method, a.max_pixel);
}

// (clip(x, 0, max) / max * 255).astype(uint8) in float32, each step rounded like numpy does (no FMA contraction)
__device__ __forceinline__ int quant255(float x, float max_val) {
  if (!(max_val > 0.f)) return 0;
  const float cl = fminf(fmaxf(x, 0.f), max_val);
  return (int)__fmul_rn(__fdiv_rn(cl, max_val), 255.0f);
}

__constant__ double c_gauss[11];   // cv2.getGaussianKernel(11, 1.5)

constexpr int SM_T = 16;           // outputs per block side
constexpr int SM_H = SM_T + 10;    // halo tile side

// blockIdx.z = channel.  Every pixel contributes to the squared-difference and Poisson sums; pixels of the "valid"
// region [5, h-5) x [5, w-5) additionally get an SSIM value (utils_image.ssim).
__global__ void __launch_bounds__(SM_T* SM_T) spect_metrics_kernel(const kair_spect_metric_args a) {
  __shared__ unsigned char te[SM_H][SM_H + 2], th[SM_H][SM_H + 2];
  __shared__ double red[3][SM_T * SM_T / 32];
  const int ch = blockIdx.z;
  const float* E = a.E + (size_t)ch * a.h * a.w;
  const float* H = a.H + (size_t)ch * a.h * a.w;
  const int y0 = blockIdx.y * SM_T, x0 = blockIdx.x * SM_T;   // this block's outputs: rows y0.., cols x0..
  const int tid = threadIdx.y * SM_T + threadIdx.x;
  for (int e = tid; e < SM_H * SM_H; e += SM_T * SM_T) {
    const int ty = e / SM_H, tx = e - ty * SM_H;
    const int y = y0 - 5 + ty, x = x0 - 5 + tx;
    unsigned char qe = 0, qh = 0;
    if (y >= 0 && y < a.h && x >= 0 && x < a.w) {
      qe = (unsigned char)quant255(E[(size_t)y * a.w + x], a.max_val);
      qh = (unsigned char)quant255(H[(size_t)y * a.w + x], a.max_val);
    }
    te[ty][tx] = qe;
    th[ty][tx] = qh;
  }
  __syncthreads();
  const int y = y0 + threadIdx.y, x = x0 + threadIdx.x;
  double sq = 0.0, ss = 0.0, pl = 0.0;
  if (y < a.h && x < a.w) {
    const double d = (double)te[threadIdx.y + 5][threadIdx.x + 5] - (double)th[threadIdx.y + 5][threadIdx.x + 5];
    sq = d * d;
    // utils_image.pll: k = round-half-even(H), mu = max(E, 1e-9) (float32); log pmf = k log mu - mu - lgamma(k + 1)
    const double k = rint((double)H[(size_t)y * a.w + x]);
    const double mu = (double)fmaxf(E[(size_t)y * a.w + x], 1e-9f);   // clamped in float32 like np.maximum on the image
    pl = k * log(mu) - mu - lgamma(k + 1.0);
    if (y >= 5 && y < a.h - 5 && x >= 5 && x < a.w - 5) {
      double m1 = 0, m2 = 0, s11 = 0, s22 = 0, s12 = 0;
      for (int dy = 0; dy < 11; ++dy) {
        double r1 = 0, r2 = 0, r11 = 0, r22 = 0, r12 = 0;
#pragma unroll
        for (int dx = 0; dx < 11; ++dx) {
          const double g = c_gauss[dx];
          const double p1 = (double)te[threadIdx.y + dy][threadIdx.x + dx];
          const double p2 = (double)th[threadIdx.y + dy][threadIdx.x + dx];
          r1 += g * p1; r2 += g * p2; r11 += g * p1 * p1; r22 += g * p2 * p2; r12 += g * p1 * p2;
        }
        const double gy = c_gauss[dy];
        m1 += gy * r1; m2 += gy * r2; s11 += gy * r11; s22 += gy * r22; s12 += gy * r12;
      }
      const double C1 = (0.01 * 255) * (0.01 * 255), C2 = (0.03 * 255) * (0.03 * 255);
      const double v1 = s11 - m1 * m1, v2 = s22 - m2 * m2, cv = s12 - m1 * m2;
      ss = ((2 * m1 * m2 + C1) * (2 * cv + C2)) / ((m1 * m1 + m2 * m2 + C1) * (v1 + v2 + C2));
    }
  }
  // block reduction of the three sums
  for (int o = 16; o > 0; o >>= 1) {
    sq += __shfl_xor_sync(0xffffffffu, sq, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
    pl += __shfl_xor_sync(0xffffffffu, pl, o);
  }
  if ((tid & 31) == 0) {
    red[0][tid >> 5] = sq;
    red[1][tid >> 5] = ss;
    red[2][tid >> 5] = pl;
  }
  __syncthreads();
  if (tid < 3) {
    double t = 0;
    for (int i = 0; i < SM_T * SM_T / 32; ++i) t += red[tid][i];
    atomicAdd(a.out + 4 * ch + tid, t);
  }
}

}  // namespace

int spect_patches_launch(const kair_spect_patch_args& a, cudaStream_t st) {
  KB_CHECK(a.sample && (a.out_L || a.out_H), "kair_spect_patches: null pointer");
  KB_CHECK((!a.out_L || a.frames_L) && (!a.out_H || a.frames_H), "kair_spect_patches: output without its frame stack");
  KB_CHECK(a.B > 0 && a.P > 0 && a.P <= a.FH && a.P <= a.FW, "kair_spect_patches: patch %d does not fit %dx%d frames",
           a.P, a.FH, a.FW);
  KB_CHECK(a.method >= 0 && a.method <= 3 && a.max_pixel > 0.f, "kair_spect_patches: bad normalisation");
  const long long total = 2ll * a.B * a.P * a.P;
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  spect_patches_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int spect_normalize_launch(const kair_spect_norm_args& a, cudaStream_t st) {
  KB_CHECK(a.x && a.y && a.n > 0, "kair_spect_normalize: null pointer / empty");
  KB_CHECK(a.method >= 0 && a.method <= 3 && a.max_pixel > 0.f, "kair_spect_normalize: bad normalisation");
  long long blocks = (a.n + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  spect_normalize_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int spect_metrics_launch(const kair_spect_metric_args& a, cudaStream_t st) {
  KB_CHECK(a.E && a.H && a.out, "kair_spect_metrics: null pointer");
  KB_CHECK(a.C >= 1 && a.C <= 65535 && a.h > 10 && a.w > 10, "kair_spect_metrics: image %dx%dx%d too small", a.C, a.h, a.w);
  static bool have_gauss_dev[64] = {false};   // per device; benign race: every caller writes the same values
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  bool& have_gauss = have_gauss_dev[dev & 63];
  if (!have_gauss) {
    double g[11], s = 0;
    for (int i = 0; i < 11; ++i) {
      g[i] = exp(-((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5));
      s += g[i];
    }
    for (int i = 0; i < 11; ++i) g[i] /= s;
    KB_CUDA(cudaMemcpyToSymbolAsync(c_gauss, g, sizeof(g), 0, cudaMemcpyHostToDevice, st));
    KB_CUDA(cudaStreamSynchronize(st));
    have_gauss = true;
  }
  const dim3 grid((a.w + SM_T - 1) / SM_T, (a.h + SM_T - 1) / SM_T, a.C);
  spect_metrics_kernel<<<grid, dim3(SM_T, SM_T), 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kb
