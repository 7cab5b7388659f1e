// The serial tail of a KAIR training step, fused:
//   kair_pixel_loss   CharbonnierLoss / L1 of models/loss.py:208-218 (and nn.L1Loss, select by mode) together with
//                     d(loss)/d(x) in one pass: loss = mean(sqrt((x-y)^2 + eps)), dx = (x-y) / sqrt(.) / n
//   kair_adam_ema     torch.optim.Adam's update (the arithmetic of torch/optim/adam.py `_single_tensor_adam`, L2 weight
//                     decay, no amsgrad) for ALL parameter tensors in one launch, followed in the same pass by the
//                     exponential moving average of ModelBase.update_E (models/model_base.py:247-252):
//                     e = e * decay + p_new * (1 - decay).  The reference runs 490 x (mul_ + add_) launches for the
//                     EMA of SwinIR-M plus the optimizer's own kernels; this is one kernel over 36 bytes per element.
// Both are HBM-bound elementwise kernels: coalesced 4-byte accesses, grid sized from the element count.
#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

namespace {

__global__ void pixel_loss_kernel(const kair_pixel_loss_args a) {
  __shared__ double red[8];
  const float inv_n = 1.0f / (float)a.n;
  double acc = 0.0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (long long)gridDim.x * blockDim.x) {
    const float xv = a.x[i], yv = a.y[i];
    const float d = xv - yv;
    float l, g;
    if (a.mode == 0) {         // Charbonnier
      l = sqrtf(d * d + a.eps);
      g = d / l;
    } else if (a.mode == 1) {  // L1 (sign(0) = 0 like torch)
      l = fabsf(d);
      g = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
    } else if (a.mode == 5) {  // uint8 MSE of tensor2uint(x), tensor2uint(y): np.round = round-half-even = rintf
      const float qx = rintf(fminf(fmaxf(xv, 0.f), 1.f) * 255.0f), qy = rintf(fminf(fmaxf(yv, 0.f), 1.f) * 255.0f);
      l = (qx - qy) * (qx - qy);
      g = 0.f;
    } else {                   // Poisson log-likelihood on de-normalised counts
      float lam, k, dlam;      // dlam = d(lambda_raw)/dx
      if (a.mode == 2) {
        lam = xv * a.max_pixel; k = yv * a.max_pixel; dlam = a.max_pixel;
      } else if (a.mode == 3) {
        const float lg = logf(a.max_pixel + 1.0f);
        lam = expf(xv * lg) - 1.0f; k = expf(yv * lg) - 1.0f; dlam = lg * (lam + 1.0f);
      } else {
        const float s = sqrtf(a.max_pixel + 0.375f);   // the reference's operation order: (sqrt(M + 3/8) * v)^2 - 3/8
        const float tx = s * xv, ty = s * yv;
        lam = tx * tx - 0.375f; k = ty * ty - 0.375f; dlam = 2.0f * s * tx;
      }
      const bool live = lam >= a.eps;   // torch.clamp(min=eps) passes the gradient where the input is >= eps
      lam = fmaxf(lam, a.eps);
      k = fmaxf(k, 0.f);
      l = lam - k * logf(lam);
      g = live ? (1.0f - k / lam) * dlam : 0.f;
    }
    acc += (double)l;
    if (a.dx) a.dx[i] = g * inv_n;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    atomicAdd(a.scratch, s);
    __threadfence();
    // the last block to arrive publishes the mean and re-arms the scratch words for the next call
    const unsigned long long done = atomicAdd(reinterpret_cast<unsigned long long*>(a.scratch + 1), 1ull) + 1ull;
    if (done == (unsigned long long)gridDim.x) {
      const double total = atomicAdd(a.scratch, 0.0);
      a.loss[0] = (float)(total / (double)a.n);
      a.scratch[0] = 0.0;
      *reinterpret_cast<unsigned long long*>(a.scratch + 1) = 0ull;
    }
  }
}

constexpr int AE_CHUNK = 8192;   // elements per block

__global__ void __launch_bounds__(256) adam_ema_kernel(const kair_adam_ema_args a) {
  const int t = a.chunk_tensor[blockIdx.x];
  const long long off = a.chunk_offset[blockIdx.x];
  const long long n = a.sizes[t];
  float* p = reinterpret_cast<float*>(a.p[t]);
  const float* g = reinterpret_cast<const float*>(a.g[t]);
  float* m = reinterpret_cast<float*>(a.m[t]);
  float* v = reinterpret_cast<float*>(a.v[t]);
  float* e = a.e ? reinterpret_cast<float*>(a.e[t]) : nullptr;
  const long long end = off + AE_CHUNK < n ? off + AE_CHUNK : n;
  const float omb1 = (float)(1.0 - a.beta1), omb2 = (float)(1.0 - a.beta2), omd = (float)(1.0 - a.ema_decay);
  const float beta2 = (float)a.beta2, decay = (float)a.ema_decay;
  for (long long i = off + threadIdx.x; i < end; i += blockDim.x) {
    float pi = p[i];
    float gi = g[i];
    if (a.weight_decay != 0.f) gi = fmaf(a.weight_decay, pi, gi);
    // exp_avg.lerp_(grad, 1 - beta1); exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    const float mi = fmaf(omb1, gi - m[i], m[i]);
    const float vi = fmaf(omb2 * gi, gi, beta2 * v[i]);
    // denom = exp_avg_sq.sqrt() / bias_correction2_sqrt + eps; param.addcdiv_(exp_avg, denom, value=-step_size)
    const float denom = sqrtf(vi) / a.bc2_sqrt + a.eps;
    pi = pi - a.step_size * (mi / denom);
    p[i] = pi;
    m[i] = mi;
    v[i] = vi;
    if (e) e[i] = fmaf(omd, pi, e[i] * decay);
  }
}

}  // namespace

int pixel_loss_launch(const kair_pixel_loss_args& a, cudaStream_t st) {
  KB_CHECK(a.x && a.y && a.loss && a.scratch, "kair_pixel_loss: null pointer");
  KB_CHECK(a.n > 0 && a.mode >= 0 && a.mode <= 5, "kair_pixel_loss: empty input or unknown mode %d", a.mode);
  KB_CHECK(a.mode < 2 || a.mode == 5 || a.max_pixel > 0.f, "kair_pixel_loss: Poisson modes need max_pixel > 0");
  long long blocks = (a.n + 256 * 8 - 1) / (256 * 8);
  const long long cap = 4ll * sm_count();
  if (blocks > cap) blocks = cap;
  pixel_loss_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int adam_ema_launch(const kair_adam_ema_args& a, cudaStream_t st) {
  KB_CHECK(a.p && a.g && a.m && a.v && a.sizes && a.chunk_tensor && a.chunk_offset, "kair_adam_ema: null pointer");
  KB_CHECK(a.n_chunks > 0 && a.chunk_elems == AE_CHUNK, "kair_adam_ema: chunk table must use %d-element chunks", AE_CHUNK);
  KB_CHECK(a.bc2_sqrt > 0.f && a.beta1 >= 0.0 && a.beta1 < 1.0 && a.beta2 >= 0.0 && a.beta2 < 1.0,
           "kair_adam_ema: bad hyper-parameters");
  adam_ema_kernel<<<(unsigned)a.n_chunks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kb
