// kair_pixel_loss_noalloc: the pixel loss of kair_pixel_loss (train_step.cu; Charbonnier / L1 / Poisson log-likelihood /
// uint8 MSE and the gradient d(loss)/dx) for callers that cannot allocate: code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
//
// kair_pixel_loss takes its two scratch words from the caller.  Here the scratch is a file-scope __device__ array that
// every GPU which loads the library holds until exit (PLN_MAX_BLOCKS fp64 partial sums + one arrival counter, 64.25 KiB):
//   * a block reduces PLN_BLOCK_ELEMS consecutive elements to one fp64 partial and stores it in its own slot;
//   * the last block to arrive sums the slots in index order with a fixed tree, publishes the mean and re-arms the
//     counter.  No floating-point atomics: the loss is bit-reproducible from call to call (kair_pixel_loss adds its block
//     sums with atomicAdd in arrival order).
// Consequences of the fixed array: n <= KAIR_PIXEL_LOSS_NOALLOC_MAX_N (= slots x elements per block), larger inputs are
// rejected; ALL calls on a device share the array, so two calls that overlap on the GPU (different streams, or graphs
// replayed concurrently) race on it -- serialise them.  The partial slots need no clearing (every slot that is read was
// written by the same launch); the counter is left at zero by the kernel and is zero at load time.
//
// Host code must not pass the array's NAME to a kernel: in host code it is a host-side stand-in, not a device address.
// The address is taken once per device with cudaGetSymbolAddress (below), with that device current.
#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

namespace {

constexpr int PLN_THREADS = 256;
constexpr int PLN_PER_THREAD = 32;
constexpr int PLN_BLOCK_ELEMS = PLN_THREADS * PLN_PER_THREAD;   // 8192 elements per block
constexpr int PLN_MAX_BLOCKS = 8192;                            // partial-sum slots
static_assert((long long)PLN_BLOCK_ELEMS * PLN_MAX_BLOCKS == KAIR_PIXEL_LOSS_NOALLOC_MAX_N,
              "include/kair_b200.h states the limit the scratch array implies");

struct PlnScratch {
  double partial[PLN_MAX_BLOCKS];
  unsigned long long arrived;      // blocks of the running launch that have stored their partial; 0 between launches
  unsigned long long pad[31];
};
static __device__ __align__(256) PlnScratch pln_scratch;

// the per-element arithmetic of pixel_loss_kernel (train_step.cu), mode for mode
__device__ __forceinline__ void pln_element(const kair_pixel_loss_args& a, float xv, float yv, float& l, float& g) {
  const float d = xv - yv;
  if (a.mode == 0) {         // Charbonnier
    l = sqrtf(d * d + a.eps);
    g = d / l;
  } else if (a.mode == 1) {  // L1 (sign(0) = 0 like torch)
    l = fabsf(d);
    g = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
  } else if (a.mode == 5) {  // uint8 MSE of tensor2uint(x), tensor2uint(y)
    const float qx = rintf(fminf(fmaxf(xv, 0.f), 1.f) * 255.0f), qy = rintf(fminf(fmaxf(yv, 0.f), 1.f) * 255.0f);
    l = (qx - qy) * (qx - qy);
    g = 0.f;
  } else {                   // Poisson log-likelihood on de-normalised counts
    float lam, k, dlam;
    if (a.mode == 2) {
      lam = xv * a.max_pixel; k = yv * a.max_pixel; dlam = a.max_pixel;
    } else if (a.mode == 3) {
      const float lg = logf(a.max_pixel + 1.0f);
      lam = expf(xv * lg) - 1.0f; k = expf(yv * lg) - 1.0f; dlam = lg * (lam + 1.0f);
    } else {
      const float s = sqrtf(a.max_pixel + 0.375f);
      const float tx = s * xv, ty = s * yv;
      lam = tx * tx - 0.375f; k = ty * ty - 0.375f; dlam = 2.0f * s * tx;
    }
    const bool live = lam >= a.eps;
    lam = fmaxf(lam, a.eps);
    k = fmaxf(k, 0.f);
    l = lam - k * logf(lam);
    g = live ? (1.0f - k / lam) * dlam : 0.f;
  }
}

// fixed-shape block reduction: the same tree for every call, so the result does not depend on timing
__device__ __forceinline__ double pln_block_sum(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0) {
    for (int w = 0; w < PLN_THREADS / 32; ++w) s += red[w];
  }
  __syncthreads();
  return s;   // valid in thread 0
}

__global__ void __launch_bounds__(PLN_THREADS) pixel_loss_noalloc_kernel(const kair_pixel_loss_args a, PlnScratch* sc) {
  __shared__ double red[PLN_THREADS / 32];
  __shared__ bool is_last;
  const float inv_n = 1.0f / (float)a.n;
  const long long base = (long long)blockIdx.x * PLN_BLOCK_ELEMS;
  double acc = 0.0;
#pragma unroll 4
  for (int u = 0; u < PLN_PER_THREAD; ++u) {
    const long long i = base + (long long)u * PLN_THREADS + threadIdx.x;   // coalesced 4-byte accesses
    if (i < a.n) {
      float l, g;
      pln_element(a, a.x[i], a.y[i], l, g);
      acc += (double)l;
      if (a.dx) a.dx[i] = g * inv_n;
    }
  }
  const double bsum = pln_block_sum(acc, red);
  if (threadIdx.x == 0) {
    sc->partial[blockIdx.x] = bsum;
    __threadfence();
    is_last = atomicAdd(&sc->arrived, 1ull) + 1ull == (unsigned long long)gridDim.x;
  }
  __syncthreads();
  if (!is_last) return;
  // the last block to arrive: every partial of this launch is visible (fence above, fence below)
  __threadfence();
  double t = 0.0;
  for (int s = threadIdx.x; s < (int)gridDim.x; s += PLN_THREADS) t += __ldcg(&sc->partial[s]);
  const double total = pln_block_sum(t, red);
  if (threadIdx.x == 0) {
    a.loss[0] = (float)(total / (double)a.n);
    sc->arrived = 0ull;   // re-armed for the next call
  }
}

// device address of the scratch array, resolved once per device with that device current
PlnScratch* pln_scratch_ptr(int dev) {
  static PlnScratch* ptr[64] = {nullptr};
  static bool have[64] = {false};
  if (dev < 0 || dev >= 64) return nullptr;
  if (!have[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, pln_scratch);
    if (e != cudaSuccess) {
      set_error("kair_pixel_loss_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return nullptr;
    }
    ptr[dev] = reinterpret_cast<PlnScratch*>(p);
    have[dev] = true;   // benign race: two threads resolve the same address
  }
  return ptr[dev];
}

}  // namespace

int pixel_loss_noalloc_launch(const kair_pixel_loss_args& a, cudaStream_t st) {
  KB_CHECK(a.x && a.y && a.loss, "kair_pixel_loss_noalloc: null pointer");
  KB_CHECK(a.scratch == nullptr, "kair_pixel_loss_noalloc: `scratch` must be NULL (the library's static array is used)");
  KB_CHECK(a.n > 0 && a.mode >= 0 && a.mode <= 5, "kair_pixel_loss_noalloc: empty input or unknown mode %d", a.mode);
  KB_CHECK(a.mode < 2 || a.mode == 5 || a.max_pixel > 0.f, "kair_pixel_loss_noalloc: Poisson modes need max_pixel > 0");
  KB_CHECK(a.n <= (long long)KAIR_PIXEL_LOSS_NOALLOC_MAX_N,
           "kair_pixel_loss_noalloc: n=%lld exceeds the static scratch array's limit of %lld elements "
           "(use kair_pixel_loss with caller-owned scratch for larger inputs)",
           (long long)a.n, (long long)KAIR_PIXEL_LOSS_NOALLOC_MAX_N);
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  KB_CHECK(dev >= 0 && dev < 64, "kair_pixel_loss_noalloc: device index %d out of range", dev);
  PlnScratch* sc = pln_scratch_ptr(dev);
  if (!sc) return -2;
  const unsigned blocks = (unsigned)((a.n + PLN_BLOCK_ELEMS - 1) / PLN_BLOCK_ELEMS);
  pixel_loss_noalloc_kernel<<<blocks, PLN_THREADS, 0, st>>>(a, sc);
  KB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kb

extern "C" int kair_pixel_loss_noalloc(const kair_pixel_loss_args* args, void* stream) {
  if (!args) {
    kb::set_error("kair_pixel_loss_noalloc: null args");
    return -1;
  }
  return kb::pixel_loss_noalloc_launch(*args, (cudaStream_t)stream);
}
