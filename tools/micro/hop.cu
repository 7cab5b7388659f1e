// Development micro-benchmark 3: mbarrier hop latency between two warps of a CTA (try_wait vs test_wait polling)
#include <stdio.h>
#include "common.cuh"
using namespace kb;
__device__ __forceinline__ bool mbar_test_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred P;\n\tmbarrier.test_wait.parity.shared::cta.b64 P, [%1], %2;\n\tselp.u32 %0, 1, 0, P;\n\t}\n"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
template <int MODE>
__device__ __forceinline__ void waitbar(uint64_t* bar, uint32_t parity) {
  if (MODE == 0) { while (!mbar_try_wait(bar, parity)) {} }
  else { while (!mbar_test_wait(bar, parity)) {} }
}
template <int MODE>
__global__ void k(int n, long long* out) {
  __shared__ uint64_t a, b;
  if (threadIdx.x == 0) { mbar_init(&a, 1); mbar_init(&b, 1); fence_barrier_init(); }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  long long t0 = clock64();
  if (warp == 0 && lane == 0) {
    uint32_t ph = 0;
    for (int i = 0; i < n; ++i) { mbar_arrive(&a); waitbar<MODE>(&b, ph); ph ^= 1; }
  } else if (warp == 1 && lane == 0) {
    uint32_t ph = 0;
    for (int i = 0; i < n; ++i) { waitbar<MODE>(&a, ph); ph ^= 1; mbar_arrive(&b); }
  }
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = clock64() - t0;
}
int main() {
  long long* d; cudaMalloc(&d, 8 * 148);
  long long h[148];
  int n = 10000;
  k<0><<<148, 128>>>(n, d); cudaMemcpy(h, d, 8 * 148, cudaMemcpyDeviceToHost);
  printf("try_wait : %.1f cycles per round trip (2 hops)\n", (double)h[0] / n);
  k<1><<<148, 128>>>(n, d); cudaMemcpy(h, d, 8 * 148, cudaMemcpyDeviceToHost);
  printf("test_wait: %.1f cycles per round trip (2 hops)\n", (double)h[0] / n);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
