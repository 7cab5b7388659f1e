// Development micro-benchmark 2: what limits the per-thread TMA issue rate?  Box = 32 rows x 128 B (4 KiB), L2-resident.
#include <stdio.h>
#include <string.h>
#include "common.cuh"
#include "host.h"
using namespace kb;
struct P { CUtensorMap tm; const CUtensorMap* tm_g; int boxes_per_cta, variant, producers, group; };

__global__ void __launch_bounds__(256, 1) k(const __grid_constant__ P p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + 200 * 1024);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { for (int i = 0; i < 64; ++i) mbar_init(&full[i], 1); fence_barrier_init(); }
  __syncthreads();
  const CUtensorMap* tm = (p.variant == 2) ? p.tm_g : &p.tm;
  // each producer warp owns 8 slots of 4 KiB; group = boxes per expect_tx; waits for its own previous group only
  if (warp < p.producers && lane == 0) {
    int row = (blockIdx.x * p.producers + warp) * 32 * p.boxes_per_cta;
    uint32_t ph = 0;
    for (int b = 0; b < p.boxes_per_cta; b += p.group) {
      uint64_t* bar = &full[warp];
      mbar_arrive_expect_tx(bar, 4096u * p.group);
      for (int g = 0; g < p.group; ++g)
        tma_load_2d(smem + (warp * 8 + (g & 7)) * 4096, tm, bar, 0, (row + (b + g) * 32) & 0xFFFF);
      if (p.variant != 1) { mbar_wait(bar, ph); ph ^= 1; }   // variant 1: fire and forget except at the end
      else if (b + p.group >= p.boxes_per_cta) { }
    }
    if (p.variant == 1) {  // drain: wait until all phases completed (each group completes a phase; poll last)
      // phases complete in order; after issuing G groups the barrier has flipped G times; just spin a while
      for (int i = 0; i < 2000; ++i) __nanosleep(100);
    }
  }
}

int main() {
  void* buf; cudaMalloc(&buf, 64 << 20); cudaMemset(buf, 1, 64 << 20);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
  P p; memset(&p, 0, sizeof(p));
  uint64_t dims[2] = {64, 65536 * 8}; uint64_t str[1] = {128}; uint32_t box[2] = {64, 32};
  if (make_tmap(&p.tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, nullptr, CU_TENSOR_MAP_SWIZZLE_128B)) return 1;
  CUtensorMap* tmg; cudaMalloc(&tmg, sizeof(CUtensorMap)); cudaMemcpy(tmg, &p.tm, sizeof(CUtensorMap), cudaMemcpyHostToDevice);
  p.tm_g = tmg;
  printf("%-8s %-5s %-6s %10s %12s\n", "variant", "prod", "group", "us", "ns/op/thread");
  for (int variant = 0; variant < 3; ++variant)
  for (int prod = 1; prod <= 8; prod *= 2)
  for (int group = 1; group <= 8; group *= 2) {
    if (variant == 1) continue;
    p.variant = variant; p.producers = prod; p.group = group; p.boxes_per_cta = 256;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<<<148, 256, 210 * 1024>>>(p); k<<<148, 256, 210 * 1024>>>(p);
    cudaEventRecord(e0);
    for (int r = 0; r < 10; ++r) k<<<148, 256, 210 * 1024>>>(p);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (cudaGetLastError() != cudaSuccess) { printf("cuda error\n"); return 1; }
    double us = ms * 1e3 / 10;
    printf("%-8d %-5d %-6d %10.1f %12.1f   (%.1f B/ns/SM)\n", variant, prod, group, us, us * 1e3 / p.boxes_per_cta, 4096.0 * p.boxes_per_cta * prod / (us * 1e3));
  }
  return 0;
}
