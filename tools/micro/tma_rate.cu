// Development micro-benchmark: per-SM TMA load rate vs box shape / pitch / stages / number of producer threads.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../kair_b200/csrc tma_rate.cu ../../kair_b200/csrc/host.cu -o tma_rate -lcuda
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"
#include "host.h"
using namespace kb;

struct P {
  CUtensorMap tm;
  int rows_per_box, boxes_per_cta, stages, producers, colblocks;
  long long total_rows;
};

__global__ void __launch_bounds__(128, 1) tma_rate_kernel(const __grid_constant__ P p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int stage_bytes = p.rows_per_box * 128;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.stages * p.producers * stage_bytes);
  uint64_t* empty = full + 64;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < p.stages * p.producers; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    fence_barrier_init();
  }
  __syncthreads();
  // warps 0..producers-1: producer lane 0; warp 2,3: consumers for producer (warp-2)
  if (warp < p.producers && lane == 0) {
    int stage = 0; uint32_t phase = 0;
    long long box0 = (long long)blockIdx.x * p.boxes_per_cta + warp;
    int rb = (int)(box0 / p.colblocks), cb = (int)(box0 % p.colblocks);   // advanced incrementally below
    for (int b = warp; b < p.boxes_per_cta; b += p.producers) {
      int si = warp * p.stages + stage;
      mbar_wait(&empty[si], phase ^ 1);
      mbar_arrive_expect_tx(&full[si], stage_bytes);
      tma_load_2d(smem + si * stage_bytes, &p.tm, &full[si], cb * 64, rb * p.rows_per_box);
      cb += p.producers;
      while (cb >= p.colblocks) { cb -= p.colblocks; ++rb; }
      if (++stage == p.stages) { stage = 0; phase ^= 1; }
    }
  } else if (warp >= 2 && warp - 2 < p.producers && lane == 0) {
    const int pw = warp - 2;
    int stage = 0; uint32_t phase = 0;
    for (int b = pw; b < p.boxes_per_cta; b += p.producers) {
      int si = pw * p.stages + stage;
      mbar_wait(&full[si], phase);
      mbar_arrive(&empty[si]);
      if (++stage == p.stages) { stage = 0; phase ^= 1; }
    }
  }
}

int main(int argc, char** argv) {
  const int nsm = 148;
  size_t bytes = 512ull << 20;
  void* buf; cudaMalloc(&buf, bytes); cudaMemset(buf, 1, bytes);
  cudaFuncSetAttribute(tma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
  printf("%-8s %-6s %-6s %-5s %-5s %-9s %10s %10s %9s\n", "pitchB", "rows", "stages", "prod", "L2res", "MB", "us", "GB/s", "B/ns/SM");
  int pitches[] = {128, 384, 1152};
  int rowsv[] = {16, 32, 64, 128, 256};
  int stagesv[] = {2, 4, 8};
  for (int l2res = 0; l2res < 2; ++l2res)
  for (int pi = 0; pi < 3; ++pi)
  for (int ri = 0; ri < 5; ++ri)
  for (int si = 0; si < 3; ++si)
  for (int prod = 1; prod <= 2; ++prod) {
    int pitch = pitches[pi], rows = rowsv[ri], stages = stagesv[si];
    if ((long long)stages * prod * rows * 128 > 200 * 1024) continue;
    int colblocks = pitch / 128;
    size_t use = l2res ? (48ull << 20) : bytes;          // L2-resident working set vs HBM streaming
    long long total_rows = use / pitch;
    long long boxes = total_rows / rows * colblocks;
    int per_cta = (int)(boxes / nsm);
    if (per_cta < 8) continue;
    P p; memset(&p, 0, sizeof(p));
    uint64_t dims[2] = {(uint64_t)pitch / 2, (uint64_t)total_rows};
    uint64_t str[1] = {(uint64_t)pitch};
    uint32_t box[2] = {64, (uint32_t)rows};
    if (make_tmap(&p.tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, str, box, nullptr, CU_TENSOR_MAP_SWIZZLE_128B)) { printf("tmap fail %s\n", get_error()); return 1; }
    p.rows_per_box = rows; p.boxes_per_cta = per_cta; p.stages = stages; p.producers = prod; p.colblocks = colblocks; p.total_rows = total_rows;
    size_t smem = 1024 + (size_t)stages * prod * rows * 128 + 2048;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int reps = l2res ? 20 : 3;
    for (int w = 0; w < 2; ++w) tma_rate_kernel<<<nsm, 128, smem>>>(p);
    cudaEventRecord(e0);
    for (int r = 0; r < reps; ++r) tma_rate_kernel<<<nsm, 128, smem>>>(p);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (cudaGetLastError() != cudaSuccess) { printf("cuda error\n"); return 1; }
    double us = ms * 1e3 / reps;
    double moved = (double)per_cta * nsm * rows * 128;
    printf("%-8d %-6d %-6d %-5d %-5d %-9.1f %10.1f %10.0f %9.1f\n", pitch, rows, stages, prod, l2res, moved / 1e6, us, moved / us / 1e3, moved / us / 1e3 / nsm);
  }
  return 0;
}
