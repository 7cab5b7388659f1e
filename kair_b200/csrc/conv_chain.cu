// kair_conv_chain: L consecutive 3x3 convolutions (64 -> 64 channels, padding 1, bias, leaky-ReLU) as ONE persistent
// launch -- the body of DnCNN / FDnCNN / IRCNN-style stacks at inference (models/network_dncnn.py:40-71: the
// `sequential(m_head, *m_body, m_tail)` loop; models/basicblock.py:58-108 builds every body entry as conv 'C' [+ folded
// 'B'] + 'R').
//
// Why: at batch 1 a 256 x 256 image is 512 output tiles of 128 pixels -- 3.5 per SM.  One launch per layer costs more in
// launch gap + prologue (barrier init, TMEM allocation, descriptor fetch, 72 KiB of weights) than the ~3.5 us of tensor
// work it carries.  Here one CTA per SM stays resident for the whole stack:
//   * the layer's 9 x [64 x 64] weight tiles sit in shared memory (72 KiB, re-loaded per layer while the grid barrier
//     of the previous layer is pending), the A operand streams through a 6-stage ring of [18 x 8 pixel x 64 channel]
//     halo columns exactly as in gemm_tn's convolution mode (three x-shifted columns per 8 x 16 pixel tile, the three
//     y taps are 1 KiB offsets into the same stage);
//   * accumulators ping-pong in TMEM (2 x 64 columns), eight epilogue warps add the bias, apply the activation and
//     TMA-store bf16 chunks; the pipelines (ring stages, accumulator stages, parities) run on ACROSS layers;
//   * layers are separated by a grid-wide barrier on a global counter (release after the CTA's stores have completed,
//     acquire + fence.proxy.async before the next layer's TMA loads): a tile's halo is written by other CTAs.
// The launch is cooperative, so all CTAs are co-resident by construction (the barrier cannot deadlock on residency);
// every spin carries the library's time-out trap.
//
//   warp 0      weights + A-operand producer (TMA), grid-barrier wait
//   warp 1      MMA issuer (tcgen05.mma 128 x 64 x 16, 36 per tile), barrier init, TMEM allocation
//   warps 2..9  epilogue: warp (q, half) owns rows 32q..32q+31, columns 32*half..+31 of every tile
//
// Arithmetic is that of kair_gemm_tn (conv mode, LINEAR epilogue) launched once per layer: same operand tiles, same
// accumulation order, same fp32 epilogue -- the results are bit-identical (tests/test_conv_chain_gpu.py).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

namespace {

constexpr int CC_THREADS = 320;
constexpr int CC_C = 64;                       // channels in = channels out
constexpr int CC_TW = 8, CC_TH = 16;           // output tile: 8 x 16 pixels = 128 rows
constexpr int CC_A_STAGE = (CC_TH + 2) * CC_TW * 128;   // 18 KiB halo column
constexpr int CC_STAGES = 6;
constexpr int CC_B_TILE = CC_C * 64 * 2;       // 8 KiB: one tap
constexpr int CC_W_BYTES = 9 * CC_B_TILE;      // 72 KiB
constexpr int CC_STG_WARP = 2 * 2048;          // two 32 x 32 bf16 chunks per epilogue warp
constexpr int CC_MAX_L = KAIR_CONV_CHAIN_MAX_LAYERS;
constexpr size_t CC_SMEM = 1024 + CC_W_BYTES + CC_STAGES * CC_A_STAGE + 8 * CC_STG_WARP + CC_MAX_L * (CC_C + 1) * 4 + 256;

struct ChainParams {
  CUtensorMap tmIn[2], tmOut[2], tmW;
  const float* bias;
  const float* slope;
  unsigned int* sync;   // [0] layer arrivals (monotonic within a launch), [1] exit arrivals; both zero between launches
  int L, H, W, tiles_w, tiles_h, m_tiles;
  int dbg;   // development ablations (devel build, env KB_CC_DBG): 1 = no A-operand traffic, 2 = no MMAs, 4 = no stores,
             // 8 = epilogue reduced to the accumulator hand-shake
};

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

#ifdef KB_DEVEL
#define CC_DBG(p) ((p).dbg)
#else
#define CC_DBG(p) 0
#endif

struct Tile {
  int b, h0, w0;
};
__device__ __forceinline__ Tile tile_of(const ChainParams& p, int t) {
  Tile r;
  const int per_img = p.tiles_w * p.tiles_h;
  r.b = t / per_img;
  const int i = t - r.b * per_img;
  const int th = i / p.tiles_w;
  r.h0 = th * CC_TH;
  r.w0 = (i - th * p.tiles_w) * CC_TW;
  return r;
}

__global__ void __launch_bounds__(CC_THREADS, 1) conv_chain_kernel(const __grid_constant__ ChainParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* w_base = smem;                                  // 9 x 8 KiB, SWIZZLE_128B K-major [64 n][64 k]
  uint8_t* a_base = w_base + CC_W_BYTES;                   // 6 x 18 KiB
  uint8_t* stg_base = a_base + CC_STAGES * CC_A_STAGE;     // 8 x 4 KiB
  float* bias_s = reinterpret_cast<float*>(stg_base + 8 * CC_STG_WARP);   // [L][64]
  float* slope_s = bias_s + CC_MAX_L * CC_C;               // [L]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(slope_s + CC_MAX_L);
  uint64_t* empty_bar = full_bar + CC_STAGES;
  uint64_t* tfull_bar = empty_bar + CC_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* wfull_bar = tempty_bar + 2;
  uint64_t* wfree_bar = wfull_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfree_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const unsigned int G = gridDim.x;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmIn[0]);
    tma_prefetch_desc(&p.tmIn[1]);
    tma_prefetch_desc(&p.tmOut[0]);
    tma_prefetch_desc(&p.tmOut[1]);
    tma_prefetch_desc(&p.tmW);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < CC_STAGES; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tfull_bar[i], 1);
        mbar_init(&tempty_bar[i], 8);
      }
      mbar_init(wfull_bar, 1);
      mbar_init(wfree_bar, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 128);
    tmem_relinquish();
  }
  // the per-layer biases and slopes of the whole stack (constant inputs of the launch)
  for (int i = threadIdx.x; i < p.L * CC_C; i += CC_THREADS) bias_s[i] = p.bias ? __ldg(p.bias + i) : 0.f;
  for (int i = threadIdx.x; i < p.L; i += CC_THREADS) slope_s[i] = __ldg(p.slope + i);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================ producer ================================
    int stage = 0;
    uint32_t phase = 0;
    for (int l = 0; l < p.L; ++l) {
      if (l > 0) mbar_wait(wfree_bar, (l - 1) & 1);        // the MMAs of layer l-1 have read the weight tiles
      if (elect_one()) {
        mbar_arrive_expect_tx(wfull_bar, (uint32_t)CC_W_BYTES);
        for (int tap = 0; tap < 9; ++tap) tma_load_2d(w_base + tap * CC_B_TILE, &p.tmW, wfull_bar, tap * 64, l * CC_C);
      }
      __syncwarp();
      if (l > 0) {
        // grid barrier: every CTA has completed the stores of layer l-1 (acquire), then order the async-proxy reads
        // that follow after it
        const unsigned int target = (unsigned int)l * G;
        if (ld_acquire_gpu(p.sync) < target) {
          const long long t0 = clock64();
          while (ld_acquire_gpu(p.sync) < target) {
            if (clock64() - t0 > KB_SPIN_TIMEOUT_CYCLES) {
              if (lane == 0) printf("kair_b200: conv_chain grid barrier timeout block %d layer %d\n", (int)blockIdx.x, l);
              __trap();
            }
          }
        }
        fence_proxy_async_all();
        __syncwarp();
      }
      const CUtensorMap* tm = &p.tmIn[l & 1];
      for (int t = blockIdx.x; t < p.m_tiles; t += G) {
        const Tile tt = tile_of(p, t);
        for (int dx = 0; dx < 3; ++dx) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (elect_one()) {
            if (CC_DBG(p) & 1) {
              mbar_arrive(&full_bar[stage]);
            } else {
              mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)CC_A_STAGE);
              tma_load_4d(a_base + stage * CC_A_STAGE, tm, &full_bar[stage], 0, tt.w0 + dx - 1, tt.h0 - 1, tt.b);
            }
          }
          __syncwarp();
          if (++stage == CC_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    constexpr uint32_t idesc = make_idesc_bf16(128, CC_C, 0, 0);
    int stage = 0;
    uint32_t phase = 0;
    uint32_t it = 0;
    for (int l = 0; l < p.L; ++l) {
      mbar_wait(wfull_bar, l & 1);
      tcgen05_fence_after();
      for (int t = blockIdx.x; t < p.m_tiles; t += G, ++it) {
        const int as = it & 1;
        mbar_wait(&tempty_bar[as], ((it >> 1) & 1) ^ 1);
        tcgen05_fence_after();
        const uint32_t tacc = tmem_base + as * CC_C;
        const bool last_tile = t + (int)G >= p.m_tiles;
        for (int dx = 0; dx < 3; ++dx) {
          mbar_wait(&full_bar[stage], phase);
          tcgen05_fence_after();
          const uint32_t a_addr = smem_u32(a_base + stage * CC_A_STAGE);
          if (elect_one()) {
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
              const uint32_t b_addr = smem_u32(w_base + (dy * 3 + dx) * CC_B_TILE);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (!(CC_DBG(p) & 2)) umma_f16_ss(tacc, make_desc_kmajor_sw128(a_addr + dy * 1024 + k * 32),
                            make_desc_kmajor_sw128(b_addr + k * 32), idesc, (dx | dy | k) != 0 ? 1u : 0u);
            }
            umma_commit(&empty_bar[stage]);
            if (dx == 2) {
              umma_commit(&tfull_bar[as]);
              if (last_tile) umma_commit(wfree_bar);
            }
          }
          __syncwarp();
          if (++stage == CC_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      if ((int)blockIdx.x >= p.m_tiles && elect_one()) mbar_arrive(wfree_bar);   // a CTA without tiles (G <= tiles: never)
      __syncwarp();
    }
  } else {
    // ================================ epilogue ================================
    const int ew = warp - 2;
    const int q = warp & 3;     // TMEM lane quarter this warp may read
    const int hf = ew >> 2;     // column half
    uint8_t* stg = stg_base + ew * CC_STG_WARP;
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    uint32_t it = 0;
    int sbuf = 0;
    for (int l = 0; l < p.L; ++l) {
      const CUtensorMap* tm = &p.tmOut[(l + 1) & 1];
      const float4* bp = reinterpret_cast<const float4*>(bias_s + l * CC_C + hf * 32);
      const float slope = slope_s[l];
      for (int t = blockIdx.x; t < p.m_tiles; t += G, ++it) {
        const int as = it & 1;
        const Tile tt = tile_of(p, t);
        mbar_wait(&tfull_bar[as], (it >> 1) & 1);
        tcgen05_fence_after();
        float v[32];
        {
          uint32_t r[32];
          tmem_ld32(tmem_base + lane_sel + as * CC_C + hf * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
        }
        // the accumulator stage is free as soon as it is in registers
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tempty_bar[as]);
        if (CC_DBG(p) & 8) continue;   // ablation: accumulator hand-shake only
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 b4 = bp[i];   // warp-uniform address: shared-memory broadcast
          add2(v[4 * i], v[4 * i + 1], v[4 * i], v[4 * i + 1], b4.x, b4.y);
          add2(v[4 * i + 2], v[4 * i + 3], v[4 * i + 2], v[4 * i + 3], b4.z, b4.w);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * slope;
        uint8_t* buf = stg + sbuf * 2048;
        if (elect_one()) tma_store_wait_read<1>();   // the store that last used this buffer has read it out
        __syncwarp();
        {  // 64-byte rows, SWIZZLE_64B: 16-byte unit j of row r lives at unit j ^ ((r >> 1) & 3)
          const uint32_t base = smem_u32(buf) + lane * 64;
          const uint32_t sw = (lane >> 1) & 3;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(base + ((j ^ sw) << 4)),
                         "r"(pack_bf16x2(v[8 * j + 0], v[8 * j + 1])), "r"(pack_bf16x2(v[8 * j + 2], v[8 * j + 3])),
                         "r"(pack_bf16x2(v[8 * j + 4], v[8 * j + 5])), "r"(pack_bf16x2(v[8 * j + 6], v[8 * j + 7]))
                         : "memory");
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (elect_one()) {
          if (!(CC_DBG(p) & 4)) tma_store_4d(tm, buf, hf * 32, tt.w0, tt.h0 + q * 4, tt.b);   // rows 32q.. = pixel rows 4q..4q+3 of the tile
          tma_store_commit();
        }
        sbuf ^= 1;
      }
      if (l + 1 < p.L) {
        // grid barrier, arrival side: this CTA's outputs of layer l are complete and visible before the counter moves
        if (elect_one()) {
          tma_store_wait_all<0>();
          __threadfence();
        }
        named_bar_sync(1, 256);
        if (threadIdx.x == 64) red_release_gpu_add(p.sync, 1u);
      }
    }
    if (elect_one()) tma_store_wait_all<0>();
    __syncwarp();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
  // exit protocol: the last CTA to get here re-arms the two counters (every CTA has passed every barrier wait by then)
  if (threadIdx.x == 0) {
    const unsigned int done = atomicAdd(p.sync + 1, 1u) + 1u;
    if (done == G) {
      p.sync[0] = 0u;
      __threadfence();
      p.sync[1] = 0u;
    }
  }
}

}  // namespace

int conv_chain_launch(const kair_conv_chain_args& a, cudaStream_t st) {
  KB_CHECK(a.act0 && a.act1 && a.Wp && a.slope && a.sync, "kair_conv_chain: null pointer");
  KB_CHECK(a.L >= 1 && a.L <= CC_MAX_L, "kair_conv_chain: L=%d layers (1..%d)", a.L, CC_MAX_L);
  KB_CHECK(a.batch > 0 && a.H > 0 && a.W > 0, "kair_conv_chain: empty geometry");
  KB_CHECK(a.pitch >= CC_C && a.pitch % 8 == 0, "kair_conv_chain: pitch=%lld must be >= 64 and a multiple of 8",
           (long long)a.pitch);
  if (check_device() != 0) return -3;

  ChainParams p;
  memset(&p, 0, sizeof(p));
  p.bias = a.bias;
  p.slope = a.slope;
  p.sync = reinterpret_cast<unsigned int*>(a.sync);
  {
    const char* e = KB_DEV_ENV("KB_CC_DBG");
    p.dbg = e ? atoi(e) : 0;
  }
  p.L = a.L;
  p.H = a.H;
  p.W = a.W;
  p.tiles_w = (a.W + CC_TW - 1) / CC_TW;
  p.tiles_h = (a.H + CC_TH - 1) / CC_TH;
  const long long m_tiles = (long long)a.batch * p.tiles_w * p.tiles_h;
  KB_CHECK(m_tiles < (1ll << 30), "kair_conv_chain: too many tiles");
  p.m_tiles = (int)m_tiles;
  void* acts[2] = {a.act0, a.act1};
  for (int i = 0; i < 2; ++i) {
    uint64_t dims[4] = {(uint64_t)CC_C, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
    uint64_t str[3] = {(uint64_t)a.pitch * 2, (uint64_t)a.pitch * 2 * a.W, (uint64_t)a.pitch * 2 * a.W * a.H};
    uint32_t box_in[4] = {64, CC_TW, CC_TH + 2, 1};
    uint32_t box_out[4] = {32, CC_TW, 32 / CC_TW, 1};
    if (make_tmap(&p.tmIn[i], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, acts[i], dims, str, box_in, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
    if (make_tmap(&p.tmOut[i], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, acts[i], dims, str, box_out, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_64B))
      return -2;
  }
  {
    uint64_t dims[2] = {(uint64_t)9 * CC_C, (uint64_t)a.L * CC_C};
    uint64_t str[1] = {(uint64_t)9 * CC_C * 2};
    uint32_t box[2] = {64, (uint32_t)CC_C};
    if (make_tmap(&p.tmW, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.Wp, dims, str, box, nullptr, CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  }
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(conv_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CC_SMEM));
    attr_set[dev & 63] = true;
  }
  const int grid = p.m_tiles < sm_count() ? p.m_tiles : sm_count();
  // cooperative launch: the runtime guarantees that all CTAs are resident together (or fails the launch)
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(CC_THREADS);
  cfg.dynamicSmemBytes = CC_SMEM;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;
  at[0].val.cooperative = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  KB_CUDA(cudaLaunchKernelEx(&cfg, conv_chain_kernel, p));
  return 0;
}

}  // namespace kb
