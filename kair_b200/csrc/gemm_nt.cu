// kair_gemm_nt: weight-gradient GEMM  D[n1, n2] += sum_t A[t, n1] * B[t, n2]   (contraction over tokens)
//
// Both operands are read in their natural token-major layout, i.e. "MN-major" UMMA operands: TMA brings
// [64 tokens x 64 channels] SWIZZLE_128B atoms and tcgen05.mma consumes them with a_major = b_major = MN.
// The token range is split across CTAs; every CTA keeps a BM x BN accumulator (BM = 128, or 256 as two 128-row
// accumulators that share each B tile) in TMEM for its whole token range and finally adds it into the fp32 result with a
// TMA reduce-add (cp.reduce.async.bulk.tensor ... .add).  The kernel is bound by the L2 -> shared-memory fill rate
// (cycle counters, KB_NT_TS: producer and MMA issuer each wait half of the time; ~38 B/clk per SM whatever the tile
// shape), so BM = 256 is used when it makes n1 a single tile: every B tile is then read by half as many CTAs.  The
// epilogue staging aliases the pipeline stages (all loads are consumed by then).
// Conv mode: B is an NHWC image read through 9 shifted 4-D boxes (TMA zero fill = zero padding), the result column
// block is tap*K2.
//
// Replaces (reference file:line): autograd's weight gradients for nn.Linear (models/network_swinir.py:19-21,
// 105-107) and nn.Conv2d 3x3 (:465, :729; models/basicblock.py:65).
#include <string.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

#ifdef KB_DEVEL
#define NT_CLOCK() clock64()
#define NT_TS(p) ((p).ts)
#else   // production builds: no clock reads / timeline stores in the pipelines
#define NT_CLOCK() 0ll
#define NT_TS(p) ((long long*)nullptr)
#endif

namespace kb {

constexpr int NT_THREADS = 320;
constexpr int NT_BKT = 64;                    // tokens per pipeline stage
constexpr int NT_A_ATOM2 = 128 * NT_BKT * 2;  // 16 KiB: two [64 tok x 64 ch] atoms = one 128-row A operand
constexpr int NT_STG_PER_WARP = 8192;         // 2 x 4 KiB fp32 chunks
constexpr int NT_STG_BYTES = 8 * NT_STG_PER_WARP;
constexpr int NT_MAX_STAGES = 8;
constexpr int NT_MISC_BYTES = 1024;

struct NtParams {
  CUtensorMap tmA, tmB, tmD;
  int T, kb_total, kb_per_split, splits, tiles1, tiles2, taps, stages;
  int H, W, TW, TH, tiles_w, tiles_h;  // conv: a k-block is a TW x TH pixel patch (TW*TH = 64)
  int k2;                              // channels of B per tap (conv: column block stride in D)
  int dil;                             // conv: tap dilation (1 = dense 3x3)
  int swap;                            // conv: tmA = the INPUT image (M rows = (tap, input channel): one tap-shifted
                                       // [64 tok x 64 ch] atom per 64 rows), tmB = the output gradient; D transposed; k2 = n2 / 64
  int ring;                            // bytes of the stage ring (>= the epilogue staging that aliases it)
  long long* ts;                       // development aid (env KB_NT_TS): cycle counters of CTA 0
};

template <int BN, bool kConv, int BM>
__global__ void __launch_bounds__(NT_THREADS, 1) gemm_nt_kernel(const __grid_constant__ NtParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int B_STAGE = BN * NT_BKT * 2;
  constexpr int NT_A_STAGE = (BM / 128) * NT_A_ATOM2;
  constexpr int NACC = BM / 128;
  uint8_t* stg_base = smem;   // aliases the pipeline stages: used only after the last MMA has completed
  uint8_t* a_base = smem;
  uint8_t* b_base = a_base + p.stages * NT_A_STAGE;
  uint8_t* misc = smem + p.ring;   // ring >= NT_STG_BYTES, so the aliased staging never reaches the barriers
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(misc);
  uint64_t* empty_bar = full_bar + NT_MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + NT_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  constexpr int ACC_COLS = NACC * BN;
  constexpr uint32_t TMEM_COLS = ACC_COLS <= 64 ? 64 : (ACC_COLS <= 128 ? 128 : (ACC_COLS <= 256 ? 256 : 512));

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    tma_prefetch_desc(&p.tmD);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < p.stages; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      mbar_init(tfull_bar, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  pdl_wait();  // no global access above this point
  pdl_launch_dependents();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // work item: (tile1, tile2, tap, split)
  int w = blockIdx.x;
  const int split = w % p.splits;
  w /= p.splits;
  const int tap = w % p.taps;
  w /= p.taps;
  const int tile2 = w % p.tiles2;
  const int tile1 = w / p.tiles2;
  const int kb0 = split * p.kb_per_split;
  int kb1 = kb0 + p.kb_per_split;
  if (kb1 > p.kb_total) kb1 = p.kb_total;
  const int nkb = kb1 - kb0;  // may be <= 0 for trailing splits

  if (warp == 0) {
    {   // whole warp; the elected thread issues (elect_one, common.cuh)
      int stage = 0;
      uint32_t phase = 0;
      const int dy = tap / 3, dx = tap - dy * 3;
      // swap mode: channel offset and pixel shift of every row atom of this CTA's M tile, computed once (inside the
      // k-block loop the divisions sat on the producer's critical path: +35 % on a fill-bound kernel)
      int sw_ch[BM / 64], sw_dx[BM / 64], sw_dy[BM / 64];
      int sw_n = 0;
      if (kConv && p.swap) {
        const int at0 = tile1 * (BM / 64);
        sw_n = 9 * p.k2 - at0 < BM / 64 ? 9 * p.k2 - at0 : BM / 64;
#pragma unroll
        for (int i = 0; i < BM / 64; ++i) {
          const int at = at0 + i, tp = at / p.k2, ty = tp / 3;
          sw_ch[i] = (at - tp * p.k2) * 64;
          sw_dx[i] = (tp - ty * 3 - 1) * p.dil;
          sw_dy[i] = (ty - 1) * p.dil;
        }
      }
      // conv: (image, tile row, tile column) of the k-block, walked incrementally -- the producer thread is a serial chain
      // (wait, coordinates, issue) in a fill-bound kernel and two integer divisions per k-block were a visible part of it
      int cb = 0, cth = 0, ctw = 0;
      if (kConv) {
        const int per_img = p.tiles_w * p.tiles_h;
        cb = kb0 / per_img;
        const int r = kb0 - cb * per_img;
        cth = r / p.tiles_w;
        ctw = r - cth * p.tiles_w;
      }
      long long t_wait = 0, t_begin = NT_CLOCK();
      for (int kb = kb0; kb < kb1; ++kb) {
        const long long tw0 = NT_CLOCK();
        mbar_wait(&empty_bar[stage], phase ^ 1);
        t_wait += NT_CLOCK() - tw0;
        if (elect_one()) {
        uint8_t* a_s = a_base + stage * NT_A_STAGE;
        uint8_t* b_s = b_base + stage * B_STAGE;
        if (kConv && p.swap) {
          // swapped roles: M rows = (tap, input channel); atoms past the ninth tap are not loaded (their rows of D are
          // outside the result tensor and clipped by the TMA store)
          const int b = cb, h0 = cth * p.TH, w0 = ctw * p.TW;
          // row atom r of the M side = (tap r / cpt, input channels 64 (r % cpt) ..), cpt = n2 / 64 atoms per tap
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)(sw_n * 8192 + B_STAGE));
#pragma unroll
          for (int i = 0; i < BM / 64; ++i) {
            if (i < sw_n) tma_load_4d(a_s + i * 8192, &p.tmA, &full_bar[stage], sw_ch[i], w0 + sw_dx[i], h0 + sw_dy[i], b);
          }
#pragma unroll
          for (int i = 0; i < BN / 64; ++i)
            tma_load_4d(b_s + i * 8192, &p.tmB, &full_bar[stage], tile2 * BN + i * 64, w0, h0, b);
        } else if (kConv) {
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)(NT_A_STAGE + B_STAGE));
          const int b = cb, h0 = cth * p.TH, w0 = ctw * p.TW;
#pragma unroll
          for (int i = 0; i < BM / 64; ++i)
            tma_load_4d(a_s + i * 8192, &p.tmA, &full_bar[stage], tile1 * BM + i * 64, w0, h0, b);
#pragma unroll
          for (int i = 0; i < BN / 64; ++i)
            tma_load_4d(b_s + i * 8192, &p.tmB, &full_bar[stage], tile2 * BN + i * 64, w0 + (dx - 1) * p.dil, h0 + (dy - 1) * p.dil, b);
        } else {
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)(NT_A_STAGE + B_STAGE));
#pragma unroll
          for (int i = 0; i < BM / 64; ++i)
            tma_load_2d(a_s + i * 8192, &p.tmA, &full_bar[stage], tile1 * BM + i * 64, kb * NT_BKT);
#pragma unroll
          for (int i = 0; i < BN / 64; ++i)
            tma_load_2d(b_s + i * 8192, &p.tmB, &full_bar[stage], tile2 * BN + i * 64, kb * NT_BKT);
        }
        }
        if (kConv && ++ctw == p.tiles_w) {
          ctw = 0;
          if (++cth == p.tiles_h) {
            cth = 0;
            ++cb;
          }
        }
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (NT_TS(p) && blockIdx.x == 0 && lane == 0) {
        p.ts[0] = t_wait;
        p.ts[1] = NT_CLOCK() - t_begin;
        p.ts[2] = nkb;
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (nkb > 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BN, 1, 1);
      int stage = 0;
      uint32_t phase = 0;
      long long t_wait = 0, t_begin = NT_CLOCK();
      for (int kb = 0; kb < nkb; ++kb) {
        const long long tw0 = NT_CLOCK();
        mbar_wait(&full_bar[stage], phase);
        t_wait += NT_CLOCK() - tw0;
        tcgen05_fence_after();
        const uint32_t a_addr = smem_u32(a_base + stage * NT_A_STAGE);
        const uint32_t b_addr = smem_u32(b_base + stage * B_STAGE);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < NT_BKT / 16; ++k) {
#pragma unroll
            for (int acc = 0; acc < NACC; ++acc)
              umma_f16_ss(tmem_base + acc * BN, make_desc_mnmajor_sw128(a_addr + acc * NT_A_ATOM2 + k * 2048, NT_BKT),
                          make_desc_mnmajor_sw128(b_addr + k * 2048, NT_BKT), idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (kb == nkb - 1) umma_commit(tfull_bar);
        }
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (NT_TS(p) && blockIdx.x == 0 && lane == 0) {
        p.ts[3] = t_wait;
        p.ts[4] = NT_CLOCK() - t_begin;
      }
    }
    __syncwarp();
  } else if (nkb > 0) {
    const int ew = warp - 2;
    const int q = warp & 3;
    const int hf = ew >> 2;
    constexpr int HALF = BN / 2;
    constexpr int NCH = HALF / 32;
    uint8_t* stg = stg_base + ew * NT_STG_PER_WARP;
    const long long te0 = NT_CLOCK();
    mbar_wait(tfull_bar, 0);
    const long long te1 = NT_CLOCK();
    tcgen05_fence_after();
    const int col_base = tap * p.k2 + tile2 * BN + hf * HALF;
    int sbuf = 0;
#pragma unroll 1
    for (int cc = 0; cc < NACC * NCH; ++cc) {
      const int acc = cc / NCH, c = cc - acc * NCH;
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN + hf * HALF;
      uint32_t r[32];
      tmem_ld32(tacc + c * 32, r);
      tmem_ld_wait();
      if (elect_one()) tma_store_wait_read<1>();
      __syncwarp();
      uint8_t* buf = stg + sbuf * 4096;
      uint32_t base = smem_u32(buf) + lane * 128;
      uint32_t sw = lane & 7;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint32_t a = base + ((j ^ sw) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(r[4 * j]), "r"(r[4 * j + 1]),
                     "r"(r[4 * j + 2]), "r"(r[4 * j + 3])
                     : "memory");
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (elect_one()) {
        tma_reduce_add_2d(&p.tmD, buf, col_base + c * 32, tile1 * BM + acc * 128 + q * 32);
        tma_store_commit();
      }
      sbuf ^= 1;
    }
    if (elect_one()) tma_store_wait_all<0>();
    __syncwarp();
    if (NT_TS(p) && blockIdx.x == 0 && ew == 0 && lane == 0) {
      p.ts[5] = te1 - te0;
      p.ts[6] = NT_CLOCK() - te1;
    }
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int BN, bool kConv, int BM>
static int nt_launch_cfg(NtParams p, int grid, size_t smem, cudaStream_t st) {
  auto kern = gemm_nt_kernel<BN, kConv, BM>;
  static long long* ts_dev = nullptr;
  const bool want_ts = KB_DEV_ENV("KB_NT_TS") != nullptr;
  if (want_ts) {
    if (!ts_dev) cudaMalloc(&ts_dev, 8 * 8);
    cudaMemset(ts_dev, 0, 8 * 8);
    p.ts = ts_dev;
  }
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_set[dev & 63] = true;
  }
  KB_CUDA(launch_pdl(kern, grid, NT_THREADS, smem, st, p));
  KB_CUDA(cudaGetLastError());
  if (want_ts) {
    long long h[8];
    cudaMemcpy(h, ts_dev, sizeof(h), cudaMemcpyDeviceToHost);
    fprintf(stderr, "gemm_nt<%d,%d,%d> grid %d stages %d splits %d kb/split %d: producer wait %lld of %lld cycles (%lld k-blocks); "
            "mma wait %lld of %lld; epilogue waited %lld, ran %lld\n", BN, (int)kConv, BM, grid, p.stages, p.splits,
            p.kb_per_split, h[0], h[1], h[2], h[3], h[4], h[5], h[6]);
  }
  return 0;
}

int gemm_nt_launch(const kair_gemm_nt_args& a, cudaStream_t st) {
  KB_CHECK(a.A && a.B && a.D, "kair_gemm_nt: null pointer");
  KB_CHECK(a.bn == 64 || a.bn == 128 || a.bn == 192 || a.bn == 256, "kair_gemm_nt: bn=%d unsupported", a.bn);
  if (a.swap) {
    KB_CHECK(a.conv && a.n2 > 0 && a.n2 % 64 == 0, "kair_gemm_nt: swap needs conv = 1 and n2 a multiple of 64 (got n2=%d)", a.n2);
    KB_CHECK(a.n1 % a.bn == 0, "kair_gemm_nt: swap: n1=%d must be a multiple of bn=%d", a.n1, a.bn);
  } else {
    KB_CHECK(a.n2 > 0 && a.n2 % a.bn == 0, "kair_gemm_nt: n2=%d must be a multiple of bn=%d", a.n2, a.bn);
  }
  KB_CHECK(a.n1 > 0 && a.n1 % 64 == 0, "kair_gemm_nt: n1=%d must be a multiple of 64", a.n1);
  KB_CHECK(a.a_pitch % 8 == 0 && a.b_pitch % 8 == 0 && a.d_pitch % 4 == 0, "kair_gemm_nt: misaligned pitch");
  KB_CHECK(a.T > 0, "kair_gemm_nt: empty token range");
  if (check_device() != 0) return -3;

  NtParams p;
  memset(&p, 0, sizeof(p));
  p.T = a.T;
  p.taps = a.conv ? 9 : 1;
  p.dil = (a.conv && a.dilation > 1) ? a.dilation : 1;
  // measured at cfg2 sizes (tools/scratch/nt_ts.py): one 256-row tile beats two 128-row tiles by 13-14 % (n1 = 192);
  // with more than one tile the uneven last tile and the doubled, un-overlapped epilogue cancel the gain (n1 = 384, 576)
  int bm = (a.n1 > 128 && a.n1 <= 256) ? 256 : 128;
  {
    const char* e = KB_DEV_ENV("KB_NT_BM");   // development aid: force the accumulator height
    if (e && (atoi(e) == 128 || atoi(e) == 256)) bm = atoi(e);
  }
  p.tiles1 = (a.n1 + bm - 1) / bm;
  p.tiles2 = a.n2 / a.bn;
  p.k2 = a.n2;
  // swapped roles (64 input channels): M = (tap, input channel) = 576 rows in 128-row tiles, N = output channels
  const void* opA = a.A;
  const void* opB = a.B;
  long long pitchA = a.a_pitch, pitchB = a.b_pitch;
  int chA = a.n1, chB = a.n2;
  if (a.swap) {
    p.swap = 1;
    bm = 128;
    p.taps = 1;
    p.tiles1 = (9 * a.n2 + bm - 1) / bm;
    p.tiles2 = a.n1 / a.bn;
    p.k2 = a.n2 / 64;   // swap: atoms of 64 input channels per tap (the epilogue's tap * k2 column term is 0: taps = 1)
    opA = a.B; pitchA = a.b_pitch; chA = a.n2;
    opB = a.A; pitchB = a.a_pitch; chB = a.n1;
  }
  if (a.conv) {
    KB_CHECK((long long)a.batch * a.H * a.W == a.T, "kair_gemm_nt: conv geometry does not match T");
    p.H = a.H;
    p.W = a.W;
    int tw = 64;
    while (tw > 8 && tw > a.W) tw >>= 1;
    p.TW = tw;
    p.TH = 64 / tw;
    p.tiles_w = (a.W + p.TW - 1) / p.TW;
    p.tiles_h = (a.H + p.TH - 1) / p.TH;
    p.kb_total = a.batch * p.tiles_w * p.tiles_h;
    uint32_t box[4] = {64, (uint32_t)p.TW, (uint32_t)p.TH, 1};
    {
      uint64_t dims[4] = {(uint64_t)chA, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
      uint64_t str[3] = {(uint64_t)pitchA * 2, (uint64_t)pitchA * 2 * a.W, (uint64_t)pitchA * 2 * a.W * a.H};
      if (make_tmap(&p.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, opA, dims, str, box, nullptr,
                    CU_TENSOR_MAP_SWIZZLE_128B))
        return -2;
    }
    {
      uint64_t dims[4] = {(uint64_t)chB, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
      uint64_t str[3] = {(uint64_t)pitchB * 2, (uint64_t)pitchB * 2 * a.W, (uint64_t)pitchB * 2 * a.W * a.H};
      if (make_tmap(&p.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, opB, dims, str, box, nullptr,
                    CU_TENSOR_MAP_SWIZZLE_128B))
        return -2;
    }
  } else {
    p.kb_total = (a.T + NT_BKT - 1) / NT_BKT;
    uint32_t box[2] = {64, 64};
    {
      uint64_t dims[2] = {(uint64_t)a.n1, (uint64_t)a.T};
      uint64_t str[1] = {(uint64_t)a.a_pitch * 2};
      if (make_tmap(&p.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.A, dims, str, box, nullptr,
                    CU_TENSOR_MAP_SWIZZLE_128B))
        return -2;
    }
    {
      uint64_t dims[2] = {(uint64_t)a.n2, (uint64_t)a.T};
      uint64_t str[1] = {(uint64_t)a.b_pitch * 2};
      if (make_tmap(&p.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.B, dims, str, box, nullptr,
                    CU_TENSOR_MAP_SWIZZLE_128B))
        return -2;
    }
  }
  {
    uint64_t dims[2] = {a.swap ? (uint64_t)a.n1 : (uint64_t)a.n2 * p.taps, a.swap ? (uint64_t)9 * a.n2 : (uint64_t)a.n1};
    uint64_t str[1] = {(uint64_t)a.d_pitch * 4};
    uint32_t box[2] = {32, 32};
    if (make_tmap(&p.tmD, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, a.D, dims, str, box, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  }

  const int b_stage = a.bn * NT_BKT * 2;
  const int a_stage = (bm / 128) * NT_A_ATOM2;
  int stages = (232448 - 1024 - NT_MISC_BYTES) / (a_stage + b_stage);
  if (stages > NT_MAX_STAGES) stages = NT_MAX_STAGES;
  KB_CHECK(stages >= 2, "kair_gemm_nt: not enough shared memory");
  p.stages = stages;
  size_t ring = (size_t)stages * (a_stage + b_stage);
  if (ring < (size_t)NT_STG_BYTES) ring = NT_STG_BYTES;   // the epilogue staging aliases the ring
  p.ring = (int)ring;
  size_t smem = 1024 + ring + NT_MISC_BYTES;

  const int items = p.tiles1 * p.tiles2 * p.taps;
  int splits = a.splits > 0 ? a.splits : sm_count() / items;
  if (splits < 1) splits = 1;
  if (splits > p.kb_total) splits = p.kb_total;
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.splits = splits;
  const int grid = items * splits;

#define KB_NT_CASE(BN_, CONV_)                                                              \
  if (a.bn == BN_ && (a.conv != 0) == CONV_)                                                \
    return bm == 256 ? nt_launch_cfg<BN_, CONV_, 256>(p, grid, smem, st) : nt_launch_cfg<BN_, CONV_, 128>(p, grid, smem, st);
  KB_NT_CASE(64, true)
  KB_NT_CASE(128, true)
  KB_NT_CASE(192, true)
  KB_NT_CASE(256, true)
  KB_NT_CASE(64, false)
  KB_NT_CASE(128, false)
  KB_NT_CASE(192, false)
  KB_NT_CASE(256, false)
#undef KB_NT_CASE
  return -1;
}

}  // namespace kb
