// kair_mlp_fused: the token-local MLP half of a Swin block as ONE persistent tcgen05 kernel
//
//     u = X W1^T + b1 ;  g = gelu_erf(u) ;  v = (g W2^T + b2) * drop_path + resid ;  [x' = LayerNorm(v)]
//
// (Mlp.forward, models/network_swinir.py:24-30, plus "x = x + drop_path(mlp(norm2(x)))" and the next block's norm1,
// :277 / :245; norm2's affine is folded into W1 by kair_pack_linear.)  The unfused plan runs this as two kair_gemm_tn
// launches (fc1 + GELU epilogue, fc2 + residual + LayerNorm epilogue) and round-trips the 2C-wide hidden activations
// u and g through HBM: 2 x 768 B written and 768 B re-read per token.  Here the hidden tile never leaves the SM:
//
//   warp 0      TMA producer: the 128-token X tile once (resident for the tile), then W1 / W2 tiles streamed from L2
//               through one ring in exactly the order the MMA warp consumes them
//   warp 1      MMA issuer:   per 128-wide hidden chunk c   acc1[c & 1] = X W1[c]^T          (UMMA 128 x 128 x 16)
//                             then, one chunk behind,        acc2 += G[c-1] W2[:, c-1]^T      (UMMA 128 x N/2 x 16)
//               so the GELU epilogue of chunk c overlaps the first GEMM of chunk c + 1
//   warps 2..9  epilogue:     acc1 -> +b1 -> GELU -> bf16 -> G tile in shared memory, written directly in the UMMA
//                             K-major SWIZZLE_128B layout (it is the A operand of the second GEMM);
//                             at the end of the tile acc2 -> +b2, DropPath scale, fp32 residual (TMA-prefetched),
//                             LayerNorm statistics, TMA stores of v (fp32) and x' (bf16) -- the RESID_LN epilogue of
//                             gemm_tn.cu on a 2 x 4 KiB staging ring per warp
//               Measured at 131 072 tokens (tools/bench_mlp_fused.py): 112 us forward-only, 120 us with u / g written,
//               against 124 us for the unfused pair back to back (whose g re-read is served by the 126 MB L2).  The
//               kernel is bound by its epilogue latency chains (tcgen05.ld -> math -> st.shared -> TMA), not by the
//               weight ring (2 -> 4 stages: 115 -> 112 us) nor by the tensor pipe.  A variant with the two epilogue
//               roles on separate warps (18 warps, 96 registers) was tried and is slower (150 us: spills, and the
//               single acc2 buffer still serialises the output phase against the next tile's second GEMM).
//   TMEM: acc1 2 x 128 columns, acc2 N columns (N <= 256)  ->  512 columns
// Training plans additionally get u and g written once (bf16, for the backward pass); g leaves straight from the G tile.
// Same rounding points and the same k-order as the unfused pair, so the results are bit-identical to it.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

namespace {

constexpr int MF_THREADS = 320;
constexpr int MF_BM = 128;
constexpr int MF_HC = 128;                 // hidden columns per chunk
constexpr int MF_STAGE = 16384;            // one weight tile of the ring ([128 x 64] W1 or [N/2 x 64] W2, bf16)
constexpr int MF_MAX_STAGES = 8;
constexpr int MF_STG_WARP = 8192;          // per epilogue warp: two 4 KiB chunk buffers
constexpr int MF_G_BYTES = 2 * MF_BM * 64 * 2;   // G tile: [128 x 128] bf16 as two K-major slices
constexpr int MF_MISC = 8192;              // barriers, row-exchange buffer, biases

struct MfParams {
  CUtensorMap tmA, tmW1, tmW2, tmResid, tmOutF, tmOutB, tmU, tmG;
  const float* bias1;
  const float* bias2;
  const float* acc_scale;
  float* rowstat;
  int tokens_per_sample, M, m_tiles, stages;
  int n_valid, ones_col, ones_col_g, ln;
  int has_resid, has_outF, has_outB, has_u, has_g;
  float ln_eps;
  long long* ts;   // development timeline (devel build, env KB_MF_TS): clock64 stamps of epilogue warp 2 / lane 0 of CTA 0
};

#ifdef KB_DEVEL
#define MF_STAMP()                                                                                   \
  do {                                                                                               \
    if (p.ts && blockIdx.x == 0 && threadIdx.x == 64 && tsn < 512) p.ts[tsn++] = clock64();          \
  } while (0)
#else
#define MF_STAMP() \
  do {             \
  } while (0)
#endif

__device__ __forceinline__ void mf_stage_bf16(uint8_t* buf, int lane, const float (&v)[32]) {
  // 32 rows x 32 bf16 (64-byte rows), SWIZZLE_64B: 16-byte unit j of row r lives at unit j ^ ((r >> 1) & 3)
  uint32_t base = smem_u32(buf) + lane * 64;
  uint32_t sw = (lane >> 1) & 3;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t a = base + ((j ^ sw) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pack_bf16x2(v[8 * j + 0], v[8 * j + 1])),
                 "r"(pack_bf16x2(v[8 * j + 2], v[8 * j + 3])), "r"(pack_bf16x2(v[8 * j + 4], v[8 * j + 5])),
                 "r"(pack_bf16x2(v[8 * j + 6], v[8 * j + 7]))
                 : "memory");
  }
}
__device__ __forceinline__ void mf_stage_f32(uint8_t* buf, int lane, const float (&v)[32]) {
  // 32 rows x 32 fp32 (128-byte rows), SWIZZLE_128B: 16-byte unit j of row r lives at unit j ^ (r & 7)
  uint32_t base = smem_u32(buf) + lane * 128;
  uint32_t sw = lane & 7;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    uint32_t a = base + ((j ^ sw) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(__float_as_uint(v[4 * j + 0])),
                 "r"(__float_as_uint(v[4 * j + 1])), "r"(__float_as_uint(v[4 * j + 2])),
                 "r"(__float_as_uint(v[4 * j + 3]))
                 : "memory");
  }
}

// kTrain: u and g are written for the backward pass; kLN: LayerNorm epilogue.  Compile-time, because the epilogue loops
// are instruction / branch bound (the same finding as gemm_tn's lean LINEAR instantiation).
template <int N, bool kTrain, bool kLN>
__global__ void __launch_bounds__(MF_THREADS, 1) mlp_fused_kernel(const __grid_constant__ MfParams p) {
  constexpr int KS1 = N / 64;              // k-slices of the first GEMM
  constexpr int F = 2 * N;                 // hidden width (mlp_ratio 2, padded)
  constexpr int NC = F / MF_HC;            // hidden chunks per tile
  constexpr int NH = N / 2;                // rows of one W2 half tile == columns of one second-GEMM UMMA
  constexpr int HALF = N / 2;              // output columns per epilogue warp
  constexpr int NCH = HALF / 32;           // 32-column chunks per epilogue warp
  constexpr int A_BYTES = KS1 * MF_BM * 64 * 2;
  static_assert(F % MF_HC == 0 && N % 64 == 0 && N <= 256, "unsupported width");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* stg_base = smem;                              // 8 x 8 KiB
  uint8_t* a_base = stg_base + 8 * MF_STG_WARP;          // X tile
  uint8_t* g_base = a_base + A_BYTES;                    // G tile
  uint8_t* w_base = g_base + MF_G_BYTES;                 // weight ring
  uint8_t* misc = w_base + p.stages * MF_STAGE;
  uint64_t* w_full = reinterpret_cast<uint64_t*>(misc);
  uint64_t* w_empty = w_full + MF_MAX_STAGES;
  uint64_t* a_full = w_empty + MF_MAX_STAGES;
  uint64_t* a_empty = a_full + 1;
  uint64_t* acc1_full = a_empty + 1;      // [2]
  uint64_t* acc1_empty = acc1_full + 2;   // [2]
  uint64_t* g_full = acc1_empty + 2;
  uint64_t* g_empty = g_full + 1;
  uint64_t* acc2_full = g_empty + 1;
  uint64_t* acc2_empty = acc2_full + 1;
  uint64_t* ld_bar = acc2_empty + 1;      // [8 warps][2 buffers]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(ld_bar + 16);
  float* xch = reinterpret_cast<float*>(misc + 512);     // [2 slots][2 halves][2 values][128 rows] = 4 KiB
  float* bias1_s = reinterpret_cast<float*>(misc + 512 + 4096);   // [F] then [N]: up to 3 KiB
  float* bias2_s = bias1_s + F;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmW1);
    tma_prefetch_desc(&p.tmW2);
    if (p.has_resid) tma_prefetch_desc(&p.tmResid);
    if (p.has_outF) tma_prefetch_desc(&p.tmOutF);
    if (p.has_outB) tma_prefetch_desc(&p.tmOutB);
    if (kTrain) tma_prefetch_desc(&p.tmU);
    if (kTrain) tma_prefetch_desc(&p.tmG);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < MF_MAX_STAGES; ++i) {
        mbar_init(&w_full[i], 1);
        mbar_init(&w_empty[i], 1);
      }
      mbar_init(a_full, 1);
      mbar_init(a_empty, 1);
      for (int i = 0; i < 2; ++i) {
        mbar_init(&acc1_full[i], 1);
        mbar_init(&acc1_empty[i], 8);
      }
      mbar_init(g_full, 8);
      mbar_init(g_empty, 1);
      mbar_init(acc2_full, 1);
      mbar_init(acc2_empty, 8);
      for (int i = 0; i < 16; ++i) mbar_init(&ld_bar[i], 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  // barriers / TMEM slot visible before the dependency wait; after it the producer and the issuer start at once and
  // only the epilogue warps (the biases' sole readers) wait for their global load
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // no global access above this point
  pdl_launch_dependents();
  if (warp >= 2) {
    for (int i = threadIdx.x - 64; i < F + N; i += 256) {
      const float* src = i < F ? p.bias1 : p.bias2;
      bias1_s[i] = src ? __ldg(src + (i < F ? i : i - F)) : 0.f;
    }
    named_bar_sync(4, 256);
  }
  const uint32_t tmem_acc1 = tmem_base;          // 2 x 128 columns
  const uint32_t tmem_acc2 = tmem_base + 256;    // N columns

  if (warp == 0) {
    // ================================ TMA producer ================================
    // whole warp runs the loops; the elected thread issues (elect_one, common.cuh)
    {
      int stage = 0;
      uint32_t phase = 0;
      auto next_stage = [&]() {
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      };
      auto load_w1 = [&](int c, int ks) {
        mbar_wait(&w_empty[stage], phase ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(&w_full[stage], (uint32_t)(MF_HC * 128));
          tma_load_2d(w_base + stage * MF_STAGE, &p.tmW1, &w_full[stage], ks * 64, c * MF_HC);
        }
        next_stage();
      };
      auto load_w2 = [&](int c) {
        for (int hs = 0; hs < 2; ++hs)
          for (int nh = 0; nh < 2; ++nh) {
            mbar_wait(&w_empty[stage], phase ^ 1);
            if (elect_one()) {
              mbar_arrive_expect_tx(&w_full[stage], (uint32_t)(NH * 128));
              tma_load_2d(w_base + stage * MF_STAGE, &p.tmW2, &w_full[stage], (c * 2 + hs) * 64, nh * NH);
            }
            next_stage();
          }
      };
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < p.m_tiles; tile += gridDim.x, ++it) {
        mbar_wait(a_empty, (it & 1) ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(a_full, (uint32_t)A_BYTES);
          for (int ks = 0; ks < KS1; ++ks)
            tma_load_2d(a_base + ks * (MF_BM * 128), &p.tmA, a_full, ks * 64, tile * MF_BM);
        }
        for (int c = 0; c < NC; ++c) {
          for (int ks = 0; ks < KS1; ++ks) load_w1(c, ks);
          if (c >= 1) load_w2(c - 1);
        }
        load_w2(NC - 1);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    {
      constexpr uint32_t idesc1 = make_idesc_bf16(MF_BM, MF_HC, 0, 0);
      constexpr uint32_t idesc2 = make_idesc_bf16(MF_BM, NH, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      auto next_stage = [&]() {
        if (++stage == p.stages) {
          stage = 0;
          phase ^= 1;
        }
      };
      uint32_t it = 0;
      uint32_t gq = 0;   // chunks whose second GEMM has been issued (parity of g_full / g_empty)
      uint32_t gc = 0;   // chunks whose first GEMM has been issued (acc1 buffer + parity)
      for (int tile = blockIdx.x; tile < p.m_tiles; tile += gridDim.x, ++it) {
        auto gemm2 = [&](int cc) {
          mbar_wait(g_full, gq & 1);
          tcgen05_fence_after();
          if (kTrain && elect_one()) {   // the hidden activations of a training plan leave from the operand tile itself
            tma_store_2d(&p.tmG, g_base, (cc * 2) * 64, tile * MF_BM);
            tma_store_2d(&p.tmG, g_base + MF_BM * 128, (cc * 2 + 1) * 64, tile * MF_BM);
            tma_store_commit();
          }
          if (cc == 0) {
            mbar_wait(acc2_empty, (it & 1) ^ 1);
            tcgen05_fence_after();
          }
          for (int hs = 0; hs < 2; ++hs)
            for (int nh = 0; nh < 2; ++nh) {
              mbar_wait(&w_full[stage], phase);
              tcgen05_fence_after();
              const uint32_t a_addr = smem_u32(g_base + hs * (MF_BM * 128));
              const uint32_t b_addr = smem_u32(w_base + stage * MF_STAGE);
              if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma_f16_ss(tmem_acc2 + nh * NH, make_desc_kmajor_sw128(a_addr + k * 32),
                              make_desc_kmajor_sw128(b_addr + k * 32), idesc2, (cc | hs | k) != 0 ? 1u : 0u);
                umma_commit(&w_empty[stage]);
              }
              next_stage();
            }
          if (elect_one()) {
            if (kTrain) tma_store_wait_read<0>();   // the G tile is re-written once both its readers are done
            umma_commit(g_empty);
          }
          ++gq;
        };
        mbar_wait(a_full, it & 1);
        tcgen05_fence_after();
        for (int c = 0; c < NC; ++c, ++gc) {
          const uint32_t b1 = gc & 1;
          mbar_wait(&acc1_empty[b1], ((gc >> 1) & 1) ^ 1);
          tcgen05_fence_after();
          for (int ks = 0; ks < KS1; ++ks) {
            mbar_wait(&w_full[stage], phase);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_base + ks * (MF_BM * 128));
            const uint32_t b_addr = smem_u32(w_base + stage * MF_STAGE);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < 4; ++k)
                umma_f16_ss(tmem_acc1 + b1 * MF_HC, make_desc_kmajor_sw128(a_addr + k * 32),
                            make_desc_kmajor_sw128(b_addr + k * 32), idesc1, (ks | k) != 0 ? 1u : 0u);
              umma_commit(&w_empty[stage]);
              if (ks == KS1 - 1) {
                umma_commit(&acc1_full[b1]);
                if (c == NC - 1) umma_commit(a_empty);   // the X tile may be replaced once the last first-GEMM has read it
              }
            }
            next_stage();
          }
          if (c >= 1) gemm2(c - 1);
        }
        gemm2(NC - 1);
        if (elect_one()) umma_commit(acc2_full);
      }
      if (kTrain && elect_one()) tma_store_wait_all<0>();
    }
    __syncwarp();
  } else {
    // ================================ epilogue ================================
    const int ew = warp - 2;
    int tsn = 0;
    (void)tsn;
    const int q = warp & 3;    // TMEM lane quarter this warp may access
    const int hf = ew >> 2;    // column half
    const int row = q * 32 + lane;
    uint8_t* stg = stg_base + ew * MF_STG_WARP;
    uint64_t* ldbar = ld_bar + ew * 2;
    uint32_t ldphase = 0;
    int ridx = 0;
    int xslot = 0;
    const float inv_n = 1.0f / (float)p.n_valid;
    const int col_half = hf * HALF;
    const bool use_r = p.has_resid != 0;
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;

    auto exchange2 = [&](float a, float b, float& oa, float& ob) {
      float* s = xch + xslot * 512;
      s[(hf * 2 + 0) * 128 + row] = a;
      s[(hf * 2 + 1) * 128 + row] = b;
      named_bar_sync(1, 256);
      oa = a + s[((hf ^ 1) * 2 + 0) * 128 + row];
      ob = b + s[((hf ^ 1) * 2 + 1) * 128 + row];
      xslot ^= 1;
    };
    auto issue_load = [&](int bi, int tile_, int col) {
      mbar_arrive_expect_tx(&ldbar[bi], 4096u);
      tma_load_2d(stg + bi * 4096, &p.tmResid, &ldbar[bi], col, tile_ * MF_BM + q * 32);
    };
    // begin a slot of the 2-buffer ring: free the other buffer (its TMA store must have been read out), prefetch the
    // next slot's residual chunk into it, then wait for this slot's chunk
    auto slot_begin = [&](bool have_next, int ntile, int ncol) -> uint8_t* {
      if (elect_one()) {
        if (have_next && use_r) {
          tma_store_wait_read<0>();
          issue_load(ridx ^ 1, ntile, ncol);
        } else {
          tma_store_wait_read<1>();
        }
      }
      __syncwarp();
      if (use_r) {
        mbar_wait(&ldbar[ridx], (ldphase >> ridx) & 1u);
        ldphase ^= 1u << ridx;
      }
      return stg + ridx * 4096;
    };
    auto slot_end = [&](bool stored) {
      if (stored && elect_one()) tma_store_commit();
      ridx ^= 1;
    };
    auto add_row_f32 = [&](float (&v)[32], const uint8_t* buf) {
      const uint32_t base = smem_u32(buf) + lane * 128;
      const uint32_t sw = lane & 7;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float4 r4;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(r4.x), "=f"(r4.y), "=f"(r4.z), "=f"(r4.w)
                     : "r"(base + ((j ^ sw) << 4)));
        add2(v[4 * j], v[4 * j + 1], v[4 * j], v[4 * j + 1], r4.x, r4.y);
        add2(v[4 * j + 2], v[4 * j + 3], v[4 * j + 2], v[4 * j + 3], r4.z, r4.w);
      }
    };

    // training plans stage u through the same per-warp buffers during the hidden chunks, so the residual chunk of the
    // NEXT tile cannot be prefetched across the tile boundary there; it is requested when the output phase begins
    const bool xprefetch = !kTrain;
    int tile = blockIdx.x;
    if (use_r && xprefetch && tile < p.m_tiles && elect_one()) issue_load(0, tile, col_half);
    __syncwarp();
    uint32_t it = 0, gc = 0;
    for (; tile < p.m_tiles; tile += gridDim.x, ++it) {
      const int row0 = tile * MF_BM + q * 32;
      const long long tok = (long long)tile * MF_BM + row;
      const bool live = tok < p.M;
      const bool more_tiles = (tile + (int)gridDim.x) < p.m_tiles;
      const int tile_next = more_tiles ? tile + (int)gridDim.x : tile;
      float sa = 1.0f;
      if (live && p.acc_scale) sa = __ldg(p.acc_scale + tok / p.tokens_per_sample);

      // ---------------- hidden chunks: acc1 -> +b1 -> [u] -> GELU -> G tile ----------------
#pragma unroll 1
      for (int c = 0; c < NC; ++c, ++gc) {
        const uint32_t b1 = gc & 1;
        MF_STAMP();   // 0: chunk start
        mbar_wait(&acc1_full[b1], (gc >> 1) & 1);
        tcgen05_fence_after();
        MF_STAMP();   // 1: acc1 ready
        uint32_t pk[2][16];
#pragma unroll
        for (int g2 = 0; g2 < 2; ++g2) {
          const int cc = hf * 64 + g2 * 32;          // column inside the chunk
          const int hcol = c * MF_HC + cc;           // hidden column
          float v[32];
          {
            uint32_t r[32];
            tmem_ld32(tmem_acc1 + lane_sel + b1 * MF_HC + cc, r);
            tmem_ld_wait();
            const float4* bp = reinterpret_cast<const float4*>(bias1_s + hcol);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 b4 = bp[i];
              add2(v[4 * i], v[4 * i + 1], __uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]), b4.x, b4.y);
              add2(v[4 * i + 2], v[4 * i + 3], __uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]), b4.z, b4.w);
            }
          }
          if (kTrain) {   // pre-activation of a training plan: 2 KiB sub-buffers of this warp's staging, round robin
            uint8_t* ub = stg + ((gc * 2 + g2) & 3) * 2048;
            if (elect_one()) tma_store_wait_read<3>();
            __syncwarp();
            mf_stage_bf16(ub, lane, v);
            fence_proxy_async_smem();
            __syncwarp();
            if (elect_one()) {
              tma_store_2d(&p.tmU, ub, hcol, row0);
              tma_store_commit();
            }
          }
#pragma unroll
          for (int i = 0; i < 32; i += 2) gelu_erf2(v[i], v[i + 1]);
#pragma unroll
          for (int i = 0; i < 16; ++i) pk[g2][i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc1_empty[b1]);
        // G tile: K-major slice `hf` (64 hidden columns), 128-byte rows in 8-row SWIZZLE_128B atoms
        MF_STAMP();   // 2: bias + u + GELU done
        mbar_wait(g_empty, (gc & 1) ^ 1);
        MF_STAMP();   // 3: G tile free
        {
          const uint32_t base = smem_u32(g_base + hf * (MF_BM * 128)) + row * 128;
          const uint32_t sw = row & 7;
#pragma unroll
          for (int g2 = 0; g2 < 2; ++g2)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t a = base + (((g2 * 4 + j) ^ sw) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pk[g2][4 * j]),
                           "r"(pk[g2][4 * j + 1]), "r"(pk[g2][4 * j + 2]), "r"(pk[g2][4 * j + 3])
                           : "memory");
            }
        }
        {
          // the ones column of g (it carries fc2's bias gradient through the weight-gradient GEMM): one 2-byte store into
          // the operand tile by the thread that owns the row, after its own vector stores.  (As `if (col == ones) v = 1`
          // over the registers the compiler emitted 64 compare + select pairs per chunk: 12 % of an issue-bound loop.)
          const int oc = p.ones_col_g - (c * MF_HC + hf * 64);   // column inside this thread's 64-column slice
          if (oc >= 0 && oc < 64) {
            const uint32_t a = smem_u32(g_base + hf * (MF_BM * 128)) + row * 128 + ((((uint32_t)oc >> 3) ^ (row & 7)) << 4) +
                               ((uint32_t)oc & 7) * 2;
            asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((unsigned short)0x3F80) : "memory");
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(g_full);
      }
      if (kTrain) {   // the staging buffers are about to hold residual chunks again
        if (elect_one()) {
          tma_store_wait_read<0>();
          if (use_r) issue_load(ridx, tile, col_half);
        }
        __syncwarp();
      }

      // ---------------- output: acc2 -> +b2, DropPath, residual, [LayerNorm] ----------------
      MF_STAMP();   // 12: hidden chunks done
      mbar_wait(acc2_full, it & 1);
      tcgen05_fence_after();
      MF_STAMP();   // 13: acc2 ready
      const uint32_t tacc = tmem_acc2 + lane_sel + col_half;
      float s1 = 0.f;
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const int col = col_half + c * 32;
        const bool last = (c == NCH - 1);
        uint8_t* buf = slot_begin(!last || (more_tiles && xprefetch), last ? tile_next : tile, last ? col_half : col + 32);
        float v[32];
        {
          uint32_t r[32];
          tmem_ld32(tacc + c * 32, r);
          tmem_ld_wait();
          const float4* bp = reinterpret_cast<const float4*>(bias2_s + col);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 b4 = bp[i];
            add2(v[4 * i], v[4 * i + 1], __uint_as_float(r[4 * i]), __uint_as_float(r[4 * i + 1]), b4.x, b4.y);
            add2(v[4 * i + 2], v[4 * i + 3], __uint_as_float(r[4 * i + 2]), __uint_as_float(r[4 * i + 3]), b4.z, b4.w);
            mul2(v[4 * i], v[4 * i + 1], v[4 * i], v[4 * i + 1], sa, sa);
            mul2(v[4 * i + 2], v[4 * i + 3], v[4 * i + 2], v[4 * i + 3], sa, sa);
          }
        }
        if (use_r) add_row_f32(v, buf);
        if (kLN) {
          float p0 = 0.f, p1 = 0.f, p2 = 0.f, p3 = 0.f;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            add2(p0, p1, p0, p1, v[i], v[i + 1]);
            add2(p2, p3, p2, p3, v[i + 2], v[i + 3]);
          }
          s1 += (p0 + p1) + (p2 + p3);
          uint32_t r[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(v[i]);
          tmem_st32(tacc + c * 32, r);
        }
        bool stored = false;
        if (p.has_outF) {
          __syncwarp();   // every lane has consumed its residual row before the chunk is overwritten in place
          mf_stage_f32(buf, lane, v);
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) tma_store_2d(&p.tmOutF, buf, col, row0);
          stored = true;
        } else if (!kLN && p.has_outB) {
          // no LayerNorm (the block that closes an RSTB): v itself goes out as the bf16 operand of the next convolution
          if (p.ones_col >= col && p.ones_col < col + 32) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (col + i == p.ones_col) v[i] = 1.0f;
          }
          __syncwarp();
          mf_stage_bf16(buf, lane, v);
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) tma_store_2d(&p.tmOutB, buf, col, row0);
          stored = true;
        }
        slot_end(stored);
      }
      if (kLN) {
        tmem_st_wait();
        MF_STAMP();   // 14: pass A done
        float tot, dummy;
        exchange2(s1, 0.f, tot, dummy);
        MF_STAMP();   // 15: first exchange
        const float mean = tot * inv_n;
        float s2 = 0.f;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          uint32_t r[32];
          tmem_ld32(tacc + c * 32, r);
          tmem_ld_wait();
          float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            float d0, d1, d2, d3;
            add2(d0, d1, __uint_as_float(r[i]), __uint_as_float(r[i + 1]), -mean, -mean);
            add2(d2, d3, __uint_as_float(r[i + 2]), __uint_as_float(r[i + 3]), -mean, -mean);
            fma2(q0, q1, d0, d1, d0, d1, q0, q1);
            fma2(q2, q3, d2, d3, d2, d3, q2, q3);
          }
          s2 += (q0 + q1) + (q2 + q3);
        }
        float tot2;
        exchange2(s2, 0.f, tot2, dummy);
        tot2 -= (float)(N - p.n_valid) * mean * mean;   // padding columns hold exact zeros
        const float rstd = rsqrtf(fmaxf(tot2, 0.f) * inv_n + p.ln_eps);
        if (hf == 0 && live && p.rowstat) p.rowstat[tok] = rstd;
        if (p.has_outB) {
          // x-hat -> bf16 through the two 2 KiB halves of the ring buffer that is NOT receiving the next tile's residual
          uint8_t* free_buf = stg + (ridx ^ 1) * 4096;
          if (elect_one()) tma_store_wait_read<0>();
          __syncwarp();
#pragma unroll 1
          for (int c = 0; c < NCH; ++c) {
            const int col = col_half + c * 32;
            float v[32];
            {
              uint32_t r[32];
              tmem_ld32(tacc + c * 32, r);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                add2(v[i], v[i + 1], __uint_as_float(r[i]), __uint_as_float(r[i + 1]), -mean, -mean);
                mul2(v[i], v[i + 1], v[i], v[i + 1], rstd, rstd);
              }
              if (col + 32 > p.n_valid) {   // only the chunk that holds the padding columns pays for the selects
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                  const int cg = col + i;
                  if (cg >= p.n_valid) v[i] = cg == p.ones_col ? 1.0f : 0.0f;
                }
              }
            }
            uint8_t* bb = free_buf + (c & 1) * 2048;
            if (elect_one()) tma_store_wait_read<1>();
            __syncwarp();
            mf_stage_bf16(bb, lane, v);
            fence_proxy_async_smem();
            __syncwarp();
            if (elect_one()) {
              tma_store_2d(&p.tmOutB, bb, col, row0);
              tma_store_commit();
            }
          }
        }
      }
      if (kTrain) {   // the next tile stages u through these buffers with a relaxed wait: drain this tile's stores first
        if (elect_one()) tma_store_wait_read<0>();
        __syncwarp();
      }
      MF_STAMP();   // 16: output phase done
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc2_empty);
    }
    if (elect_one()) tma_store_wait_all<0>();
    __syncwarp();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

template <int N, bool kTrain, bool kLN>
int mf_launch_cfg(const MfParams& p, int grid, size_t smem, cudaStream_t st) {
  auto kern = mlp_fused_kernel<N, kTrain, kLN>;
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_set[dev & 63] = true;
  }
  static long long* ts_dev = nullptr;
  MfParams q = p;
  const bool want_ts = KB_DEV_ENV("KB_MF_TS") != nullptr;
  if (want_ts) {
    if (!ts_dev) cudaMalloc(&ts_dev, 512 * 8);
    cudaMemset(ts_dev, 0, 512 * 8);
    q.ts = ts_dev;
  }
  KB_CUDA(launch_pdl(kern, grid, MF_THREADS, smem, st, q));
  if (want_ts) {
    static long long h[512];
    cudaMemcpy(h, ts_dev, sizeof(h), cudaMemcpyDeviceToHost);
    fprintf(stderr, "mlp_fused:");
    for (int i = 0; i < 120 && h[i]; ++i) fprintf(stderr, " %lld", h[i] - h[0]);
    fprintf(stderr, "\n");
  }
  return 0;
}

template <int N>
int mf_launch(const MfParams& p, int grid, size_t smem, cudaStream_t st) {
  const bool train = p.has_u != 0;
  if (train && p.ln) return mf_launch_cfg<N, true, true>(p, grid, smem, st);
  if (train) return mf_launch_cfg<N, true, false>(p, grid, smem, st);
  if (p.ln) return mf_launch_cfg<N, false, true>(p, grid, smem, st);
  return mf_launch_cfg<N, false, false>(p, grid, smem, st);
}

}  // namespace

int mlp_fused_launch(const kair_mlp_fused_args& a, cudaStream_t st) {
  KB_CHECK(a.X && a.W1 && a.W2, "kair_mlp_fused: null operand");
  KB_CHECK(a.N == 64 || a.N == 192 || a.N == 256, "kair_mlp_fused: N=%d (padded channels) must be 64, 192 or 256", a.N);
  KB_CHECK(a.F == 2 * a.N, "kair_mlp_fused: hidden width F=%d must be 2*N (mlp_ratio 2, padded)", a.F);
  KB_CHECK(a.M > 0, "kair_mlp_fused: empty M");
  KB_CHECK(a.x_pitch % 8 == 0 && a.w1_pitch % 8 == 0 && a.w2_pitch % 8 == 0, "kair_mlp_fused: operand pitches must be "
           "multiples of 8 elements");
  KB_CHECK(a.outF || a.outB, "kair_mlp_fused: no output");
  KB_CHECK(a.ln || !(a.outF && a.outB), "kair_mlp_fused: without LayerNorm exactly one of outF / outB is written");
  KB_CHECK(!a.acc_scale || a.tokens_per_sample > 0, "kair_mlp_fused: tokens_per_sample required");
  KB_CHECK((a.outU != nullptr) == (a.outG != nullptr), "kair_mlp_fused: outU and outG come together (training plans)");
  if (check_device() != 0) return -3;

  MfParams p;
  memset(&p, 0, sizeof(p));
  p.bias1 = a.bias1;
  p.bias2 = a.bias2;
  p.acc_scale = a.acc_scale;
  p.rowstat = a.rowstat;
  p.tokens_per_sample = a.tokens_per_sample > 0 ? a.tokens_per_sample : 1;
  p.M = a.M;
  p.m_tiles = (a.M + MF_BM - 1) / MF_BM;
  p.n_valid = a.n_valid > 0 ? a.n_valid : a.N;
  p.ones_col = a.ones_col;
  p.ones_col_g = a.ones_col_g;
  p.ln = a.ln;
  p.ln_eps = a.ln_eps;
  p.has_resid = a.resid != nullptr;
  p.has_outF = a.outF != nullptr;
  p.has_outB = a.outB != nullptr;
  p.has_u = a.outU != nullptr;
  p.has_g = a.outG != nullptr;

  auto map2d = [&](CUtensorMap* tm, const void* ptr, uint64_t cols, uint64_t rows, long long pitch, int esz,
                   uint32_t bc, uint32_t br, CUtensorMapSwizzle sw) -> int {
    KB_CHECK((pitch * esz) % 16 == 0, "kair_mlp_fused: pitches must be multiples of 16 bytes");
    uint64_t dims[2] = {cols, rows};
    uint64_t str[1] = {(uint64_t)pitch * esz};
    uint32_t box[2] = {bc, br};
    return make_tmap(tm, esz == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, dims,
                     str, box, nullptr, sw);
  };
  const uint64_t M = (uint64_t)a.M, N = (uint64_t)a.N, F = (uint64_t)a.F;
  if (map2d(&p.tmA, a.X, N, M, a.x_pitch, 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return -2;
  if (map2d(&p.tmW1, a.W1, N, F, a.w1_pitch, 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return -2;
  if (map2d(&p.tmW2, a.W2, F, N, a.w2_pitch, 2, 64, (uint32_t)(a.N / 2), CU_TENSOR_MAP_SWIZZLE_128B)) return -2;
  if (a.resid && map2d(&p.tmResid, a.resid, N, M, a.resid_pitch, 4, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B)) return -2;
  if (a.outF && map2d(&p.tmOutF, a.outF, N, M, a.outF_pitch, 4, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B)) return -2;
  if (a.outB && map2d(&p.tmOutB, a.outB, N, M, a.outB_pitch, 2, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) return -2;
  if (a.outU && map2d(&p.tmU, a.outU, F, M, a.outU_pitch, 2, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) return -2;
  if (a.outG && map2d(&p.tmG, a.outG, F, M, a.outG_pitch, 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return -2;

  const int a_bytes = (a.N / 64) * MF_BM * 128;
  const int fixed = 8 * MF_STG_WARP + a_bytes + MF_G_BYTES + MF_MISC;
  int stages = (232448 - 1024 - fixed) / MF_STAGE;
  if (stages > MF_MAX_STAGES) stages = MF_MAX_STAGES;
  {
    const char* e = KB_DEV_ENV("KB_MF_STAGES");   // development aid: shrink the weight ring
    if (e && atoi(e) >= 2 && atoi(e) < stages) stages = atoi(e);
  }
  KB_CHECK(stages >= 2, "kair_mlp_fused: not enough shared memory for N=%d", a.N);
  p.stages = stages;
  const size_t smem = 1024 + (size_t)fixed + (size_t)stages * MF_STAGE;
  const int nsm = sm_count();
  const int grid = p.m_tiles < nsm ? p.m_tiles : nsm;
  if (a.N == 64) return mf_launch<64>(p, grid, smem, st);
  if (a.N == 192) return mf_launch<192>(p, grid, smem, st);
  return mf_launch<256>(p, grid, smem, st);
}

}  // namespace kb
