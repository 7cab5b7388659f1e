// kair_gemm_tn: persistent, warp-specialised tcgen05 GEMM  D[M,N] = epilogue(A[M,K] * B[N,K]^T)
//
//   warp 0      TMA producer   (A tiles [128 x 64] bf16, SWIZZLE_128B; conv mode: 9 shifted 4-D boxes per k-slice,
//                                zero padding comes from TMA out-of-bounds fill)
//   warp 1      MMA issuer     (one lane issues tcgen05.mma 128 x BN x 16, accumulators double-buffered in TMEM)
//   warp 10     second issuer  (narrow tiles only, BN <= 64: a 128 x 64 x 16 UMMA costs its issuing thread ~85 cycles but
//                                the tensor pipe ~48, so ONE thread cannot keep the pipe busy; the two issuers take
//                                alternate tiles -- issuer i owns TMEM accumulator stage i -- and walk the same operand
//                                rings with independently computed stage / phase sequences)
//   warps 2..9  epilogue       (tcgen05.ld -> registers -> fused bias / GELU / residual / LayerNorm /
//                                LayerNorm-backward -> swizzled smem staging -> TMA store)
//
// One thread of the epilogue owns one row of the 128-row tile (TMEM lane == row), split in two column halves across
// two warps, so LayerNorm statistics are thread-local plus one smem exchange.
//
// Replaces (reference file:line): nn.Linear qkv/proj/fc1/fc2 and their input-gradient GEMMs
// (models/network_swinir.py:25-28, 121, 143), LayerNorm norm1/norm2 (:245, :277), the residual adds (:276-277) and
// nn.Conv2d 3x3 (:465, :729; models/basicblock.py:65).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

constexpr int TN_THREADS = 352;       // 11 warps: TMA, MMA issuer 0, 8 epilogue, MMA issuer 1
constexpr int TN_BM = 128;
constexpr int TN_BK = 64;
constexpr int TN_A_STAGE = TN_BM * TN_BK * 2;  // 16 KiB
// per-epilogue-warp staging ring (runtime, TnParams::stg_warp): 12 KiB = 2 x (4 KiB fp32 + 2 KiB bf16 chunk) for the
// epilogues with in-place residual traffic, 6-8 KiB for the light ones so that more A stages fit
constexpr int TN_MAX_STAGES = 8;
constexpr int TN_EPI_LEAN = 5;                 // internal epilogue id (see the epilogue section)
constexpr int TN_MISC_BYTES = 6144;            // 1 KiB barriers + tmem ptr, 4 KiB row-exchange buffer

struct TnParams {
  CUtensorMap tmA, tmB, tmOutB, tmOutB2, tmOutF, tmResid, tmAux;
  const float* bias;
  const float* resid;
  const bf16* aux;
  float* rowstat;
  const float* acc_scale;
  const float* outB_scale;
  long long resid_pitch, aux_pitch;
  int tokens_per_sample;
  int M, m_tiles, n_tiles, kc, taps, stages;
  int stages_b;  // conv with streamed weights: depth of the separate B ring
  int a_stage;  // bytes of one A stage: 16 KiB (dense), 18 KiB (conv: 8 x 18 pixel halo column of one 64-channel slice)
  int n_valid, ones_col, act;
  int stg_warp, ringD, ringS, ringDist;
  int issuers;  // MMA issuer warps in use (1 or 2); the block has 320 + 32 * (issuers - 1) threads
  int ring_split;  // 2 issuers: 1 = each issuer owns one half of the operand rings, 0 = one shared ring (needs
                   // stages > stages per tile so that the parity waits of the issuer that runs ahead stay unambiguous)
  int dbg;  // development ablations (env KB_TN_DBG): 1 = skip tcgen05.ld in LINEAR, 2 = skip the MMAs,
            // 4 = LINEAR epilogue does the barrier handshake only, 8 = dense producer moves no A tiles,
            // 16 = DGELU without the gelu' arithmetic, 32 = DGELU without its stores
  float act_slope, ln_eps;
  int has_outB, has_outB2, has_outF;
  // conv geometry
  int H, W, TW, TH, tiles_w, tiles_h;
  int dil;  // tap dilation (1 = dense 3x3): halo of dil rows / columns, y taps are dil whole-atom offsets apart
};

struct TileCoord {
  int b, h0, w0;
};

template <bool kConv>
__device__ __forceinline__ TileCoord tile_coord(const TnParams& p, int m_tile) {
  TileCoord t;
  if (kConv) {
    int per_img = p.tiles_w * p.tiles_h;
    t.b = m_tile / per_img;
    int r = m_tile - t.b * per_img;
    int th = r / p.tiles_w;
    t.h0 = th * p.TH;
    t.w0 = (r - th * p.tiles_w) * p.TW;
  } else {
    t.b = 0;
    t.h0 = 0;
    t.w0 = m_tile * TN_BM;
  }
  return t;
}

// global token index of tile row `row` (or -1 when the row falls outside the tensor)
template <bool kConv>
__device__ __forceinline__ long long row_token(const TnParams& p, const TileCoord& t, int row) {
  if (kConv) {
    int h = t.h0 + row / p.TW;
    int w = t.w0 + row % p.TW;
    if (h >= p.H || w >= p.W) return -1;
    return ((long long)t.b * p.H + h) * p.W + w;
  } else {
    long long tok = (long long)t.w0 + row;
    return tok < p.M ? tok : -1;
  }
}

// ---- staging + TMA store of one 32-row x 32-col chunk owned by a warp --------------------------------------
__device__ __forceinline__ void stage_bf16(uint8_t* buf, int lane, const float (&v)[32]) {
  // 64-byte rows, SWIZZLE_64B: 16-byte unit j of row r lives at unit j ^ ((r >> 1) & 3)
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
__device__ __forceinline__ void stage_f32(uint8_t* buf, int lane, const float (&v)[32]) {
  // 128-byte rows, SWIZZLE_128B: 16-byte unit j of row r lives at unit j ^ (r & 7)
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
template <bool kConv>
__device__ __forceinline__ void tma_store_chunk(const TnParams& p, const CUtensorMap* tm, const uint8_t* buf,
                                                const TileCoord& t, int q, int col) {
  if (kConv) {
    int r0 = q * 32;
    tma_store_4d(tm, buf, col, t.w0 + r0 % p.TW, t.h0 + r0 / p.TW, t.b);
  } else {
    tma_store_2d(tm, buf, col, t.w0 + q * 32);
  }
}

#ifdef KB_DEVEL
#define TN_DBG(p) ((p).dbg)
#else
#define TN_DBG(p) 0
#endif

template <int BN, int EPI, bool kConv, bool kBRes>
__global__ void __launch_bounds__(TN_THREADS, 1) gemm_tn_kernel(const __grid_constant__ TnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);

  constexpr int B_STAGE = BN * TN_BK * 2;
  const int k_total = p.kc * p.taps;
  uint8_t* stg_base = smem;
  // conv mode: a stage is one x-shifted halo column [18 rows][8 px][64 ch] of the 8 x 16 pixel tile; its three
  // y-shifted 128-row windows (start + dy * 1024 B) feed the taps (dy, dx), so every input element is brought into
  // shared memory 3 times instead of 9 (the 64-channel-wide layers were bound by that fill rate)
  // Streamed (non-resident) weights in conv mode travel through their own ring of single [BN x 64] tiles, so the A
  // ring (stages x 18 KiB) and the B ring (stages_b x B_STAGE) are sized independently.
  uint8_t* a_base = smem + 8 * p.stg_warp;
  uint8_t* b_base = a_base + p.stages * p.a_stage;
  uint8_t* misc = b_base + (kBRes ? k_total : (kConv ? p.stages_b : p.stages)) * B_STAGE;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(misc);
  uint64_t* empty_bar = full_bar + TN_MAX_STAGES;
  uint64_t* tfull_bar = empty_bar + TN_MAX_STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint64_t* bres_bar = tempty_bar + 2;
  uint64_t* ld_bar = bres_bar + 1;  // [8 epilogue warps][up to 6 ring buffers]
  uint64_t* fullb_bar = ld_bar + 48;
  uint64_t* emptyb_bar = fullb_bar + TN_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(emptyb_bar + TN_MAX_STAGES);
  float* xch = reinterpret_cast<float*>(misc + 1024);  // [2 slots][2 halves][2 values][128 rows] fp32 = 4 KiB
  float* bias_s = reinterpret_cast<float*>(misc + 5120);  // this CTA's BN bias values (1 KiB)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  constexpr uint32_t TMEM_COLS = (2 * BN <= 128) ? 128 : (2 * BN <= 256 ? 256 : 512);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA);
    tma_prefetch_desc(&p.tmB);
    if (p.has_outB) tma_prefetch_desc(&p.tmOutB);
    if (p.has_outB2) tma_prefetch_desc(&p.tmOutB2);
    if (p.has_outF) tma_prefetch_desc(&p.tmOutF);
    if (p.resid) tma_prefetch_desc(&p.tmResid);
    if (p.aux) tma_prefetch_desc(&p.tmAux);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < p.stages; ++i) {
        mbar_init(&full_bar[i], 1);
        mbar_init(&empty_bar[i], 1);
      }
      for (int i = 0; i < TN_MAX_STAGES; ++i) {
        mbar_init(&fullb_bar[i], 1);
        mbar_init(&emptyb_bar[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&tfull_bar[i], 1);
        mbar_init(&tempty_bar[i], 8);
      }
      mbar_init(bres_bar, 1);
      for (int i = 0; i < 48; ++i) mbar_init(&ld_bar[i], 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  // barriers and the TMEM slot become visible here, BEFORE the wait on the previous kernel: the producer and the
  // issuers then start the moment the dependency resolves, and only the epilogue warps (the bias' sole readers) wait
  // for its global load (it used to sit in front of this barrier and delayed the first TMA load by a DRAM round trip)
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlaps the previous kernel's tail; no global access before this point
  pdl_launch_dependents();
  if (threadIdx.x >= 64 && threadIdx.x < 320) {
    if (p.bias != nullptr && threadIdx.x < 64 + BN)
      bias_s[threadIdx.x - 64] = __ldg(p.bias + (blockIdx.x % p.n_tiles) * BN + (threadIdx.x - 64));
    named_bar_sync(4, 256);
  }

  // static persistent schedule: this CTA keeps one n-tile and walks m-tiles with a fixed stride
  const int n_tile = blockIdx.x % p.n_tiles;
  const int m_first = blockIdx.x / p.n_tiles;
  const int m_stride = gridDim.x / p.n_tiles;

  if (warp == 0) {
    // ================================ TMA producer ================================
    // the whole warp runs the loops, one elected thread issues (see elect_one in common.cuh: `lane == 0` costs a
    // waterfall loop per TMA / MMA instruction)
    {
      if (kBRes && elect_one()) {
        mbar_arrive_expect_tx(bres_bar, (uint32_t)(k_total * B_STAGE));
        for (int kb = 0; kb < k_total; ++kb)
          tma_load_2d(b_base + kb * B_STAGE, &p.tmB, bres_bar, kb * TN_BK, n_tile * BN);
      }
      // With two MMA issuers the operand rings are split in halves: tiles alternate between the halves, each half is
      // consumed by one issuer in order, so the mbarrier parity protocol stays single-producer / single-consumer.
      const int nr = p.ring_split ? 2 : 1;
      const int sh = p.stages / nr, shb = (kConv && !kBRes) ? p.stages_b / nr : 0;   // stages per sub-ring
      int stage_r[2] = {0, 0}, stage_br[2] = {0, 0};
      uint32_t phase_r[2] = {0, 0}, phase_br[2] = {0, 0};
      int it = 0;
      for (int m_tile = m_first; m_tile < p.m_tiles; m_tile += m_stride, ++it) {
        const TileCoord t = tile_coord<kConv>(p, m_tile);
        const int r = nr == 2 ? (it & 1) : 0;
        int stage = stage_r[r], stage_b = stage_br[r];
        uint32_t phase = phase_r[r], phase_b = phase_br[r];
        const int s0 = r * sh, sb0 = r * shb;
        if (kConv) {
          for (int kb = 0; kb < p.kc * 3; ++kb) {
            const int kcc = kb / 3, dx = kb - kcc * 3;
            mbar_wait(&empty_bar[s0 + stage], phase ^ 1);
            if (elect_one()) {
              mbar_arrive_expect_tx(&full_bar[s0 + stage], (uint32_t)p.a_stage);
              tma_load_4d(a_base + (s0 + stage) * p.a_stage, &p.tmA, &full_bar[s0 + stage], kcc * TN_BK,
                          t.w0 + (dx - 1) * p.dil, t.h0 - p.dil, t.b);
            }
            if (!kBRes) {
              for (int dy = 0; dy < 3; ++dy) {
                mbar_wait(&emptyb_bar[sb0 + stage_b], phase_b ^ 1);
                if (elect_one()) {
                  mbar_arrive_expect_tx(&fullb_bar[sb0 + stage_b], (uint32_t)B_STAGE);
                  tma_load_2d(b_base + (sb0 + stage_b) * B_STAGE, &p.tmB, &fullb_bar[sb0 + stage_b],
                              ((dy * 3 + dx) * p.kc + kcc) * TN_BK, n_tile * BN);
                }
                if (++stage_b == shb) {
                  stage_b = 0;
                  phase_b ^= 1;
                }
              }
            }
            if (++stage == sh) {
              stage = 0;
              phase ^= 1;
            }
          }
        } else {
          for (int kb = 0; kb < k_total; ++kb) {
            mbar_wait(&empty_bar[s0 + stage], phase ^ 1);
            if (TN_DBG(p) & 8) {  // ablation: no A traffic
              if (elect_one()) mbar_arrive(&full_bar[s0 + stage]);
              if (++stage == sh) {
                stage = 0;
                phase ^= 1;
              }
              continue;
            }
            if (elect_one()) {
              mbar_arrive_expect_tx(&full_bar[s0 + stage], (uint32_t)(TN_A_STAGE + (kBRes ? 0 : B_STAGE)));
              tma_load_2d(a_base + (s0 + stage) * TN_A_STAGE, &p.tmA, &full_bar[s0 + stage], kb * TN_BK, t.w0);
              if (!kBRes)
                tma_load_2d(b_base + (s0 + stage) * B_STAGE, &p.tmB, &full_bar[s0 + stage], kb * TN_BK, n_tile * BN);
            }
            if (++stage == sh) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        stage_r[r] = stage; stage_br[r] = stage_b;
        phase_r[r] = phase; phase_br[r] = phase_b;
      }
    }
    __syncwarp();
  } else if (warp == 1 || warp == 10) {
    // ================================ MMA issuers ================================
    const int issuer = warp == 1 ? 0 : 1;
    if (issuer < p.issuers) {   // whole warp; the elected thread issues
      constexpr uint32_t idesc = make_idesc_bf16(TN_BM, BN, 0, 0);
      if (kBRes) {
        mbar_wait(bres_bar, 0);
        tcgen05_fence_after();
      }
      const int nr = p.ring_split ? 2 : 1;
      const int sh = p.stages / nr, shb = (kConv && !kBRes) ? p.stages_b / nr : 0;
      const int s0 = p.ring_split ? issuer * sh : 0, sb0 = p.ring_split ? issuer * shb : 0;
      const int a_per_tile = kConv ? p.kc * 3 : k_total;       // A ring stages of one tile
      const int b_per_tile = (kConv && !kBRes) ? p.kc * 9 : 0; // B ring stages of one tile (streamed conv weights)
      const bool shared2 = p.issuers == 2 && !p.ring_split;    // shared ring: step over the other issuer's tiles
      const uint32_t dil_bytes = (uint32_t)p.dil * 1024u;      // y taps: dil whole 8-pixel atoms apart
      int stage = 0, stage_b = 0;
      uint32_t phase = 0, phase_b = 0;
      auto skip_tile = [&]() {
        stage += a_per_tile;
        while (stage >= sh) {
          stage -= sh;
          phase ^= 1;
        }
        if (shb) {
          stage_b += b_per_tile;
          while (stage_b >= shb) {
            stage_b -= shb;
            phase_b ^= 1;
          }
        }
      };
      if (shared2 && issuer == 1) skip_tile();
      for (int it = issuer, m_tile = m_first + issuer * m_stride; m_tile < p.m_tiles;
           it += p.issuers, m_tile += p.issuers * m_stride) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty_bar[as], aphase ^ 1);
        tcgen05_fence_after();
        const uint32_t tacc = tmem_base + as * BN;
        if (kConv) {
          for (int kb = 0; kb < p.kc * 3; ++kb) {
            const int kcc = kb / 3, dx = kb - kcc * 3;
            mbar_wait(&full_bar[s0 + stage], phase);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_base + (s0 + stage) * p.a_stage);
#pragma unroll
            for (int dy = 0; dy < 3; ++dy) {
              const uint32_t a_dy = a_addr + dy * dil_bytes;
              uint32_t b_addr;
              if (kBRes) {
                b_addr = smem_u32(b_base + ((dy * 3 + dx) * p.kc + kcc) * B_STAGE);
              } else {
                mbar_wait(&fullb_bar[sb0 + stage_b], phase_b);
                tcgen05_fence_after();
                b_addr = smem_u32(b_base + (sb0 + stage_b) * B_STAGE);
              }
              if (elect_one()) {
#pragma unroll
                for (int k = 0; k < TN_BK / 16; ++k) {
                  if (TN_DBG(p) & 2) continue;
                  umma_f16_ss(tacc, make_desc_kmajor_sw128(a_dy + k * 32),
                              make_desc_kmajor_sw128(b_addr + k * 32), idesc, (kb | dy | k) != 0 ? 1u : 0u);
                }
                if (!kBRes) umma_commit(&emptyb_bar[sb0 + stage_b]);
                if (dy == 2) umma_commit(&empty_bar[s0 + stage]);
              }
              if (!kBRes) {
                if (++stage_b == shb) {
                  stage_b = 0;
                  phase_b ^= 1;
                }
              }
            }
            if (++stage == sh) {
              stage = 0;
              phase ^= 1;
            }
          }
        } else {
          for (int kb = 0; kb < k_total; ++kb) {
            mbar_wait(&full_bar[s0 + stage], phase);
            tcgen05_fence_after();
            const uint32_t a_addr = smem_u32(a_base + (s0 + stage) * TN_A_STAGE);
            const uint32_t b_addr = smem_u32(b_base + (kBRes ? kb : s0 + stage) * B_STAGE);
            if (elect_one()) {
#pragma unroll
              for (int k = 0; k < TN_BK / 16; ++k) {
                if (TN_DBG(p) & 2) continue;
                umma_f16_ss(tacc, make_desc_kmajor_sw128(a_addr + k * 32), make_desc_kmajor_sw128(b_addr + k * 32), idesc,
                            (kb | k) != 0 ? 1u : 0u);
              }
              umma_commit(&empty_bar[s0 + stage]);
              if (kb == k_total - 1) umma_commit(&tfull_bar[as]);
            }
            if (++stage == sh) {
              stage = 0;
              phase ^= 1;
            }
          }
        }
        if (kConv && elect_one()) umma_commit(&tfull_bar[as]);
        if (shared2) skip_tile();
      }
    }
    __syncwarp();
  } else if (warp < 10) {
    // ================================ epilogue ================================
    // Each epilogue warp owns two 6 KiB staging sets (4 KiB fp32 chunk + 2 KiB bf16 chunk of 32 rows x 32 columns).
    // Global operands of the epilogue (fp32 residual rows, bf16 aux rows) are TMA-loaded into the set one chunk
    // ("slot") ahead -- across tile boundaries too --, consumed from shared memory, overwritten in place with the
    // results and TMA-stored from the same set.  `resid` / `aux` must not alias any output of this launch.
    const int ew = warp - 2;
    const int q = warp & 3;    // TMEM lane quarter this warp may access
    const int hf = ew >> 2;    // column half
    const int row = q * 32 + lane;
    constexpr int HALF = BN / 2;
    constexpr int NCH = HALF / 32;
    uint8_t* stg = stg_base + ew * p.stg_warp;
    uint64_t* ldbar = ld_bar + ew * 6;
    uint32_t ldphase = 0;      // bit b: parity the next wait on ring buffer b expects
    int ridx = 0;              // ring buffer of the current slot
    int xslot = 0;
    const float inv_n = 1.0f / (float)p.n_valid;
    const int col_half = n_tile * BN + hf * HALF;
    // TN_EPI_LEAN: LINEAR with bias / ReLU (optional) and the bf16 output only -- qkv, the proj input gradient, the
    // DnCNN body convolutions: the option tests below fold away at compile time (the epilogue is instruction / branch bound:
    // ncu showed 14 % branch_resolving + 12 % no_inst stalls in the generic LINEAR loop)
    constexpr bool kLean = (EPI == TN_EPI_LEAN);
    const bool use_r = !kLean && p.resid != nullptr && (EPI == KAIR_EPI_LINEAR || EPI == KAIR_EPI_RESID_LN || EPI == KAIR_EPI_LNBWD);
    // LINEAR with aux: activation backward, v *= (aux > 0 ? 1 : act_slope) with aux = the saved activation output
    const bool use_a = !kLean && (EPI == KAIR_EPI_DGELU || EPI == KAIR_EPI_LNBWD || (EPI == KAIR_EPI_LINEAR && p.aux != nullptr));
    const bool has_ld = use_r || use_a;
    // The warp's 12 KiB of staging memory is a ring of ringD buffers of ringS bytes; slot k uses buffer k % ringD and
    // its operands are TMA-prefetched ringDist slots ahead.
    //   6 KiB x 2: fp32 chunk at +0 (residual in / outF out), bf16 chunk at +4096 (aux in / outB out)
    //   4 KiB x 3: GELU, u at +0 and gelu(u) at +2048          2 KiB x 6: a single bf16 chunk (store-only or in place)
    const int ringD = p.ringD, ringS = p.ringS, ringDist = p.ringDist;
    const int boff = (ringS == 6144) ? 4096 : 0;  // offset of the bf16 chunk inside a ring buffer
    auto ring_at = [&](int d) { int i = ridx + d; return i >= ringD ? i - ringD : i; };
    auto wait_stores = [&](int n) {  // at most n of this thread's latest TMA-store groups may still be reading smem
      if (n <= 0) tma_store_wait_read<0>();
      else if (n == 1) tma_store_wait_read<1>();
      else if (n == 2) tma_store_wait_read<2>();
      else tma_store_wait_read<2>();  // n >= 2 (stricter than needed for deeper rings)
    };

    // two-value exchange between the two threads that share a row
    auto exchange2 = [&](float a, float b, float& oa, float& ob) {
      float* s = xch + xslot * 512;
      s[(hf * 2 + 0) * 128 + row] = a;
      s[(hf * 2 + 1) * 128 + row] = b;
      named_bar_sync(1, 256);
      oa = a + s[((hf ^ 1) * 2 + 0) * 128 + row];
      ob = b + s[((hf ^ 1) * 2 + 1) * 128 + row];
      xslot ^= 1;
    };
    auto load_acc = [&](uint32_t taddr, float (&v)[32]) {
      uint32_t r[32];
      tmem_ld32(taddr, r);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
    };
    auto add_bias = [&](float (&v)[32], int col) {
      if (p.bias) {
        const float4* bp = reinterpret_cast<const float4*>(bias_s + (col - n_tile * BN));
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float4 b4 = bp[i];  // warp-uniform address: shared-memory broadcast
          add2(v[4 * i], v[4 * i + 1], v[4 * i], v[4 * i + 1], b4.x, b4.y);
          add2(v[4 * i + 2], v[4 * i + 3], v[4 * i + 2], v[4 * i + 3], b4.z, b4.w);
        }
      }
    };
    // lane 0: TMA-load the operands of one slot into staging set `set` (fp32 residual at +0, bf16 aux at +4096)
    auto issue_load = [&](int bi, bool want_r, bool want_a, const TileCoord& tt, int col) {
      uint8_t* b = stg + bi * ringS;
      mbar_arrive_expect_tx(&ldbar[bi], (want_r ? 4096u : 0u) + (want_a ? 2048u : 0u));
      if (kConv) {
        const int r0 = q * 32;
        if (want_r) tma_load_4d(b, &p.tmResid, &ldbar[bi], col, tt.w0 + r0 % p.TW, tt.h0 + r0 / p.TW, tt.b);
        if (want_a) tma_load_4d(b + boff, &p.tmAux, &ldbar[bi], col, tt.w0 + r0 % p.TW, tt.h0 + r0 / p.TW, tt.b);
      } else {
        if (want_r) tma_load_2d(b, &p.tmResid, &ldbar[bi], col, tt.w0 + q * 32);
        if (want_a) tma_load_2d(b + boff, &p.tmAux, &ldbar[bi], col, tt.w0 + q * 32);
      }
    };
    // Begin a slot: make sure the buffer that is about to be (re)filled is free (the TMA store that last used it has
    // been read out), prefetch the slot ringDist ahead into it, then wait for this slot's operands.
    auto slot_begin = [&](bool have_next, bool nr, bool na, const TileCoord& nt, int ncol, bool wait_mine) -> uint8_t* {
      if (elect_one()) {
        if (have_next && (nr || na)) {
          wait_stores(ringD - ringDist - 1);
          issue_load(ring_at(ringDist), nr, na, nt, ncol);
        } else {
          wait_stores(ringD - 1);
        }
      }
      __syncwarp();
      if (wait_mine) {
        mbar_wait(&ldbar[ridx], (ldphase >> ridx) & 1u);
        ldphase ^= 1u << ridx;
      }
      return stg + ridx * ringS;
    };
    auto slot_end = [&](bool stored) {
      if (stored && elect_one()) tma_store_commit();
      ridx = ring_at(1);
    };
    // this thread's row of a landed fp32 / bf16 chunk (same swizzles as stage_f32 / stage_bf16)
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
    auto read_row_bf16 = [&](float (&x)[32], const uint8_t* buf) {
      const uint32_t base = smem_u32(buf) + lane * 64;
      const uint32_t sw = (lane >> 1) & 3;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t w0, w1, w2, w3;
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3)
                     : "r"(base + ((j ^ sw) << 4)));
        const uint32_t w[4] = {w0, w1, w2, w3};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float2 f2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[e]));
          x[8 * j + 2 * e] = f2.x;
          x[8 * j + 2 * e + 1] = f2.y;
        }
      }
    };

    int it = 0;
    int m_tile = m_first;
    if (has_ld && m_tile < p.m_tiles && elect_one()) {
      const TileCoord t0 = tile_coord<kConv>(p, m_tile);
      if (EPI == KAIR_EPI_DGELU) {
        for (int c = 0; c < ringDist; ++c) issue_load(c, false, true, t0, col_half + c * 32);  // needs ringDist <= NCH
      } else {
        issue_load(0, use_r && EPI != KAIR_EPI_LNBWD, use_a, t0, col_half);              // slot 0 of the first tile
      }
    }
    __syncwarp();

    for (; m_tile < p.m_tiles; m_tile += m_stride, ++it) {
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const TileCoord t = tile_coord<kConv>(p, m_tile);
      const long long tok = row_token<kConv>(p, t, row);
      const bool more_tiles = (m_tile + m_stride) < p.m_tiles;
      const TileCoord t_next = tile_coord<kConv>(p, more_tiles ? m_tile + m_stride : m_tile);
      float sa = 1.0f, sb = 1.0f;  // DropPath per-sample scales
      if (!kLean && tok >= 0) {
        if (p.acc_scale) sa = __ldg(p.acc_scale + tok / p.tokens_per_sample);
        if (p.outB_scale) sb = __ldg(p.outB_scale + tok / p.tokens_per_sample);
      }
      mbar_wait(&tfull_bar[as], aphase);
      tcgen05_fence_after();
      const uint32_t tacc = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + hf * HALF;

      if (EPI == KAIR_EPI_LINEAR || kLean) {
#pragma unroll 1
        for (int c = 0; c < ((TN_DBG(p) & 4) ? 0 : NCH); ++c) {  // dbg 4: ablation, epilogue = barrier handshake only
          const int col = col_half + c * 32;
          const bool last = (c == NCH - 1);
          uint8_t* buf = slot_begin(!last || more_tiles, use_r, use_a, last ? t_next : t, last ? col_half : col + 32,
                                    use_r || use_a);
          float v[32];
          if (TN_DBG(p) & 1) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.f;
          } else {
            load_acc(tacc + c * 32, v);
          }
          add_bias(v, col);
          if (!kLean && p.acc_scale) {
#pragma unroll
            for (int i = 0; i < 32; i += 2) mul2(v[i], v[i + 1], v[i], v[i + 1], sa, sa);
          }
          if (use_a) {
            float u[32];
            read_row_bf16(u, buf + boff);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = u[i] > 0.f ? v[i] : v[i] * p.act_slope;
          }
          if (use_r) add_row_f32(v, buf);
          if (p.act == KAIR_ACT_LRELU) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = v[i] > 0.f ? v[i] : v[i] * p.act_slope;
          }
          if (!kLean && p.has_outF) stage_f32(buf, lane, v);
          if (kLean || p.has_outB) {
            if (!kLean && p.outB_scale) {
#pragma unroll
              for (int i = 0; i < 32; i += 2) mul2(v[i], v[i + 1], v[i], v[i + 1], sb, sb);
            }
            if (!kLean && p.ones_col >= col && p.ones_col < col + 32) {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (col + i == p.ones_col) v[i] = 1.0f;
            }
            stage_bf16(buf + boff, lane, v);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) {
            if (!kLean && p.has_outF) tma_store_chunk<kConv>(p, &p.tmOutF, buf, t, q, col);
            if (kLean || p.has_outB) tma_store_chunk<kConv>(p, &p.tmOutB, buf + boff, t, q, col);
          }
          slot_end(true);
        }
      } else if (EPI == KAIR_EPI_GELU) {
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          const int col = col_half + c * 32;
          uint8_t* buf = slot_begin(false, false, false, t, 0, false);
          float v[32];
          load_acc(tacc + c * 32, v);
          add_bias(v, col);
          if (p.has_outB2) stage_bf16(buf, lane, v);  // pre-activation u
#pragma unroll
          for (int i = 0; i < 32; i += 2) gelu_erf2(v[i], v[i + 1]);
          if (p.ones_col >= col && p.ones_col < col + 32) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (col + i == p.ones_col) v[i] = 1.0f;
          }
          stage_bf16(buf + 2048, lane, v);
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) {
            if (p.has_outB2) tma_store_chunk<kConv>(p, &p.tmOutB2, buf, t, q, col);
            tma_store_chunk<kConv>(p, &p.tmOutB, buf + 2048, t, q, col);
          }
          slot_end(true);
        }
      } else if (EPI == KAIR_EPI_RESID_LN) {
        // pass A: v = acc + bias + resid -> outF (in place over the landed residual chunk), keep v in TMEM
        float s1 = 0.f;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          const int col = col_half + c * 32;
          const bool last = (c == NCH - 1);
          uint8_t* buf = slot_begin(!last || more_tiles, use_r, false, last ? t_next : t, last ? col_half : col + 32,
                                    use_r);
          float v[32];
          load_acc(tacc + c * 32, v);
          add_bias(v, col);
          if (p.acc_scale) {
#pragma unroll
            for (int i = 0; i < 32; i += 2) mul2(v[i], v[i + 1], v[i], v[i + 1], sa, sa);
          }
          if (use_r) add_row_f32(v, buf);
          {  // four independent partial sums: a single 32-long FADD chain costs 32 x 4 dependent cycles
            float p0 = 0.f, p1 = 0.f, p2 = 0.f, p3 = 0.f;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              add2(p0, p1, p0, p1, v[i], v[i + 1]);
              add2(p2, p3, p2, p3, v[i + 2], v[i + 3]);
            }
            s1 += (p0 + p1) + (p2 + p3);
          }
          {
            uint32_t r[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(v[i]);
            tmem_st32(tacc + c * 32, r);
          }
          if (p.has_outF) {
            stage_f32(buf, lane, v);
            fence_proxy_async_smem();
            __syncwarp();
            if (elect_one()) tma_store_chunk<kConv>(p, &p.tmOutF, buf, t, q, col);
          }
          slot_end(p.has_outF);
        }
        tmem_st_wait();
        float tot, dummy;
        exchange2(s1, 0.f, tot, dummy);
        const float mean = tot * inv_n;
        // pass B: centred second moment (padding columns hold exact zeros -> subtract their mean^2 contribution)
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
        tot2 -= (float)(p.n_tiles * BN - p.n_valid) * mean * mean;
        const float rstd = rsqrtf(fmaxf(tot2, 0.f) * inv_n + p.ln_eps);
        if (hf == 0 && tok >= 0 && p.rowstat) p.rowstat[tok] = rstd;
        // pass C: xhat -> bf16, staged in the bf16 halves of the two sets (the fp32 halves may already be receiving
        // the next tile's residual chunk, so the slot counter is left alone)
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
          uint8_t* bb = stg + (c & 1) * 6144 + 4096;
          if (elect_one()) tma_store_wait_read<1>();
          __syncwarp();
          stage_bf16(bb, lane, v);
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) {
            tma_store_chunk<kConv>(p, &p.tmOutB, bb, t, q, col);
            tma_store_commit();
          }
        }
      } else if (EPI == KAIR_EPI_DGELU) {
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          const int col = col_half + c * 32;
          const int cn = c + ringDist;  // chunk prefetched now: ringDist slots ahead, possibly in the next tile
          const bool wrap = cn >= NCH;
          uint8_t* buf = slot_begin(!wrap || more_tiles, false, true, wrap ? t_next : t,
                                    col_half + (wrap ? cn - NCH : cn) * 32, true);
          float v[32], u[32];
          load_acc(tacc + c * 32, v);
          read_row_bf16(u, buf);
#pragma unroll
          if (TN_DBG(p) & 16) {   // ablation: no gelu' arithmetic
#pragma unroll
            for (int i = 0; i < 32; i += 2) mul2(v[i], v[i + 1], v[i], v[i + 1], u[i], u[i + 1]);
          } else {
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float g0, g1;
              gelu_erf_grad2(u[i], u[i + 1], g0, g1);
              mul2(v[i], v[i + 1], v[i], v[i + 1], g0, g1);
            }
          }
          stage_bf16(buf, lane, v);
          fence_proxy_async_smem();
          __syncwarp();
          if (!(TN_DBG(p) & 32) && elect_one()) tma_store_chunk<kConv>(p, &p.tmOutB, buf, t, q, col);   // 32: no store
          slot_end(true);
        }
      } else if (EPI == KAIR_EPI_LNBWD) {
        // slots of a tile: A0..A(NCH-1) need xhat (aux); B0..B(NCH-1) need xhat and the residual gradient
        float s1 = 0.f, s2 = 0.f;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          const int col = col_half + c * 32;
          const bool last = (c == NCH - 1);
          uint8_t* buf = slot_begin(true, last ? use_r : false, true, t, last ? col_half : col + 32, true);
          float x[32];
          read_row_bf16(x, buf + 4096);
          uint32_t r[32];
          tmem_ld32(tacc + c * 32, r);
          tmem_ld_wait();
          float p0 = 0.f, p1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float a0 = __uint_as_float(r[i]), a1 = __uint_as_float(r[i + 1]);
            add2(p0, p1, p0, p1, a0, a1);
            fma2(q0, q1, a0, a1, x[i], x[i + 1], q0, q1);
          }
          s1 += p0 + p1;
          s2 += q0 + q1;
          slot_end(false);
        }
        float t1, t2;
        exchange2(s1, s2, t1, t2);
        const float m1 = t1 * inv_n, m2 = t2 * inv_n;
        const float rstd = (tok >= 0) ? p.rowstat[tok] : 0.f;
#pragma unroll 1
        for (int c = 0; c < NCH; ++c) {
          const int col = col_half + c * 32;
          const bool last = (c == NCH - 1);
          uint8_t* buf = slot_begin(!last || more_tiles, last ? false : use_r, true, last ? t_next : t,
                                    last ? col_half : col + 32, true);
          float v[32];
          {
            float x[32];
            read_row_bf16(x, buf + 4096);
            uint32_t r[32];
            tmem_ld32(tacc + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; i += 2) {   // rstd * fma(-x, m2, dy - m1)
              float a0, a1;
              add2(a0, a1, __uint_as_float(r[i]), __uint_as_float(r[i + 1]), -m1, -m1);
              fma2(a0, a1, x[i], x[i + 1], -m2, -m2, a0, a1);
              mul2(v[i], v[i + 1], a0, a1, rstd, rstd);
            }
          }
          if (use_r) add_row_f32(v, buf);
#pragma unroll
          if (col + 32 > p.n_valid) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (col + i >= p.n_valid) v[i] = 0.f;
          }
          if (p.has_outF) stage_f32(buf, lane, v);
          if (p.has_outB) {
            if (p.outB_scale) {
#pragma unroll
              for (int i = 0; i < 32; i += 2) mul2(v[i], v[i + 1], v[i], v[i + 1], sb, sb);
            }
            stage_bf16(buf + 4096, lane, v);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) {
            if (p.has_outF) tma_store_chunk<kConv>(p, &p.tmOutF, buf, t, q, col);
            if (p.has_outB) tma_store_chunk<kConv>(p, &p.tmOutB, buf + 4096, t, q, col);
          }
          slot_end(true);
        }
      }
      // release the accumulator stage back to the MMA warp
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);  // plain SYNCS, no uniform operands
    }
    if (elect_one()) tma_store_wait_all<0>();
    __syncwarp();
  }

  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------ host side

template <int BN, int EPI, bool kConv, bool kBRes>
static int launch_cfg(const TnParams& p, int grid, size_t smem, cudaStream_t st) {
  auto kern = gemm_tn_kernel<BN, EPI, kConv, kBRes>;
  static bool attr_set[64] = {false};  // per instantiation and device; benign race (idempotent)
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_set[dev & 63] = true;
  }
  KB_CUDA(launch_pdl(kern, grid, p.issuers == 2 ? TN_THREADS : TN_THREADS - 32, smem, st, p));
  return 0;
}

template <int BN, int EPI, bool kConv>
static int launch_res(const TnParams& p, bool bres, int grid, size_t smem, cudaStream_t st) {
  if (bres) return launch_cfg<BN, EPI, kConv, true>(p, grid, smem, st);
  return launch_cfg<BN, EPI, kConv, false>(p, grid, smem, st);
}

// Instantiated configurations (kept explicit to bound compile time):
//   dense: every epilogue at BN=192 (SwinIR-M, C=180 padded to 192), BN=64 (SwinIR-S, C=60) and BN=256 (SwinIR-L,
//          C=240); LINEAR at BN 64/128/256
//   conv : LINEAR at BN 64/128/192/256; RESID_LN at BN=192, 64 and 256
static int launch_dispatch(const TnParams& p, int bn, int epi, bool conv, bool bres, int grid, size_t smem,
                           cudaStream_t st) {
#define KB_CASE(BN_, EPI_, CONV_) \
  if (bn == BN_ && epi == EPI_ && conv == CONV_) return launch_res<BN_, EPI_, CONV_>(p, bres, grid, smem, st);
  KB_CASE(192, TN_EPI_LEAN, false)
  KB_CASE(64, TN_EPI_LEAN, false)
  KB_CASE(192, TN_EPI_LEAN, true)
  KB_CASE(64, TN_EPI_LEAN, true)
  KB_CASE(192, KAIR_EPI_LINEAR, false)
  KB_CASE(192, KAIR_EPI_GELU, false)
  KB_CASE(192, KAIR_EPI_RESID_LN, false)
  KB_CASE(192, KAIR_EPI_DGELU, false)
  KB_CASE(192, KAIR_EPI_LNBWD, false)
  KB_CASE(64, KAIR_EPI_LINEAR, false)
  KB_CASE(128, KAIR_EPI_LINEAR, false)
  KB_CASE(256, KAIR_EPI_LINEAR, false)
  KB_CASE(64, KAIR_EPI_LINEAR, true)
  KB_CASE(128, KAIR_EPI_LINEAR, true)
  KB_CASE(192, KAIR_EPI_LINEAR, true)
  KB_CASE(256, KAIR_EPI_LINEAR, true)
  KB_CASE(192, KAIR_EPI_RESID_LN, true)
  // SwinIR-S (lightweight SR, embed_dim 60 -> 64 padded channels)
  KB_CASE(64, KAIR_EPI_GELU, false)
  KB_CASE(64, KAIR_EPI_RESID_LN, false)
  KB_CASE(64, KAIR_EPI_DGELU, false)
  KB_CASE(64, KAIR_EPI_LNBWD, false)
  KB_CASE(64, KAIR_EPI_RESID_LN, true)
  KB_CASE(256, KAIR_EPI_GELU, false)
  KB_CASE(256, KAIR_EPI_RESID_LN, false)
  KB_CASE(256, KAIR_EPI_DGELU, false)
  KB_CASE(256, KAIR_EPI_LNBWD, false)
  KB_CASE(256, KAIR_EPI_RESID_LN, true)
#undef KB_CASE
  set_error("kair_gemm_tn: configuration bn=%d mode=%d conv=%d is not instantiated", bn, epi, (int)conv);
  return -1;
}

static bool tn_is_lean(const kair_gemm_tn_args& a) {
  return a.mode == KAIR_EPI_LINEAR && a.outB && !a.outF && !a.resid && !a.aux && !a.acc_scale && !a.outB_scale &&
         a.ones_col < 0 && (a.bn == 192 || a.bn == 64);
}

int gemm_tn_launch(const kair_gemm_tn_args& a, cudaStream_t st) {
  KB_CHECK(a.A && a.B, "kair_gemm_tn: null operand");
  KB_CHECK(a.K > 0 && a.K % 64 == 0, "kair_gemm_tn: K=%d must be a positive multiple of 64", a.K);
  KB_CHECK(a.bn == 64 || a.bn == 128 || a.bn == 192 || a.bn == 256, "kair_gemm_tn: bn=%d unsupported", a.bn);
  KB_CHECK(a.N > 0 && a.N % a.bn == 0, "kair_gemm_tn: N=%d must be a multiple of bn=%d", a.N, a.bn);
  KB_CHECK(a.a_pitch % 8 == 0 && a.b_pitch % 8 == 0, "kair_gemm_tn: operand pitches must be multiples of 8 elements");
  KB_CHECK(a.M > 0, "kair_gemm_tn: empty M");
  if (a.mode == KAIR_EPI_RESID_LN || a.mode == KAIR_EPI_LNBWD)
    KB_CHECK(a.N == a.bn, "kair_gemm_tn: LayerNorm epilogues need the whole row in one tile (N == bn)");
  if (a.mode == KAIR_EPI_LNBWD || a.mode == KAIR_EPI_DGELU) KB_CHECK(a.aux != nullptr, "kair_gemm_tn: aux required");
  if (a.mode == KAIR_EPI_LNBWD) KB_CHECK(a.rowstat != nullptr, "kair_gemm_tn: rowstat required");
  if (a.mode == KAIR_EPI_LINEAR && a.aux) KB_CHECK(a.act == KAIR_ACT_NONE, "kair_gemm_tn: aux (activation backward) excludes act");
  if (a.acc_scale || a.outB_scale) KB_CHECK(a.tokens_per_sample > 0, "kair_gemm_tn: tokens_per_sample required");
  if (check_device() != 0) return -3;

  TnParams p;
  memset(&p, 0, sizeof(p));
  const int taps = a.conv ? 9 : 1;
  p.bias = a.bias;
  p.resid = a.resid;
  p.aux = (const bf16*)a.aux;
  p.rowstat = a.rowstat;
  p.acc_scale = a.acc_scale;
  p.outB_scale = a.outB_scale;
  p.tokens_per_sample = a.tokens_per_sample > 0 ? a.tokens_per_sample : 1;
  p.resid_pitch = a.resid_pitch;
  p.aux_pitch = a.aux_pitch;
  p.M = a.M;
  p.n_tiles = a.N / a.bn;
  p.kc = a.K / 64;
  p.taps = taps;
  p.n_valid = a.n_valid > 0 ? a.n_valid : a.N;
  p.ones_col = a.ones_col;
  p.act = a.act;
  p.act_slope = a.act_slope;
  p.ln_eps = a.ln_eps;
  p.has_outB = a.outB != nullptr;
  p.has_outB2 = a.outB2 != nullptr;
  p.has_outF = a.outF != nullptr;
  {
    const char* e = KB_DEV_ENV("KB_TN_DBG");
    p.dbg = e ? atoi(e) : 0;
  }


  p.dil = (a.conv && a.dilation > 1) ? a.dilation : 1;
  KB_CHECK(p.dil <= 8, "kair_gemm_tn: dilation=%d (max 8)", p.dil);
  if (a.conv) {
    KB_CHECK(a.batch > 0 && a.H > 0 && a.W > 0 && (long long)a.batch * a.H * a.W == a.M,
             "kair_gemm_tn: conv geometry %dx%dx%d does not match M=%d", a.batch, a.H, a.W, a.M);
    p.H = a.H;
    p.W = a.W;
    p.TW = 8;    // 8 x 16 pixel tiles: each 8-row UMMA atom is one tile row, so y shifts are whole-atom offsets
    p.TH = 16;
    p.tiles_w = (a.W + p.TW - 1) / p.TW;
    p.tiles_h = (a.H + p.TH - 1) / p.TH;
    p.m_tiles = a.batch * p.tiles_w * p.tiles_h;
    uint64_t dims[4] = {(uint64_t)a.K, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
    uint64_t str[3] = {(uint64_t)a.a_pitch * 2, (uint64_t)a.a_pitch * 2 * a.W, (uint64_t)a.a_pitch * 2 * a.W * a.H};
    uint32_t box[4] = {64, (uint32_t)p.TW, (uint32_t)(p.TH + 2 * p.dil), 1};
    if (make_tmap(&p.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a.A, dims, str, box, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  } else {
    p.m_tiles = (a.M + TN_BM - 1) / TN_BM;
    uint64_t dims[2] = {(uint64_t)a.K, (uint64_t)a.M};
    uint64_t str[1] = {(uint64_t)a.a_pitch * 2};
    uint32_t box[2] = {64, 128};
    if (make_tmap(&p.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.A, dims, str, box, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  }
  {
    uint64_t dims[2] = {(uint64_t)a.K * taps, (uint64_t)a.N};
    uint64_t str[1] = {(uint64_t)a.b_pitch * 2};
    uint32_t box[2] = {64, (uint32_t)a.bn};
    if (make_tmap(&p.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a.B, dims, str, box, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  }
  // output maps: per-warp 32-row x 32-col boxes
  auto make_out = [&](CUtensorMap* tm, const void* ptr, long long pitch, bool f32) -> int {
    const int esz = f32 ? 4 : 2;
    KB_CHECK((pitch * esz) % 16 == 0, "kair_gemm_tn: output pitch must be a multiple of 16 bytes");
    CUtensorMapDataType dt = f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    CUtensorMapSwizzle sw = f32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    if (a.conv) {
      uint64_t dims[4] = {(uint64_t)a.N, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
      uint64_t str[3] = {(uint64_t)pitch * esz, (uint64_t)pitch * esz * a.W, (uint64_t)pitch * esz * a.W * a.H};
      uint32_t bw = p.TW >= 32 ? 32 : p.TW;
      uint32_t box[4] = {32, bw, 32 / bw, 1};
      return make_tmap(tm, dt, 4, ptr, dims, str, box, nullptr, sw);
    } else {
      uint64_t dims[2] = {(uint64_t)a.N, (uint64_t)a.M};
      uint64_t str[1] = {(uint64_t)pitch * esz};
      uint32_t box[2] = {32, 32};
      return make_tmap(tm, dt, 2, ptr, dims, str, box, nullptr, sw);
    }
  };
  if (a.outB && make_out(&p.tmOutB, a.outB, a.outB_pitch, false)) return -2;
  if (a.outB2 && make_out(&p.tmOutB2, a.outB2, a.outB2_pitch, false)) return -2;
  if (a.outF && make_out(&p.tmOutF, a.outF, a.outF_pitch, true)) return -2;
  if (a.resid && make_out(&p.tmResid, a.resid, a.resid_pitch, true)) return -2;
  if (a.aux && make_out(&p.tmAux, a.aux, a.aux_pitch, false)) return -2;

  // shared memory plan
  const int k_total = p.kc * taps;
  const int b_stage = a.bn * 64 * 2;
  // staging ring per epilogue warp (see the kernel's epilogue comment)
  const bool loads_r = a.resid && (a.mode == KAIR_EPI_LINEAR || a.mode == KAIR_EPI_RESID_LN || a.mode == KAIR_EPI_LNBWD);
  p.stg_warp = 12288; p.ringD = 2; p.ringS = 6144;
  const bool lin_aux = a.mode == KAIR_EPI_LINEAR && a.aux != nullptr;
  p.ringDist = (loads_r || lin_aux || a.mode == KAIR_EPI_DGELU || a.mode == KAIR_EPI_LNBWD) ? 1 : 0;
  if (a.mode == KAIR_EPI_GELU) { p.stg_warp = 8192; p.ringD = 2; p.ringS = 4096; p.ringDist = 0; }
  if (a.mode == KAIR_EPI_DGELU) {  // prefetch distance must not exceed the chunks per thread (bn / 64)
    p.stg_warp = 8192; p.ringD = 4; p.ringS = 2048; p.ringDist = a.bn >= 128 ? 2 : 1;
  }
  if (a.mode == KAIR_EPI_LINEAR && !loads_r && !a.outF && !lin_aux) { p.stg_warp = 6144; p.ringD = 3; p.ringS = 2048; p.ringDist = 0; }
  const int stg_bytes = 8 * p.stg_warp;
  const int budget = 232448 - 1024 - stg_bytes - TN_MISC_BYTES;
  p.a_stage = a.conv ? (p.TH + 2 * p.dil) * p.TW * 128 : TN_A_STAGE;
  bool bres = false;
  int stages = 0, stages_b = 0;
  {
    long long bres_bytes = (long long)k_total * b_stage;
    int st_res = (int)((budget - bres_bytes) / p.a_stage);
    if (bres_bytes < budget && st_res >= 3) {
      bres = true;
      stages = st_res;
    } else if (a.conv) {
      // separate rings: 2-3 halo stages, the rest of the budget for streamed weight tiles
      stages = 3;
      stages_b = (budget - stages * p.a_stage) / b_stage;
      if (stages_b < 3) {
        stages = 2;
        stages_b = (budget - stages * p.a_stage) / b_stage;
      }
      if (stages_b > TN_MAX_STAGES) stages_b = TN_MAX_STAGES;
      KB_CHECK(stages_b >= 2, "kair_gemm_tn: not enough shared memory for streamed conv weights at bn=%d", a.bn);
    } else {
      stages = budget / (p.a_stage + b_stage);
    }
  }
  if (stages > TN_MAX_STAGES) stages = TN_MAX_STAGES;
  KB_CHECK(stages >= 2, "kair_gemm_tn: not enough shared memory for bn=%d", a.bn);
  p.stages = stages;
  p.stages_b = stages_b;
  // narrow tiles are bound by the issue rate of a single thread: a second issuer warp on its own halves of the rings
  p.issuers = 1;
  p.ring_split = 0;
  if (a.bn <= 64) {
    const int a_tile = a.conv ? p.kc * 3 : k_total, b_tile = (a.conv && !bres) ? p.kc * 9 : 0;
    if (stages > a_tile && (!b_tile || stages_b > b_tile)) {
      p.issuers = 2;                                   // one shared ring, deeper prefetch per issuer
    } else if (stages >= 4 && (!b_tile || stages_b >= 4)) {
      p.issuers = 2;
      p.ring_split = 1;
      p.stages &= ~1;                                  // equal halves
      p.stages_b &= ~1;
    }
  }
  {
    const char* e = KB_DEV_ENV("KB_TN_ISSUERS");
    if (e && atoi(e) == 1) {
      p.issuers = 1;
      p.ring_split = 0;
      p.stages = stages;
      p.stages_b = stages_b;
    }
  }
  size_t smem = 1024 + stg_bytes + (size_t)stages * p.a_stage +
                (size_t)(bres ? k_total : (a.conv ? stages_b : stages)) * b_stage + TN_MISC_BYTES;

  int nsm = sm_count();
  long long total_tiles = (long long)p.m_tiles * p.n_tiles;
  int grid = (int)(total_tiles < nsm ? total_tiles : nsm);
  grid = (grid / p.n_tiles) * p.n_tiles;
  KB_CHECK(grid >= p.n_tiles, "kair_gemm_tn: n_tiles=%d exceeds the SM count", p.n_tiles);

  return launch_dispatch(p, a.bn, tn_is_lean(a) ? TN_EPI_LEAN : a.mode, a.conv != 0, bres, grid, smem, st);
}

}  // namespace kb
