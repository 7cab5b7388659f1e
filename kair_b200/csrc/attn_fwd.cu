// kair_window_attn_fwd: fused shifted-window attention core on tcgen05 / TMEM.
//
//   out[t, h*32+d] = sum_j softmax_j( q_i . k_j + bias[h][rel(i,j)] + mask(i,j) ) * v_j        (q pre-scaled)
//
// One CTA iteration handles a PAIR of 8x8 windows (128 tokens = one UMMA M tile).  Window partition, the cyclic
// shift (torch.roll) and window reverse are pure TMA coordinates: each window is fetched / written as four 4x4-pixel
// quadrant boxes at ((wi*8 + qy*4 + shift) mod H, (wj*8 + qx*4 + shift) mod W), which never straddle the image
// border.  The shifted-window mask is evaluated in closed form (regions are constant per quadrant) and the
// relative-position bias is gathered from the (2ws-1)^2 x nH table held in shared memory.
//
//   warp 0       TMA producer: per head pair, Q/K/V tiles [128 tok x 64 ch] (SWIZZLE_128B), double buffered
//   warp 1       MMA issuer:   per window S_h = Q_h K_h^T (UMMA 64x64x32) and O_h = P_h V_h (64x32x64); the two windows
//                              of the pair are interleaved in TMEM lanes (M=64 atoms use 16 lanes per sub-partition,
//                              the second window is issued at lane offset 16)
//   warps 2..9   softmax:      two groups (even / odd heads); thread == query row; TMEM -> regs, bias + mask, exp2,
//                              P (bf16) -> swizzled smem; final 1/rowsum scaling, bf16 conversion, TMA scatter
//
// Replaces models/network_swinir.py:124-142 (q@k^T, bias gather :127-130, mask :132-135, softmax :136-138, attn@v)
// together with window_partition/window_reverse (:33-62), torch.roll (:249-252, :269-272) and calculate_mask
// (:216-237).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

// timeline stamps: role r (0 producer, 1 mma, 2 softmax warp 2 lane 0), up to 256 events each
#ifdef KB_DEVEL
#define AF_TS(role, cnt)                                                              \
  do {                                                                                \
    if (p.ts && blockIdx.x == 0 && (cnt) < 256) p.ts[(role) * 256 + (cnt)++] = clock64(); \
  } while (0)
#define AF_DBG(p) ((p).dbg)
#else   // production builds carry no timeline stamps / ablation branches inside the pipelines
#define AF_TS(role, cnt) \
  do {                   \
    (void)(cnt);         \
  } while (0)
#define AF_DBG(p) 0
#endif

constexpr int AF_THREADS = 320;
constexpr int AF_STAGE = 3 * 16384;     // Q, K, V tiles of one head pair
constexpr int AF_P_BYTES = 16384;       // [2 windows x 64 rows][64 keys] bf16, SWIZZLE_128B
constexpr int AF_MAX_HEADS = 8;

struct AttnFwdParams {
  CUtensorMap tmQKV, tmOut;      // 4x4-pixel quadrant boxes
  CUtensorMap tmQKV2, tmOut2;    // 4-wide x 8-tall half-window boxes (two quadrants per TMA operation)
  const float* table;  // [(2*8-1)^2][nH]
  int nH, H, W, nWh, nWw, shift, total_windows, pairs;
  long long* ts;  // development timeline buffer (env KB_AF_TS): [role][event] clock64 stamps of CTA 0
  int dbg;  // development ablations (env KB_AF_DBG): 1 no bias/mask/exp, 2 no output TMA stores, 4 no epilogue math
};

__device__ __forceinline__ void window_origin(const AttnFwdParams& p, int gw, int& b, int& wi, int& wj) {
  int nW = p.nWh * p.nWw;
  b = gw / nW;
  int r = gw - b * nW;
  wi = r / p.nWw;
  wj = r - wi * p.nWw;
}

__global__ void __launch_bounds__(AF_THREADS, 1) window_attn_fwd_kernel(const __grid_constant__ AttnFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* qkv_s = smem;                        // 2 stages x 48 KiB
  uint8_t* p_s = qkv_s + 2 * AF_STAGE;          // 2 x 32 KiB
  uint8_t* out_s = p_s + 2 * AF_P_BYTES;        // [128 x nH*32] bf16 as nH/2 atoms of 16 KiB
  uint8_t* misc = out_s + (p.nH / 2) * 16384;
  uint64_t* qk_full = reinterpret_cast<uint64_t*>(misc);  // Q,K tiles of a stage (dead once S is computed)
  uint64_t* qk_empty = qk_full + 2;
  uint64_t* v_full = qk_empty + 2;                        // V tile of a stage (dead once O is accumulated)
  uint64_t* v_empty = v_full + 2;
  uint64_t* s_full = v_empty + 2;
  uint64_t* s_empty = s_full + 2;
  uint64_t* p_full = s_empty + 2;
  uint64_t* p_empty = p_full + 2;
  uint64_t* o_full = p_empty + 2;
  uint64_t* o_empty = o_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_empty + 2);
  float* table_s = reinterpret_cast<float*>(misc + 256);  // [nH][225]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nH = p.nH;
  const int hpairs = nH / 2;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmQKV);
    tma_prefetch_desc(&p.tmOut);
    tma_prefetch_desc(&p.tmQKV2);
    tma_prefetch_desc(&p.tmOut2);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < 2; ++i) {
        mbar_init(&qk_full[i], 1);
        mbar_init(&qk_empty[i], 1);
        mbar_init(&v_full[i], 1);
        mbar_init(&v_empty[i], 1);
        mbar_init(&s_full[i], 1);
        mbar_init(&s_empty[i], 4);
        mbar_init(&p_full[i], 4);
        mbar_init(&p_empty[i], 1);
      }
      for (int i = 0; i < 2; ++i) {
        mbar_init(&o_full[i], 1);
        mbar_init(&o_empty[i], 8);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  // barriers / TMEM slot visible before the dependency wait; after it the producer and the issuer start at once and
  // only the softmax warps (the table's sole readers) wait for its global load
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // no global access above this point
  pdl_launch_dependents();
  if (warp >= 2) {
    // bias table -> smem as [h][idx].  All loads of a thread are issued before the first store; values are
    // pre-multiplied by log2(e) so that the softmax works in the exp2 domain with a single FFMA per score
    const int tid = threadIdx.x - 64;
    float tv[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = tid + u * 256;
      tv[u] = i < 225 * nH ? __ldg(p.table + i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int i = tid + u * 256;
      if (i < 225 * nH) {
        const int idx = i / nH, h = i - idx * nH;
        table_s[h * 225 + idx] = tv[u] * 1.4426950408889634f;
      }
    }
    named_bar_sync(3, 256);
  }
  const uint32_t tmem_S = tmem_base;         // 2 x 64 columns
  const uint32_t tmem_O = tmem_base + 128;   // nH x 32 columns, twice when they fit (nH <= 6)
  // O is double-buffered by window-pair parity when 128 + 2 * nH * 32 <= 512 TMEM columns: the softmax warps then run
  // the output epilogue of pair n-1 AFTER the first softmax of pair n, i.e. while the tensor pipe finishes P V of pair
  // n-1 and computes the next scores -- they used to sit ~3800 cycles per pair in the o_full wait (clock64 timeline)
  const bool o_dbl = 128 + 2 * nH * 32 <= 512;

  if (warp == 0) {
    // ================================ TMA producer ================================
    // whole warp runs the loops; the elected thread issues (elect_one, common.cuh)
    {
      int stage = 0;
      uint32_t phase = 0;
      int tsn = 0;
      for (int pair = blockIdx.x; pair < p.pairs; pair += gridDim.x) {
        const int gw0 = pair * 2;
        const int gw1 = gw0 + 1 < p.total_windows ? gw0 + 1 : gw0;
        // box coordinates of the pair's four half windows (in-window order is quadrant-major with quad = qx*2 + qy, so
        // the two quadrants of a half window are adjacent in shared memory: one 4x8 box per half unless the window
        // wraps around the bottom edge, then two 4x4 boxes)
        int wb[2], hw[2], h0[2], h1[2], ww[2][2];
#pragma unroll
        for (int wsel = 0; wsel < 2; ++wsel) {
          int wi, wj;
          window_origin(p, wsel ? gw1 : gw0, wb[wsel], wi, wj);
          hw[wsel] = p.shift > 0 && wi == p.nWh - 1;
          // (conditional subtraction instead of a run-time `%`: the shifted coordinates are below 2 H / 2 W)
          auto wrap = [](int v, int n) { return v >= n ? v - n : v; };
          h0[wsel] = wrap(wi * 8 + p.shift, p.H);
          h1[wsel] = wrap(h0[wsel] + 4, p.H);
          ww[wsel][0] = wrap(wj * 8 + p.shift, p.W);
          ww[wsel][1] = wrap(wj * 8 + 4 + p.shift, p.W);
        }
        auto load_tile = [&](uint8_t* dst, uint64_t* bar, int col) {
#pragma unroll
          for (int wsel = 0; wsel < 2; ++wsel) {
#pragma unroll
            for (int qx = 0; qx < 2; ++qx) {
              uint8_t* d = dst + (wsel * 4 + qx * 2) * 2048;
              if (!hw[wsel]) {
                tma_load_4d(d, &p.tmQKV2, bar, col, ww[wsel][qx], h0[wsel], wb[wsel]);
              } else {
                tma_load_4d(d, &p.tmQKV, bar, col, ww[wsel][qx], h0[wsel], wb[wsel]);
                tma_load_4d(d + 2048, &p.tmQKV, bar, col, ww[wsel][qx], h1[wsel], wb[wsel]);
              }
            }
          }
        };
        for (int hp = 0; hp < hpairs; ++hp) {
          uint8_t* st = qkv_s + stage * AF_STAGE;
          mbar_wait(&qk_empty[stage], phase ^ 1);
          AF_TS(0, tsn);
          if (elect_one()) {
            mbar_arrive_expect_tx(&qk_full[stage], 32768);
            load_tile(st, &qk_full[stage], hp * 64);
            load_tile(st + 16384, &qk_full[stage], nH * 32 + hp * 64);
          }
          mbar_wait(&v_empty[stage], phase ^ 1);
          if (elect_one()) {
            mbar_arrive_expect_tx(&v_full[stage], 16384);
            load_tile(st + 32768, &v_full[stage], 2 * nH * 32 + hp * 64);
          }
          AF_TS(0, tsn);
          if (++stage == 2) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    // Software-pipelined by one head-pair step: S(n+1) = Q K^T is issued before waiting for the probabilities of step
    // n, so the softmax warps always find their next score tile ready while O(n) = P V is still being accumulated.
    {
      constexpr uint32_t idesc_s = make_idesc_bf16(64, 64, 0, 0);
      constexpr uint32_t idesc_pv = make_idesc_bf16(64, 32, 0, 1);
      uint32_t s_use[2] = {0, 0}, p_use[2] = {0, 0};  // per S / P buffer use counters
      int s_stage = 0;
      uint32_t s_phase = 0;
      int tsn = 0;
      auto issue_S = [&]() {
        mbar_wait(&qk_full[s_stage], s_phase);
        AF_TS(1, tsn);
        tcgen05_fence_after();
        const uint32_t q_addr = smem_u32(qkv_s + s_stage * AF_STAGE);
        const uint32_t k_addr = q_addr + 16384;
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          mbar_wait(&s_empty[b], (s_use[b] & 1) ^ 1);
          tcgen05_fence_after();
          if (elect_one()) {
#pragma unroll
            for (int w = 0; w < 2; ++w) {
#pragma unroll
              for (int k = 0; k < 2; ++k) {
                umma_f16_ss(tmem_S + b * 64 + ((uint32_t)(w * 16) << 16),
                            make_desc_kmajor_sw128(q_addr + w * 8192 + b * 64 + k * 32),
                            make_desc_kmajor_sw128(k_addr + w * 8192 + b * 64 + k * 32), idesc_s, k);
              }
            }
            umma_commit(&s_full[b]);
            if (b == 1) umma_commit(&qk_empty[s_stage]);  // Q and K of this stage may be overwritten once S is done
          }
          s_use[b]++;
        }
        AF_TS(1, tsn);
        if (++s_stage == 2) {
          s_stage = 0;
          s_phase ^= 1;
        }
      };
      int stage = 0;
      uint32_t v_phase = 0;
      uint32_t pair_it = 0;
      if (blockIdx.x < p.pairs) issue_S();
      for (int pair = blockIdx.x; pair < p.pairs; pair += gridDim.x, ++pair_it) {
        const bool last_pair = pair + (int)gridDim.x >= p.pairs;
        const int ob = o_dbl ? (int)(pair_it & 1) : 0;                       // O buffer of this pair
        const uint32_t oph = o_dbl ? ((pair_it >> 1) & 1) : (pair_it & 1);   // its barrier phase
        const uint32_t tO = tmem_O + ob * nH * 32;
        for (int hp = 0; hp < hpairs; ++hp) {
          // issue the next step's S now if its Q/K have landed, otherwise after this step's P V
          const bool more = !(last_pair && hp == hpairs - 1);
          bool s_done = !more;
          if (more && __any_sync(0xffffffffu, mbar_try_wait(&qk_full[s_stage], s_phase))) {
            issue_S();
            s_done = true;
          }
          const uint32_t v_addr = smem_u32(qkv_s + stage * AF_STAGE) + 32768;
          mbar_wait(&v_full[stage], v_phase);
          tcgen05_fence_after();
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            const int h = hp * 2 + b;
            mbar_wait(&p_full[b], p_use[b] & 1);
            AF_TS(1, tsn);
            tcgen05_fence_after();
            if (h == 0) {
              mbar_wait(&o_empty[ob], oph ^ 1);
              tcgen05_fence_after();
            }
            const uint32_t pa = smem_u32(p_s + b * AF_P_BYTES);
            if (elect_one()) {
#pragma unroll
              for (int w = 0; w < 2; ++w) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                  umma_f16_ss(tO + h * 32 + ((uint32_t)(w * 16) << 16),
                              make_desc_kmajor_sw128(pa + w * 8192 + kk * 32),
                              make_desc_mnmajor_sw128(v_addr + w * 8192 + kk * 2048 + b * 64, 64), idesc_pv, kk);
                }
              }
              umma_commit(&p_empty[b]);
              if (b == 1) {
                umma_commit(&v_empty[stage]);
                if (hp == hpairs - 1) umma_commit(&o_full[ob]);  // before a possibly blocking wait for the next pair's Q/K
              }
            }
            p_use[b]++;
          }
          AF_TS(1, tsn);
          if (!s_done) issue_S();
          if (++stage == 2) {
            stage = 0;
            v_phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ================================ softmax / epilogue ================================
    const int q = warp & 3;
    const int g = (warp - 2) >> 2;  // head parity handled by this group
    // M=64 accumulators: TMEM lane 32q + l holds row 16q + (l & 15) of window (l >> 4)
    const int wsel = lane >> 4;
    const int i = q * 16 + (lane & 15);  // in-window index, quadrant-major (== quadrant q)
    const int row = wsel * 64 + i;       // row of the 128-row smem tiles
    const int quad_i = i >> 4;           // = qx*2 + qy
    const int ry = (quad_i & 1) * 4 + ((i & 15) >> 2);
    const int rx = (quad_i >> 1) * 4 + (i & 3);
    const int base_i = ry * 15 + rx;
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    uint32_t use = 0;
    uint32_t pair_it = 0;
    int tsn = 0;
    const bool tsw = (warp == 2 && lane == 0);
    const float LOG2E = 1.4426950408889634f;

    // ---- epilogue of one window pair: O * 1/rowsum -> bf16 -> staging -> TMA scatter to the windows' pixels
    auto epilogue = [&](int e_pair, uint32_t e_it, const float (&inv)[AF_MAX_HEADS / 2]) {
      const int gw0 = e_pair * 2;
      const bool second_valid = gw0 + 1 < p.total_windows;
      const int ob = o_dbl ? (int)(e_it & 1) : 0;
      const uint32_t tO = tmem_O + ob * nH * 32;
      mbar_wait(&o_full[ob], o_dbl ? ((e_it >> 1) & 1) : (e_it & 1));
      if (tsw) AF_TS(2, tsn);
      tcgen05_fence_after();
      if (elect_one()) tma_store_wait_read<0>();  // this warp's share of the previous pair's output stores
      named_bar_sync(2, 256);
      for (int hh = 0; hh < nH / 2; ++hh) {
        if (AF_DBG(p) & 4) break;
        const int h = hh * 2 + g;
        uint32_t r[32];
        tmem_ld32(tO + lane_sel + h * 32, r);
        tmem_ld_wait();
        const float sc = inv[hh];
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; j += 2)
          mul2(v[j], v[j + 1], __uint_as_float(r[j]), __uint_as_float(r[j + 1]), sc, sc);
        const uint32_t base = smem_u32(out_s + (h >> 1) * 16384) + row * 128;
        const uint32_t sw = row & 7;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t a = base + ((((h & 1) * 4 + c) ^ sw) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pack_bf16x2(v[8 * c], v[8 * c + 1])),
                       "r"(pack_bf16x2(v[8 * c + 2], v[8 * c + 3])), "r"(pack_bf16x2(v[8 * c + 4], v[8 * c + 5])),
                       "r"(pack_bf16x2(v[8 * c + 6], v[8 * c + 7]))
                       : "memory");
        }
      }
      fence_proxy_async_smem();
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_empty[ob]);
      named_bar_sync(2, 256);
      if (elect_one()) {
        // the pair's (window, half, head-pair) output boxes are spread over the eight epilogue warps
        const int ew = warp - 2;
        const int nops = (second_valid ? 2 : 1) * 2 * hpairs;
        for (int op = ew; op < nops; op += 8) {
          const int a = op % hpairs, hq = op / hpairs;
          const int ws2 = hq >> 1, qx = hq & 1;
          int b2, wi2, wj2;
          window_origin(p, gw0 + ws2, b2, wi2, wj2);
          const bool hwrap = p.shift > 0 && wi2 == p.nWh - 1;
          int h0 = wi2 * 8 + p.shift, ww2 = wj2 * 8 + qx * 4 + p.shift;
          if (h0 >= p.H) h0 -= p.H;
          if (ww2 >= p.W) ww2 -= p.W;
          const uint8_t* src = out_s + a * 16384 + (ws2 * 4 + qx * 2) * 2048;
          if (!hwrap) {
            tma_store_4d(&p.tmOut2, src, a * 64, ww2, h0, b2);
          } else {
            tma_store_4d(&p.tmOut, src, a * 64, ww2, h0, b2);
            tma_store_4d(&p.tmOut, src + 2048, a * 64, ww2, h0 + 4 >= p.H ? h0 + 4 - p.H : h0 + 4, b2);
          }
        }
        tma_store_commit();
      }
      if (tsw) AF_TS(2, tsn);
    };
    float inv_prev[AF_MAX_HEADS / 2];
    int prev_pair = -1;
    uint32_t prev_it = 0;

    for (int pair = blockIdx.x; pair < p.pairs; pair += gridDim.x, ++pair_it) {
      const int gw0 = pair * 2;
      const bool second_valid = gw0 + 1 < p.total_windows;
      const int gw = (wsel && second_valid) ? gw0 + 1 : gw0;
      int wb, wi, wj;
      window_origin(p, gw, wb, wi, wj);
      const bool lastrow = p.shift > 0 && wi == p.nWh - 1;
      const bool lastcol = p.shift > 0 && wj == p.nWw - 1;
      float madd[4];
#pragma unroll
      for (int qj = 0; qj < 4; ++qj) {
        bool m = (lastrow && ((qj & 1) != (quad_i & 1))) || (lastcol && ((qj >> 1) != (quad_i >> 1)));
        madd[qj] = m ? -100.0f * 1.4426950408889634f : 0.0f;
      }
      float inv_sum[AF_MAX_HEADS / 2];

      for (int hh = 0; hh < nH / 2; ++hh, ++use) {
        const int h = hh * 2 + g;
        const float* tb = table_s + h * 225 + base_i;
        mbar_wait(&s_full[g], use & 1);
        if (tsw) AF_TS(2, tsn);
        tcgen05_fence_after();
        float s[64];
        {
          uint32_t r0[32], r1[32];
          tmem_ld32(tmem_S + lane_sel + g * 64, r0);
          tmem_ld32(tmem_S + lane_sel + g * 64 + 32, r1);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            s[j] = __uint_as_float(r0[j]);
            s[32 + j] = __uint_as_float(r1[j]);
          }
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_empty[g]);
        if (tsw) AF_TS(2, tsn);
        // scores in the exp2 domain: t = s*log2(e) + bias*log2(e) (+ mask); 4-way split max / sum chains for ILP
        // (fp32 math packed two lanes per instruction: FFMA2 / FADD2; the phase is issue bound)
        const bool masked = lastrow || lastcol;
#pragma unroll
        for (int j = 0; j < 64; j += 2) {
          const int qj = j >> 4;
          const int jy = (qj & 1) * 4 + ((j & 15) >> 2);
          const int jx = (qj >> 1) * 4 + (j & 3);
          const int o = (7 - jy) * 15 + (7 - jx);   // column j + 1 is one pixel to the right: table offset o - 1
          fma2(s[j], s[j + 1], s[j], s[j + 1], LOG2E, LOG2E, tb[o], tb[o - 1]);
        }
        if (masked) {
#pragma unroll
          for (int j = 0; j < 64; j += 2) add2(s[j], s[j + 1], s[j], s[j + 1], madd[j >> 4], madd[j >> 4]);
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < 64; ++j) mx4[j & 3] = fmaxf(mx4[j & 3], s[j]);
        const float nmb = -fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        float sum4[4] = {0.f, 0.f, 0.f, 0.f};
        uint32_t pk[32];
#pragma unroll
        for (int j = 0; j < 64; j += 4) {
          add2(s[j], s[j + 1], s[j], s[j + 1], nmb, nmb);
          add2(s[j + 2], s[j + 3], s[j + 2], s[j + 3], nmb, nmb);
#pragma unroll
          for (int u = 0; u < 4; ++u) s[j + u] = ex2_approx(s[j + u]);
          pk[j >> 1] = pack_bf16x2(s[j], s[j + 1]);
          pk[(j >> 1) + 1] = pack_bf16x2(s[j + 2], s[j + 3]);
          // the row sum is taken over the unrounded exponentials: against the sum of the bf16 values the tensor core
          // sees it differs by ~2^-9 / sqrt(64) relative, an order below the bf16 rounding of the output itself
          add2(sum4[0], sum4[1], sum4[0], sum4[1], s[j], s[j + 1]);
          add2(sum4[2], sum4[3], sum4[2], sum4[3], s[j + 2], s[j + 3]);
        }
        const float sum = (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
        inv_sum[hh] = 1.0f / sum;
        if (tsw) AF_TS(2, tsn);
        mbar_wait(&p_empty[g], (use & 1) ^ 1);
        if (tsw) AF_TS(2, tsn);
        {
          const uint32_t base = smem_u32(p_s + g * AF_P_BYTES) + row * 128;
          const uint32_t sw = row & 7;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint32_t a = base + ((c ^ sw) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pk[4 * c]), "r"(pk[4 * c + 1]),
                         "r"(pk[4 * c + 2]), "r"(pk[4 * c + 3])
                         : "memory");
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[g]);
        if (tsw) AF_TS(2, tsn);
        if (hh == 0 && prev_pair >= 0) {   // the previous pair's output, under this pair's tensor work
          epilogue(prev_pair, prev_it, inv_prev);
          prev_pair = -1;
        }
      }

      if (!o_dbl) {
        epilogue(pair, pair_it, inv_sum);
      } else {   // deferred: runs after the first softmax of the next pair (or after the loop)
#pragma unroll
        for (int u = 0; u < AF_MAX_HEADS / 2; ++u) inv_prev[u] = inv_sum[u];
        prev_pair = pair;
        prev_it = pair_it;
      }
    }
    if (prev_pair >= 0) epilogue(prev_pair, prev_it, inv_prev);
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

int window_attn_fwd_launch(const kair_window_attn_args& a, cudaStream_t st) {
  KB_CHECK(a.qkv && a.out && a.table, "kair_window_attn_fwd: null pointer");
  KB_CHECK(a.window == 8, "kair_window_attn_fwd: only 8x8 windows are supported by the tcgen05 path (got %d)", a.window);
  KB_CHECK(a.num_heads >= 2 && a.num_heads <= AF_MAX_HEADS && a.num_heads % 2 == 0,
           "kair_window_attn_fwd: num_heads=%d must be even and <= %d", a.num_heads, AF_MAX_HEADS);
  KB_CHECK(a.H % 8 == 0 && a.W % 8 == 0 && a.H > 0 && a.W > 0, "kair_window_attn_fwd: H,W must be multiples of 8");
  KB_CHECK(a.shift == 0 || a.shift == 4, "kair_window_attn_fwd: shift must be 0 or 4");
  KB_CHECK(a.qkv_pitch % 8 == 0 && a.out_pitch % 8 == 0, "kair_window_attn_fwd: pitches must be multiples of 8");
  KB_CHECK(a.qkv_pitch >= 3 * a.num_heads * 32 && a.out_pitch >= a.num_heads * 32, "kair_window_attn_fwd: pitch too small");
  if (check_device() != 0) return -3;

  AttnFwdParams p;
  memset(&p, 0, sizeof(p));
  p.table = a.table;
  p.nH = a.num_heads;
  p.H = a.H;
  p.W = a.W;
  p.nWh = a.H / 8;
  p.nWw = a.W / 8;
  p.shift = a.shift;
  p.total_windows = a.batch * p.nWh * p.nWw;
  p.pairs = (p.total_windows + 1) / 2;
  {
    const char* e = KB_DEV_ENV("KB_AF_DBG");
    p.dbg = e ? atoi(e) : 0;
  }
  uint32_t box[4] = {64, 4, 4, 1};
  {
    uint64_t dims[4] = {(uint64_t)3 * a.num_heads * 32, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
    uint64_t str[3] = {(uint64_t)a.qkv_pitch * 2, (uint64_t)a.qkv_pitch * 2 * a.W, (uint64_t)a.qkv_pitch * 2 * a.W * a.H};
    if (make_tmap(&p.tmQKV, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a.qkv, dims, str, box, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  }
  {
    uint64_t dims[4] = {(uint64_t)a.num_heads * 32, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
    uint64_t str[3] = {(uint64_t)a.out_pitch * 2, (uint64_t)a.out_pitch * 2 * a.W, (uint64_t)a.out_pitch * 2 * a.W * a.H};
    if (make_tmap(&p.tmOut, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a.out, dims, str, box, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  }
  uint32_t box2[4] = {64, 4, 8, 1};
  {
    uint64_t dims[4] = {(uint64_t)3 * a.num_heads * 32, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
    uint64_t str[3] = {(uint64_t)a.qkv_pitch * 2, (uint64_t)a.qkv_pitch * 2 * a.W, (uint64_t)a.qkv_pitch * 2 * a.W * a.H};
    if (make_tmap(&p.tmQKV2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a.qkv, dims, str, box2, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  }
  {
    uint64_t dims[4] = {(uint64_t)a.num_heads * 32, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
    uint64_t str[3] = {(uint64_t)a.out_pitch * 2, (uint64_t)a.out_pitch * 2 * a.W, (uint64_t)a.out_pitch * 2 * a.W * a.H};
    if (make_tmap(&p.tmOut2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a.out, dims, str, box2, nullptr,
                  CU_TENSOR_MAP_SWIZZLE_128B))
      return -2;
  }
  const size_t smem = 1024 + 2 * AF_STAGE + 2 * AF_P_BYTES + (size_t)(a.num_heads / 2) * 16384 + 256 + (size_t)a.num_heads * 225 * 4 + 64;
  KB_CHECK(smem <= 232448, "kair_window_attn_fwd: shared memory plan exceeds 227 KiB (num_heads=%d)", a.num_heads);
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(window_attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_set[dev & 63] = true;
  }
  int grid = p.pairs < sm_count() ? p.pairs : sm_count();
  static long long* ts_dev = nullptr;
  const bool want_ts = KB_DEV_ENV("KB_AF_TS") != nullptr;
  if (want_ts) {
    if (!ts_dev) cudaMalloc(&ts_dev, 3 * 256 * 8);
    cudaMemset(ts_dev, 0, 3 * 256 * 8);
    p.ts = ts_dev;
  }
  KB_CUDA(launch_pdl(window_attn_fwd_kernel, grid, AF_THREADS, smem, st, p));
  if (want_ts) {
    static long long h[3 * 256];
    cudaMemcpy(h, ts_dev, sizeof(h), cudaMemcpyDeviceToHost);
    long long t0 = h[0];
    const char* names[3] = {"producer", "mma", "softmax"};
    for (int r = 0; r < 3; ++r) {
      fprintf(stderr, "%s:", names[r]);
      for (int i = 0; i < 64 && h[r * 256 + i]; ++i) fprintf(stderr, " %lld", h[r * 256 + i] - t0);
      fprintf(stderr, "\n");
    }
  }
  return 0;
}

}  // namespace kb
