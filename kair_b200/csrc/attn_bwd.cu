// kair_window_attn_bwd: backward of the fused shifted-window attention core on tcgen05 / TMEM.
//
// Per window and head (q pre-scaled, P = softmax(q k^T + bias + mask) recomputed, nothing but q,k,v was saved):
//   dP = dO V^T,  delta_i = sum_j P_ij dP_ij,  dS = P o (dP - delta)
//   dV = P^T dO,  dQ = dS K,  dK = dS^T Q,  d_table[rel(i,j), h] += dS_ij
// Every transposed product reads the SAME shared-memory tile through an MN-major UMMA descriptor (P and dS are
// written once, K-major, and consumed both as K-major and as MN-major operands), so no transposes are materialised.
//
//   warp 0      TMA producer: Q, K, V, dO tiles [2 windows x 64 tok][64 ch = 2 heads] per (window pair, head pair)
//   warp 1, 10  MMA issuers, one per group (UMMA M=64 per window; the two windows of a pair interleave in TMEM lanes);
//               a single issuing thread was the critical path: ~85 cycles per tcgen05.mma, 32 of them per group
//   warps 2..9  two groups (even / odd head of the pair), thread == query row: softmax recompute, delta, dS,
//               P/dS -> swizzled smem, d_table partial sums kept in TMEM, dQ/dK/dV -> bf16 -> TMA scatter
//
// Loop order is head-pair outermost so the per-row d_table partial sums (64 columns per head) stay resident in
// TMEM for the whole pass over the windows.
//
// Replaces autograd's backward of models/network_swinir.py:124-142 (+ :33-62, :249-252, :269-272, :216-237).
#include <string.h>

#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

constexpr int AB_THREADS = 352;   // producer, MMA issuer g0, 8 gradient warps, MMA issuer g1
constexpr int AB_STAGE = 4 * 16384;   // Q, K, V, dO tiles of one head pair
constexpr int AB_PDS = 32768;         // per group: P (16 KiB) + dS (16 KiB); reused as dq/dk/dv staging (24 KiB)

struct AttnBwdParams {
  CUtensorMap tmQKV, tmDO, tmDQKV;
  const float* table;
  float* dtable;
  int nH, H, W, nWh, nWw, shift, total_windows, pairs;
  long long* ts;  // development timeline buffer (env KB_AB_TS): [role][event] clock64 stamps of CTA 0
};

// timeline stamps: role 0 = MMA issuer, 1 = group 0 / thread 0, 2 = group 1 / thread 0; up to 256 events each
#ifdef KB_DEVEL
#define AB_STAMP(role, cnt)                                                                    \
  do {                                                                                         \
    if (p.ts && blockIdx.x == 0 && (cnt) < 256) p.ts[(role) * 256 + (cnt)++] = clock64();      \
  } while (0)
#else   // production builds carry no timeline stamps inside the pipelines
#define AB_STAMP(role, cnt) \
  do {                      \
    (void)(cnt);            \
  } while (0)
#endif

__device__ __forceinline__ void ab_window_origin(const AttnBwdParams& p, int gw, int& b, int& wi, int& wj) {
  int nW = p.nWh * p.nWw;
  b = gw / nW;
  int r = gw - b * nW;
  wi = r / p.nWw;
  wj = r - wi * p.nWw;
}

__global__ void __launch_bounds__(AB_THREADS, 1) window_attn_bwd_kernel(const __grid_constant__ AttnBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* stage_s = smem;                   // 2 x 64 KiB
  uint8_t* pds_s = stage_s + 2 * AB_STAGE;   // 2 groups x 32 KiB
  uint8_t* misc = pds_s + 2 * AB_PDS;
  uint64_t* st_full = reinterpret_cast<uint64_t*>(misc);
  uint64_t* st_empty = st_full + 2;
  uint64_t* sdp_full = st_empty + 2;
  uint64_t* pds_full = sdp_full + 2;
  uint64_t* grad_full = pds_full + 2;
  uint64_t* w_free = grad_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_free + 2);
  float* table_s = reinterpret_cast<float*>(misc + 256);   // [nH][225]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nH = p.nH;
  const int hpairs = nH / 2;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmQKV);
    tma_prefetch_desc(&p.tmDO);
    tma_prefetch_desc(&p.tmDQKV);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < 2; ++i) {
        mbar_init(&st_full[i], 1);
        mbar_init(&st_empty[i], 2);   // one commit per MMA issuer
        mbar_init(&sdp_full[i], 1);
        mbar_init(&pds_full[i], 4);
        mbar_init(&grad_full[i], 1);
        mbar_init(&w_free[i], 4);
      }
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  // barriers / TMEM slot visible before the dependency wait; after it the producer and the issuers start at once and
  // only the gradient warps (the table's sole readers) wait for its global load
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // no global access above this point
  pdl_launch_dependents();
  if (warp >= 2 && warp < 10) {
    for (int i = threadIdx.x - 64; i < 225 * nH; i += 256) {
      int idx = i / nH, h = i - idx * nH;
      table_s[h * 225 + idx] = __ldg(p.table + i);
    }
    named_bar_sync(6, 256);
  }
  // per group g: W_g = tmem_base + g*128 (S | dP, later dV | dQ | dK), ACC_g = tmem_base + 256 + g*64

  if (warp == 0) {
    // ================================ TMA producer ================================
    // The whole warp runs the loop and one ELECTED thread issues: under `if (lane == 0)` the compiler cannot prove a
    // single active thread and wraps every UTMALDG / UTCHMMA / UTMASTG (uniform-datapath operands) in an
    // ELECT / R2UR / BRA.U.ANY waterfall loop, ~100-300 cycles per instruction; under elect.sync they issue back to back.
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int hp = 0; hp < hpairs; ++hp) {
        for (int pair = blockIdx.x; pair < p.pairs; pair += gridDim.x) {
          int gw0 = pair * 2;
          int gw1 = gw0 + 1 < p.total_windows ? gw0 + 1 : gw0;
          int wb[2], wi[2], wj[2];
          ab_window_origin(p, gw0, wb[0], wi[0], wj[0]);
          ab_window_origin(p, gw1, wb[1], wi[1], wj[1]);
          mbar_wait(&st_empty[stage], phase ^ 1);
          if (elect_one()) {
            mbar_arrive_expect_tx(&st_full[stage], AB_STAGE);
            uint8_t* st = stage_s + stage * AB_STAGE;
#pragma unroll
            for (int wq = 0; wq < 8; ++wq) {
              const int wsel = wq >> 2, quad = wq & 3;
              const int hh = (wi[wsel] * 8 + (quad >> 1) * 4 + p.shift) % p.H;
              const int ww = (wj[wsel] * 8 + (quad & 1) * 4 + p.shift) % p.W;
#pragma unroll
              for (int t3 = 0; t3 < 3; ++t3)
                tma_load_4d(st + t3 * 16384 + wq * 2048, &p.tmQKV, &st_full[stage], t3 * nH * 32 + hp * 64, ww, hh,
                            wb[wsel]);
              tma_load_4d(st + 3 * 16384 + wq * 2048, &p.tmDO, &st_full[stage], hp * 64, ww, hh, wb[wsel]);
            }
          }
          __syncwarp();
          if (++stage == 2) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1 || warp == 10) {
    // ================================ MMA issuers (warp 1: group 0, warp 10: group 1) ================================
    {
      const int g = warp == 1 ? 0 : 1;
      constexpr uint32_t idesc_s = make_idesc_bf16(64, 64, 0, 0);    // S, dP
      constexpr uint32_t idesc_dv = make_idesc_bf16(64, 32, 1, 1);   // P^T dO, dS^T Q
      constexpr uint32_t idesc_dq = make_idesc_bf16(64, 32, 0, 1);   // dS K
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;
      int tsc = 0;
#define AB_MS() do { if (g == 0) AB_STAMP(0, tsc); } while (0)
      for (int hp = 0; hp < hpairs; ++hp) {
        for (int pair = blockIdx.x; pair < p.pairs; pair += gridDim.x, ++it) {
          AB_MS();
          mbar_wait(&st_full[stage], phase);
          tcgen05_fence_after();
          AB_MS();
          const uint32_t q_addr = smem_u32(stage_s + stage * AB_STAGE);
          const uint32_t k_addr = q_addr + 16384;
          const uint32_t v_addr = q_addr + 32768;
          const uint32_t do_addr = q_addr + 49152;
          {
            mbar_wait(&w_free[g], (it & 1) ^ 1);
            tcgen05_fence_after();
            const uint32_t tw = tmem_base + g * 128;
            if (elect_one()) {
#pragma unroll
            for (int w = 0; w < 2; ++w) {
              const uint32_t lo = (uint32_t)(w * 16) << 16;
#pragma unroll
              for (int k = 0; k < 2; ++k) {
                umma_f16_ss(tw + lo, make_desc_kmajor_sw128(q_addr + w * 8192 + g * 64 + k * 32),
                            make_desc_kmajor_sw128(k_addr + w * 8192 + g * 64 + k * 32), idesc_s, k);
              }
#pragma unroll
              for (int k = 0; k < 2; ++k) {
                umma_f16_ss(tw + 64 + lo, make_desc_kmajor_sw128(do_addr + w * 8192 + g * 64 + k * 32),
                            make_desc_kmajor_sw128(v_addr + w * 8192 + g * 64 + k * 32), idesc_s, k);
              }
            }
            umma_commit(&sdp_full[g]);
            }
            __syncwarp();
            AB_MS();
          }
          {
            mbar_wait(&pds_full[g], it & 1);
            tcgen05_fence_after();
            const uint32_t tw = tmem_base + g * 128;
            const uint32_t p_addr = smem_u32(pds_s + g * AB_PDS);
            const uint32_t ds_addr = p_addr + 16384;
            if (elect_one()) {
#pragma unroll
            for (int w = 0; w < 2; ++w) {
              const uint32_t lo = (uint32_t)(w * 16) << 16;
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {  // dV[j,d] = sum_i P[i,j] dO[i,d]
                umma_f16_ss(tw + lo, make_desc_mnmajor_sw128(p_addr + w * 8192 + kk * 2048, 64),
                            make_desc_mnmajor_sw128(do_addr + w * 8192 + kk * 2048 + g * 64, 64), idesc_dv, kk);
              }
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {  // dQ[i,d] = sum_j dS[i,j] K[j,d]
                umma_f16_ss(tw + 32 + lo, make_desc_kmajor_sw128(ds_addr + w * 8192 + kk * 32),
                            make_desc_mnmajor_sw128(k_addr + w * 8192 + kk * 2048 + g * 64, 64), idesc_dq, kk);
              }
#pragma unroll
              for (int kk = 0; kk < 4; ++kk) {  // dK[j,d] = sum_i dS[i,j] Q[i,d]
                umma_f16_ss(tw + 64 + lo, make_desc_mnmajor_sw128(ds_addr + w * 8192 + kk * 2048, 64),
                            make_desc_mnmajor_sw128(q_addr + w * 8192 + kk * 2048 + g * 64, 64), idesc_dv, kk);
              }
            }
            umma_commit(&grad_full[g]);
            umma_commit(&st_empty[stage]);
            }
            __syncwarp();
            AB_MS();
          }
          if (++stage == 2) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ================================ softmax-recompute / gradient warps ================================
    const int q = warp & 3;
    const int g = (warp - 2) >> 2;
    const int gtid = ((warp - 2) & 3) * 32 + lane;  // thread index inside the group
    const int wsel = lane >> 4;
    const int i = q * 16 + (lane & 15);
    const int row = wsel * 64 + i;
    const int quad_i = q;
    const int ry = (quad_i >> 1) * 4 + ((i & 15) >> 2);
    const int rx = (quad_i & 1) * 4 + (i & 3);
    const int base_i = ry * 15 + rx;
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    const uint32_t tW = tmem_base + lane_sel + g * 128;
    const uint32_t tACC = tmem_base + lane_sel + 256 + g * 64;
    uint8_t* pds = pds_s + g * AB_PDS;
    const float LOG2E = 1.4426950408889634f;
    uint32_t it = 0;
    int tsc = 0;
#define AB_CS() do { if (gtid == 0) AB_STAMP(1 + g, tsc); } while (0)

    for (int hp = 0; hp < hpairs; ++hp) {
      const int h = hp * 2 + g;
      const float* tb = table_s + h * 225 + base_i;
      {  // zero the d_table partial sums of this pass
        uint32_t z[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) z[j] = 0u;
        tmem_st32(tACC, z);
        tmem_st32(tACC + 32, z);
        tmem_st_wait();
      }
      for (int pair = blockIdx.x; pair < p.pairs; pair += gridDim.x, ++it) {
        const int gw0 = pair * 2;
        const bool second_valid = gw0 + 1 < p.total_windows;
        const bool my_valid = (wsel == 0) || second_valid;
        const int gw = (wsel && second_valid) ? gw0 + 1 : gw0;
        int wb, wi, wj;
        ab_window_origin(p, gw, wb, wi, wj);
        // output box coordinates of this warp's quadrant in both windows, for the lane that issues the TMA stores
        // (computed here, off the critical path between grad_full and the stores; lane 16 holds the second window)
        int sy = wi * 8 + (q >> 1) * 4 + p.shift, sx = wj * 8 + (q & 1) * 4 + p.shift;
        if (sy >= p.H) sy -= p.H;
        if (sx >= p.W) sx -= p.W;
        const int sb1 = __shfl_sync(0xffffffffu, wb, 16), sy1 = __shfl_sync(0xffffffffu, sy, 16),
                  sx1 = __shfl_sync(0xffffffffu, sx, 16);
        const bool lastrow = p.shift > 0 && wi == p.nWh - 1;
        const bool lastcol = p.shift > 0 && wj == p.nWw - 1;
        float madd[4];
#pragma unroll
        for (int qj = 0; qj < 4; ++qj) {
          bool m = (lastrow && ((qj >> 1) != (quad_i >> 1))) || (lastcol && ((qj & 1) != (quad_i & 1)));
          madd[qj] = m ? -100.0f : 0.0f;
        }

        // ---- A/B: S -> normalised P (fp32 registers)
        AB_CS();
        mbar_wait(&sdp_full[g], it & 1);
        tcgen05_fence_after();
        AB_CS();
        float pr[64];
        {
          uint32_t r0[32], r1[32];
          tmem_ld32(tW, r0);
          tmem_ld32(tW + 32, r1);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            pr[j] = __uint_as_float(r0[j]);
            pr[32 + j] = __uint_as_float(r1[j]);
          }
        }
        // bias, mask (only the windows on the shifted border carry one), softmax.  All fp32 arithmetic of this phase is
        // packed two lanes per instruction (FADD2 / FFMA2 / FMUL2): the phase was issue bound (two gradient warps per
        // scheduler, ~15 issue slots per element).  Four independent max / sum / delta chains.
#pragma unroll
        for (int j = 0; j < 64; j += 2) {
          const int qj = j >> 4;
          const int jy = (qj >> 1) * 4 + ((j & 15) >> 2);
          const int jx = (qj & 1) * 4 + (j & 3);
          const int o = (7 - jy) * 15 + (7 - jx);   // column j + 1 sits one step to the right: table offset o - 1
          add2(pr[j], pr[j + 1], pr[j], pr[j + 1], tb[o], tb[o - 1]);
        }
        if (lastrow || lastcol) {
#pragma unroll
          for (int j = 0; j < 64; j += 2) add2(pr[j], pr[j + 1], pr[j], pr[j + 1], madd[j >> 4], madd[j >> 4]);
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < 64; ++j) mx4[j & 3] = fmaxf(mx4[j & 3], pr[j]);
        const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        float sum4[4] = {0.f, 0.f, 0.f, 0.f};
        const float nmb = -mx * LOG2E;
#pragma unroll
        for (int j = 0; j < 64; j += 4) {
          fma2(pr[j], pr[j + 1], pr[j], pr[j + 1], LOG2E, LOG2E, nmb, nmb);
          fma2(pr[j + 2], pr[j + 3], pr[j + 2], pr[j + 3], LOG2E, LOG2E, nmb, nmb);
#pragma unroll
          for (int u = 0; u < 4; ++u) pr[j + u] = ex2_approx(pr[j + u]);
          add2(sum4[0], sum4[1], sum4[0], sum4[1], pr[j], pr[j + 1]);
          add2(sum4[2], sum4[3], sum4[2], sum4[3], pr[j + 2], pr[j + 3]);
        }
        const float sum = (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
        // rows of the padding window of an odd window count get P = 0, hence dS = 0 and no d_table contribution
        const float inv = my_valid ? 1.0f / sum : 0.0f;
#pragma unroll
        for (int j = 0; j < 64; j += 2) mul2(pr[j], pr[j + 1], pr[j], pr[j + 1], inv, inv);

        // ---- C: delta = sum_j P_ij dP_ij
        AB_CS();
        float dl4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t r[32];
          tmem_ld32(tW + 64 + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            fma2(dl4[0], dl4[1], pr[c * 32 + j], pr[c * 32 + j + 1], __uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                 dl4[0], dl4[1]);
            fma2(dl4[2], dl4[3], pr[c * 32 + j + 2], pr[c * 32 + j + 3], __uint_as_float(r[j + 2]),
                 __uint_as_float(r[j + 3]), dl4[2], dl4[3]);
          }
        }
        const float ndelta = -((dl4[0] + dl4[1]) + (dl4[2] + dl4[3]));

        // ---- D: the P/dS tiles double as the previous iteration's output staging: wait for its TMA reads
        AB_CS();
        if (elect_one()) tma_store_wait_read<0>();   // every warp issued the stores of its own quadrant rows
        named_bar_sync(2 + g, 128);
        AB_CS();

        // ---- E: dS, d_table partial sums, P/dS -> smem
        const uint32_t prow = smem_u32(pds) + row * 128;
        const uint32_t dsrow = prow + 16384;
        const uint32_t sw = row & 7;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t r[32], a[32];
          tmem_ld32(tW + 64 + c * 32, r);
          tmem_ld32(tACC + c * 32, a);
          tmem_ld_wait();
          float ds[32];
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            float t0, t1, a0, a1;
            add2(t0, t1, __uint_as_float(r[j]), __uint_as_float(r[j + 1]), ndelta, ndelta);
            mul2(ds[j], ds[j + 1], pr[c * 32 + j], pr[c * 32 + j + 1], t0, t1);
            add2(a0, a1, __uint_as_float(a[j]), __uint_as_float(a[j + 1]), ds[j], ds[j + 1]);
            a[j] = __float_as_uint(a0);
            a[j + 1] = __float_as_uint(a1);
          }
          tmem_st32(tACC + c * 32, a);
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) {
            const uint32_t off = (((c * 4 + cc) ^ sw) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dsrow + off),
                         "r"(pack_bf16x2(ds[8 * cc], ds[8 * cc + 1])), "r"(pack_bf16x2(ds[8 * cc + 2], ds[8 * cc + 3])),
                         "r"(pack_bf16x2(ds[8 * cc + 4], ds[8 * cc + 5])), "r"(pack_bf16x2(ds[8 * cc + 6], ds[8 * cc + 7]))
                         : "memory");
            const int o = c * 32 + cc * 8;
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(prow + off),
                         "r"(pack_bf16x2(pr[o], pr[o + 1])), "r"(pack_bf16x2(pr[o + 2], pr[o + 3])),
                         "r"(pack_bf16x2(pr[o + 4], pr[o + 5])), "r"(pack_bf16x2(pr[o + 6], pr[o + 7]))
                         : "memory");
          }
        }
        // ---- F
        tmem_st_wait();
        fence_proxy_async_smem();
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&pds_full[g]);
        AB_CS();

        // ---- G: dV | dQ | dK -> bf16 -> staging (64-byte rows, SWIZZLE_64B), tensor order dq, dk, dv
        mbar_wait(&grad_full[g], it & 1);
        tcgen05_fence_after();
        AB_CS();
        {
          const uint32_t sw64 = (row >> 1) & 3;
#pragma unroll
          for (int t3 = 0; t3 < 3; ++t3) {
            // TMEM column blocks: 0 = dV, 32 = dQ, 64 = dK ; staging tensors: 0 = dq, 1 = dk, 2 = dv
            const int tcol = t3 == 0 ? 32 : (t3 == 1 ? 64 : 0);
            uint32_t r[32];
            tmem_ld32(tW + tcol, r);
            tmem_ld_wait();
            const uint32_t base = smem_u32(pds) + t3 * 8192 + row * 64;
#pragma unroll
            for (int cc = 0; cc < 4; ++cc) {
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(base + ((cc ^ sw64) << 4)),
                           "r"(pack_bf16x2(__uint_as_float(r[8 * cc]), __uint_as_float(r[8 * cc + 1]))),
                           "r"(pack_bf16x2(__uint_as_float(r[8 * cc + 2]), __uint_as_float(r[8 * cc + 3]))),
                           "r"(pack_bf16x2(__uint_as_float(r[8 * cc + 4]), __uint_as_float(r[8 * cc + 5]))),
                           "r"(pack_bf16x2(__uint_as_float(r[8 * cc + 6]), __uint_as_float(r[8 * cc + 7])))
                           : "memory");
            }
          }
        }
        // ---- H
        fence_proxy_async_smem();
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&w_free[g]);
        AB_CS();
        // the staging rows of quadrant q of both windows were written by this warp alone: its elected thread stores them
        if (elect_one()) {
#pragma unroll
          for (int t3 = 0; t3 < 3; ++t3)
            tma_store_4d(&p.tmDQKV, pds + t3 * 8192 + q * 1024, t3 * nH * 32 + h * 32, sx, sy, wb);
          if (second_valid) {
#pragma unroll
            for (int t3 = 0; t3 < 3; ++t3)
              tma_store_4d(&p.tmDQKV, pds + t3 * 8192 + (4 + q) * 1024, t3 * nH * 32 + h * 32, sx1, sy1, sb1);
          }
          tma_store_commit();
        }
        __syncwarp();
        AB_CS();
      }

      // ---- end of pass: fold the per-row partial sums into d_table[:, h].  The two windows of the pair are summed by a
      // shuffle, the 64x64 (i, j) sums go to shared memory (the P/dS tiles are idle now) and each table entry gathers
      // its (8-|dy|)(8-|dx|) contributions with plain loads: one global atomic per entry and pass.  (Shared-memory
      // atomics per (i, j) cost ~21 000 cycles per pass: fp32 atomicAdd on shared memory is a CAS loop.)
      {
        if (elect_one()) tma_store_wait_read<0>();
        named_bar_sync(2 + g, 128);
        float* M = reinterpret_cast<float*>(pds);  // [64][65]
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t a[32];
          tmem_ld32(tACC + c * 32, a);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float v = __uint_as_float(a[j]);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            if (lane < 16) M[i * 65 + c * 32 + j] = v;
          }
        }
        named_bar_sync(2 + g, 128);
        for (int idx = gtid; idx < 225; idx += 128) {
          const int dy = idx / 15 - 7, dx = idx % 15 - 7;
          // fixed 8x8 trip counts with predicated loads: all 64 loads are independent (a variable-bound loop made
          // this gather a ~7000-cycle dependent chain)
          int fx[8];
          bool vx[8];
#pragma unroll
          for (int xi = 0; xi < 8; ++xi) {
            const int xj = xi - dx;
            vx[xi] = (unsigned)xj < 8u;
            fx[xi] = (xj >> 2) * 16 + (xj & 3);
          }
          float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int yi = 0; yi < 8; ++yi) {
            const int yj = yi - dy;
            const bool vy = (unsigned)yj < 8u;
            const int fy = (yj >> 2) * 32 + (yj & 3) * 4;
#pragma unroll
            for (int xi = 0; xi < 8; ++xi) {
              const int ii = ((yi >> 2) * 2 + (xi >> 2)) * 16 + (yi & 3) * 4 + (xi & 3);
              s4[xi & 3] += (vy && vx[xi]) ? M[ii * 65 + fy + fx[xi]] : 0.f;
            }
          }
          const float s = (s4[0] + s4[1]) + (s4[2] + s4[3]);
          if (s != 0.f) atomicAdd(&p.dtable[idx * nH + h], s);
        }
        named_bar_sync(2 + g, 128);
      }
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

int window_attn_bwd_launch(const kair_window_attn_args& a, cudaStream_t st) {
  KB_CHECK(a.qkv && a.dout && a.dqkv && a.table && a.dtable, "kair_window_attn_bwd: null pointer");
  KB_CHECK(a.window == 8, "kair_window_attn_bwd: only 8x8 windows are supported by the tcgen05 path (got %d)", a.window);
  KB_CHECK(a.num_heads >= 2 && a.num_heads <= 8 && a.num_heads % 2 == 0,
           "kair_window_attn_bwd: num_heads=%d must be even and <= 8", a.num_heads);
  KB_CHECK(a.H % 8 == 0 && a.W % 8 == 0 && a.H > 0 && a.W > 0, "kair_window_attn_bwd: H,W must be multiples of 8");
  KB_CHECK(a.shift == 0 || a.shift == 4, "kair_window_attn_bwd: shift must be 0 or 4");
  KB_CHECK(a.qkv_pitch % 8 == 0 && a.dout_pitch % 8 == 0 && a.dqkv_pitch % 8 == 0,
           "kair_window_attn_bwd: pitches must be multiples of 8");
  if (check_device() != 0) return -3;

  AttnBwdParams p;
  memset(&p, 0, sizeof(p));
  p.table = a.table;
  p.dtable = a.dtable;
  p.nH = a.num_heads;
  p.H = a.H;
  p.W = a.W;
  p.nWh = a.H / 8;
  p.nWw = a.W / 8;
  p.shift = a.shift;
  p.total_windows = a.batch * p.nWh * p.nWw;
  p.pairs = (p.total_windows + 1) / 2;
  // L2 promotion 128 B: the channel (head pair) loop is the OUTER loop, so the 128-byte neighbour a 256-byte promotion
  // would drag in from DRAM belongs to a pass that runs tens of microseconds later (ncu: 300 MB read for 201 MB asked)
  auto map4 = [&](CUtensorMap* tm, const void* ptr, long long pitch, int cols, uint32_t boxc, CUtensorMapSwizzle sw) {
    uint64_t dims[4] = {(uint64_t)cols, (uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.batch};
    uint64_t str[3] = {(uint64_t)pitch * 2, (uint64_t)pitch * 2 * a.W, (uint64_t)pitch * 2 * a.W * a.H};
    uint32_t box[4] = {boxc, 4, 4, 1};
    return make_tmap(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, ptr, dims, str, box, nullptr, sw, 128);
  };
  if (map4(&p.tmQKV, a.qkv, a.qkv_pitch, 3 * a.num_heads * 32, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return -2;
  if (map4(&p.tmDO, a.dout, a.dout_pitch, a.num_heads * 32, 64, CU_TENSOR_MAP_SWIZZLE_128B)) return -2;
  if (map4(&p.tmDQKV, a.dqkv, a.dqkv_pitch, 3 * a.num_heads * 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) return -2;

  const size_t smem = 1024 + 2 * AB_STAGE + 2 * AB_PDS + 256 + (size_t)a.num_heads * 225 * 4 + 2 * 225 * 4 + 64;
  KB_CHECK(smem <= 232448, "kair_window_attn_bwd: shared memory plan exceeds 227 KiB");
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(window_attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
    attr_set[dev & 63] = true;
  }
  int grid = p.pairs < sm_count() ? p.pairs : sm_count();
  static long long* ts_dev = nullptr;
  const bool want_ts = KB_DEV_ENV("KB_AB_TS") != nullptr;
  if (want_ts) {
    if (!ts_dev) cudaMalloc(&ts_dev, 3 * 256 * 8);
    cudaMemset(ts_dev, 0, 3 * 256 * 8);
    p.ts = ts_dev;
  }
  KB_CUDA(launch_pdl(window_attn_bwd_kernel, grid, AB_THREADS, smem, st, p));
  if (want_ts) {
    static long long h[3 * 256];
    cudaMemcpy(h, ts_dev, sizeof(h), cudaMemcpyDeviceToHost);
    const long long t0 = h[0];
    const char* names[3] = {"mma", "group0", "group1"};
    for (int r = 0; r < 3; ++r) {
      fprintf(stderr, "%s:", names[r]);
      for (int i = 0; i < 72 && h[r * 256 + i]; ++i) fprintf(stderr, " %lld", h[r * 256 + i] - t0);
      fprintf(stderr, "\n");
    }
  }
  return 0;
}

}  // namespace kb
