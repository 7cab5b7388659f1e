// kair_window_attn_fwd / _bwd for window sizes other than 8 (2..7): the JPEG-artifact SwinIR models use 7x7 windows
// (main_test_swinir.py:170-175 of the reference; options/swinir/train_swinir_car_jpeg.json), whose 49-token windows do
// not fill the 64-row UMMA tiles the tcgen05 kernels (attn_fwd.cu / attn_bwd.cu) are laid out for.  The attention
// matmuls are ~7 % of the network's FLOPs, so this path runs them on the CUDA cores in fp32:
//   one CTA per (window, head), one thread per query token; K / V rows are broadcast from shared memory.
// Same contract as the 8x8 kernels (include/kair_b200.h): q pre-scaled, head_dim padded to 32 columns, window
// partition / cyclic shift / window reverse as index arithmetic, closed-form {0,-100} mask
// (network_swinir.py:114-145, 216-237, 249-272), d(table) accumulated with atomics.
#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

namespace {

constexpr int AS_MAXN = 64;   // tokens per window (window <= 8)
constexpr int AS_HD = 32;     // padded head dim
constexpr int AS_PS = 65;     // row pitch of the transposed P / dS matrices (odd: conflict-free column reads)

struct SmallGeom {
  int ws, N, H, W, nWh, nWw, nH, shift, tsz;
};

__device__ __forceinline__ int region(int p, int L, int ws, int s) { return (p >= L - ws ? 1 : 0) + (p >= L - s ? 1 : 0); }

// token index of in-window position i of window (b, wi, wj) and its mask region id (in the shifted frame)
__device__ __forceinline__ long long token_of(const SmallGeom& g, int b, int wi, int wj, int i, int& rid) {
  const int r = i / g.ws, c = i % g.ws;
  const int ph = wi * g.ws + r, pw = wj * g.ws + c;   // position in the rolled image
  rid = g.shift ? 3 * region(ph, g.H, g.ws, g.shift) + region(pw, g.W, g.ws, g.shift) : 0;
  int y = ph + g.shift, x = pw + g.shift;
  if (y >= g.H) y -= g.H;
  if (x >= g.W) x -= g.W;
  return ((long long)b * g.H + y) * g.W + x;
}

__device__ __forceinline__ void load_row32(const bf16* p, float (&f)[AS_HD]) {
  const uint4* v = reinterpret_cast<const uint4*>(p);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint4 u = __ldg(v + k);
    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 t = __bfloat1622float2(h2[e]);
      f[8 * k + 2 * e] = t.x;
      f[8 * k + 2 * e + 1] = t.y;
    }
  }
}

__device__ __forceinline__ void store_row32(bf16* p, const float (&f)[AS_HD]) {
  uint4* v = reinterpret_cast<uint4*>(p);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    uint4 u;
    u.x = pack_bf16x2(f[8 * k], f[8 * k + 1]);
    u.y = pack_bf16x2(f[8 * k + 2], f[8 * k + 3]);
    u.z = pack_bf16x2(f[8 * k + 4], f[8 * k + 5]);
    u.w = pack_bf16x2(f[8 * k + 6], f[8 * k + 7]);
    v[k] = u;
  }
}

__device__ __forceinline__ void load_row32(const float* p, float (&f)[AS_HD]) {
  const float4* v = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float4 u = __ldg(v + k);
    f[4 * k] = u.x; f[4 * k + 1] = u.y; f[4 * k + 2] = u.z; f[4 * k + 3] = u.w;
  }
}

__device__ __forceinline__ void store_row32(float* p, const float (&f)[AS_HD]) {
  float4* v = reinterpret_cast<float4*>(p);
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = make_float4(f[4 * k], f[4 * k + 1], f[4 * k + 2], f[4 * k + 3]);
}

// fp32-accurate mode uses expf (<= 2 ulp); the bf16 path keeps the MUFU approximation
template <typename IO> __device__ __forceinline__ float exp_io(float x);
template <> __device__ __forceinline__ float exp_io<bf16>(float x) { return __expf(x); }
template <> __device__ __forceinline__ float exp_io<float>(float x) { return expf(x); }

__device__ __forceinline__ float dot32(const float (&a)[AS_HD], const float* b) {
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
  for (int d = 0; d < AS_HD; d += 4) {
    s0 = fmaf(a[d], b[d], s0);
    s1 = fmaf(a[d + 1], b[d + 1], s1);
    s2 = fmaf(a[d + 2], b[d + 2], s2);
    s3 = fmaf(a[d + 3], b[d + 3], s3);
  }
  return (s0 + s1) + (s2 + s3);
}

__device__ __forceinline__ int rel_index(int i, int j, int ws) {
  return (i / ws - j / ws + ws - 1) * (2 * ws - 1) + (i % ws - j % ws + ws - 1);
}

// ------------------------------------------------------------------------------------------------ forward
template <typename IO>   // bf16 (window 2..7 of the bf16 mode) or float (every window of the fp32-accurate mode)
__global__ void __launch_bounds__(AS_MAXN) attn_small_fwd_kernel(const kair_window_attn_args a, const SmallGeom g) {
  __shared__ float Ks[AS_MAXN][AS_HD];
  __shared__ float Vs[AS_MAXN][AS_HD];
  __shared__ float tab[225];
  __shared__ int rids[AS_MAXN];
  const int head = blockIdx.x % g.nH;
  const int win = blockIdx.x / g.nH;
  const int wj = win % g.nWw, wi = (win / g.nWw) % g.nWh, b = win / (g.nWw * g.nWh);
  const int i = threadIdx.x;
  const bool live = i < g.N;
  const IO* qkv = reinterpret_cast<const IO*>(a.qkv);
  const int HP = g.nH * AS_HD;

  for (int e = i; e < g.tsz; e += AS_MAXN) tab[e] = __ldg(a.table + (long long)e * g.nH + head);
  float q[AS_HD];
  int rid = 0;
  long long tok = 0;
  if (live) {
    tok = token_of(g, b, wi, wj, i, rid);
    const IO* row = qkv + tok * a.qkv_pitch + head * AS_HD;
    float t[AS_HD];
    load_row32(row, q);
    load_row32(row + HP, t);
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) Ks[i][d] = t[d];
    load_row32(row + 2 * HP, t);
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) Vs[i][d] = t[d];
    rids[i] = rid;
  }
  __syncthreads();
  if (!live) return;

  // online softmax over the window's keys
  float m = -INFINITY, l = 0.f;
  float o[AS_HD];
#pragma unroll
  for (int d = 0; d < AS_HD; ++d) o[d] = 0.f;
  for (int j = 0; j < g.N; ++j) {
    float s = dot32(q, Ks[j]) + tab[rel_index(i, j, g.ws)];
    if (rids[j] != rid) s -= 100.0f;
    const float mn = fmaxf(m, s);
    const float corr = exp_io<IO>(m - mn), p = exp_io<IO>(s - mn);
    l = l * corr + p;
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) o[d] = fmaf(o[d], corr, p * Vs[j][d]);
    m = mn;
  }
  const float inv = 1.0f / l;
#pragma unroll
  for (int d = 0; d < AS_HD; ++d) o[d] *= inv;
  store_row32(reinterpret_cast<IO*>(a.out) + tok * a.out_pitch + head * AS_HD, o);
}

// ------------------------------------------------------------------------------------------------ backward
// phase 1 (thread = query i): P_ij, dP_ij = dO_i . V_j, delta_i, dS_ij = P_ij (dP_ij - delta_i), dQ_i = sum_j dS_ij K_j;
//                             P and dS are stored transposed ([j][i]) and d(table) is accumulated in shared memory
// phase 2 (thread = key j):   dK_j = sum_i dS_ij Q_i,  dV_j = sum_i P_ij dO_i
template <typename IO>
__global__ void __launch_bounds__(AS_MAXN) attn_small_bwd_kernel(const kair_window_attn_args a, const SmallGeom g) {
  extern __shared__ float sm[];
  float* Qs = sm;                          // [64][32]
  float* Ks = Qs + AS_MAXN * AS_HD;
  float* Vs = Ks + AS_MAXN * AS_HD;
  float* Gs = Vs + AS_MAXN * AS_HD;        // dO
  float* Pt = Gs + AS_MAXN * AS_HD;        // [64][65]
  float* St = Pt + AS_MAXN * AS_PS;
  float* tab = St + AS_MAXN * AS_PS;       // [225]
  float* dtab = tab + 225;                 // [225]
  int* rids = reinterpret_cast<int*>(dtab + 225);
  const int head = blockIdx.x % g.nH;
  const int win = blockIdx.x / g.nH;
  const int wj = win % g.nWw, wi = (win / g.nWw) % g.nWh, b = win / (g.nWw * g.nWh);
  const int i = threadIdx.x;
  const bool live = i < g.N;
  const IO* qkv = reinterpret_cast<const IO*>(a.qkv);
  const int HP = g.nH * AS_HD;

  for (int e = i; e < g.tsz; e += AS_MAXN) {
    tab[e] = __ldg(a.table + (long long)e * g.nH + head);
    dtab[e] = 0.f;
  }
  float q[AS_HD], go[AS_HD];
  int rid = 0;
  long long tok = 0;
  if (live) {
    tok = token_of(g, b, wi, wj, i, rid);
    const IO* row = qkv + tok * a.qkv_pitch + head * AS_HD;
    float t[AS_HD];
    load_row32(row, q);
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) Qs[i * AS_HD + d] = q[d];
    load_row32(row + HP, t);
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) Ks[i * AS_HD + d] = t[d];
    load_row32(row + 2 * HP, t);
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) Vs[i * AS_HD + d] = t[d];
    load_row32(reinterpret_cast<const IO*>(a.dout) + tok * a.dout_pitch + head * AS_HD, go);
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) Gs[i * AS_HD + d] = go[d];
    rids[i] = rid;
  }
  __syncthreads();

  if (live) {
    float m = -INFINITY;
    for (int j = 0; j < g.N; ++j) {
      float s = dot32(q, Ks + j * AS_HD) + tab[rel_index(i, j, g.ws)];
      if (rids[j] != rid) s -= 100.0f;
      Pt[j * AS_PS + i] = s;
      m = fmaxf(m, s);
    }
    float l = 0.f;
    for (int j = 0; j < g.N; ++j) {
      const float p = exp_io<IO>(Pt[j * AS_PS + i] - m);
      Pt[j * AS_PS + i] = p;
      l += p;
    }
    const float inv = 1.0f / l;
    float delta = 0.f;
    for (int j = 0; j < g.N; ++j) {
      const float p = Pt[j * AS_PS + i] * inv;
      const float dp = dot32(go, Vs + j * AS_HD);
      Pt[j * AS_PS + i] = p;
      St[j * AS_PS + i] = dp;
      delta = fmaf(p, dp, delta);
    }
    float dq[AS_HD];
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) dq[d] = 0.f;
    for (int j = 0; j < g.N; ++j) {
      const float ds = Pt[j * AS_PS + i] * (St[j * AS_PS + i] - delta);
      St[j * AS_PS + i] = ds;
      atomicAdd(&dtab[rel_index(i, j, g.ws)], ds);
      const float* kj = Ks + j * AS_HD;
#pragma unroll
      for (int d = 0; d < AS_HD; ++d) dq[d] = fmaf(ds, kj[d], dq[d]);
    }
    store_row32(reinterpret_cast<IO*>(a.dqkv) + tok * a.dqkv_pitch + head * AS_HD, dq);
  }
  __syncthreads();

  if (live) {
    const int j = i;
    float dk[AS_HD], dv[AS_HD];
#pragma unroll
    for (int d = 0; d < AS_HD; ++d) { dk[d] = 0.f; dv[d] = 0.f; }
    for (int r = 0; r < g.N; ++r) {
      const float ds = St[j * AS_PS + r], p = Pt[j * AS_PS + r];
      const float* qr = Qs + r * AS_HD;
      const float* gr = Gs + r * AS_HD;
#pragma unroll
      for (int d = 0; d < AS_HD; ++d) {
        dk[d] = fmaf(ds, qr[d], dk[d]);
        dv[d] = fmaf(p, gr[d], dv[d]);
      }
    }
    IO* drow = reinterpret_cast<IO*>(a.dqkv) + tok * a.dqkv_pitch + head * AS_HD;
    store_row32(drow + HP, dk);
    store_row32(drow + 2 * HP, dv);
  }
  for (int e = i; e < g.tsz; e += AS_MAXN) atomicAdd(a.dtable + (long long)e * g.nH + head, dtab[e]);
}

int small_geom(const kair_window_attn_args& a, SmallGeom& g, const char* who) {
  KB_CHECK(a.window >= 2 && a.window <= 8, "%s: window=%d must be in 2..8", who, a.window);
  KB_CHECK(a.num_heads >= 1 && a.num_heads <= 16, "%s: num_heads=%d", who, a.num_heads);
  KB_CHECK(a.H > 0 && a.W > 0 && a.batch > 0 && a.H % a.window == 0 && a.W % a.window == 0,
           "%s: H,W must be positive multiples of the window (%d)", who, a.window);
  KB_CHECK(a.shift >= 0 && a.shift < a.window, "%s: shift=%d out of range", who, a.shift);
  KB_CHECK(a.qkv_pitch % 8 == 0 && a.qkv_pitch >= 3 * a.num_heads * AS_HD, "%s: bad qkv pitch", who);
  g.ws = a.window; g.N = a.window * a.window; g.H = a.H; g.W = a.W;
  g.nWh = a.H / a.window; g.nWw = a.W / a.window; g.nH = a.num_heads; g.shift = a.shift;
  g.tsz = (2 * a.window - 1) * (2 * a.window - 1);
  return 0;
}

}  // namespace

int window_attn_small_fwd_launch(const kair_window_attn_args& a, cudaStream_t st) {
  KB_CHECK(a.qkv && a.out && a.table, "kair_window_attn_fwd: null pointer");
  SmallGeom g;
  if (int rc = small_geom(a, g, "kair_window_attn_fwd")) return rc;
  KB_CHECK(a.out_pitch % 8 == 0 && a.out_pitch >= a.num_heads * AS_HD, "kair_window_attn_fwd: bad out pitch");
  const long long blocks = (long long)a.batch * g.nWh * g.nWw * g.nH;
  KB_CHECK(blocks < (1ll << 31), "kair_window_attn_fwd: too many windows");
  if (a.fp32) attn_small_fwd_kernel<float><<<(unsigned)blocks, AS_MAXN, 0, st>>>(a, g);
  else attn_small_fwd_kernel<bf16><<<(unsigned)blocks, AS_MAXN, 0, st>>>(a, g);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int window_attn_small_bwd_launch(const kair_window_attn_args& a, cudaStream_t st) {
  KB_CHECK(a.qkv && a.dout && a.dqkv && a.dtable && a.table, "kair_window_attn_bwd: null pointer");
  SmallGeom g;
  if (int rc = small_geom(a, g, "kair_window_attn_bwd")) return rc;
  KB_CHECK(a.dout_pitch % 8 == 0 && a.dout_pitch >= a.num_heads * AS_HD && a.dqkv_pitch % 8 == 0 &&
               a.dqkv_pitch >= 3 * a.num_heads * AS_HD, "kair_window_attn_bwd: bad gradient pitch");
  const long long blocks = (long long)a.batch * g.nWh * g.nWw * g.nH;
  KB_CHECK(blocks < (1ll << 31), "kair_window_attn_bwd: too many windows");
  const size_t smem = (4 * AS_MAXN * AS_HD + 2 * AS_MAXN * AS_PS + 2 * 225) * sizeof(float) + AS_MAXN * sizeof(int);
  static bool attr_set[64] = {false};
  int dev = 0;
  KB_CUDA(cudaGetDevice(&dev));
  if (!attr_set[dev & 63]) {
    KB_CUDA(cudaFuncSetAttribute(attn_small_bwd_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    KB_CUDA(cudaFuncSetAttribute(attn_small_bwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set[dev & 63] = true;
  }
  if (a.fp32) attn_small_bwd_kernel<float><<<(unsigned)blocks, AS_MAXN, smem, st>>>(a, g);
  else attn_small_bwd_kernel<bf16><<<(unsigned)blocks, AS_MAXN, smem, st>>>(a, g);
  KB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kb
