// Tile / ensemble drivers on the device: the data movement of the reference's large-image
// inference loops as index arithmetic, one launch per chunk of tiles instead of one eager slice / cat / add_ / rot90
// per tile.
//   kair_tiles_gather      patches[n] = D_mode(img[:, y:y+t, x:x+t])  -- the tile cut of main_test_swinir.py:266-276
//                          (`test`) and, with mode = 0..7, the eight flip / rotation inputs of utils_model.test_x8
//                          (utils/utils_model.py:176-178, utils_image.augment_img_tensor4 :408-426) in one batch
//   kair_tiles_accumulate  E[:, sy:sy+ts, sx:sx+ts] += w * D_mode^-1(tiles[n]),  W[...] += w  -- the overlap-add of
//                          main_test_swinir.py:277-281 and the inverse transforms + mean of test_x8 (:179-185).
//                          One thread per OUTPUT pixel walks the chunk's tiles in index order, so the floating-point
//                          sum order is fixed (no atomics): results are deterministic and equal the sequential loop.
// Both are pure copies / adds: HBM-bound, 4 B in + 4 B out per element.
#include "common.cuh"
#include "host.h"
#include "../../include/kair_b200.h"

namespace kb {

namespace {

// augment_img_tensor4(img, mode)[i][j] = img[r][c] for a source of size (hs, ws); output size is (hs, ws) for the
// modes without a quarter turn and (ws, hs) otherwise
__device__ __forceinline__ void dihedral_src(int m, int i, int j, int hs, int ws, int& r, int& c) {
  switch (m) {
    case 0: r = i; c = j; break;
    case 1: r = j; c = i; break;                       // rot90(1).flip(rows)  = transpose
    case 2: r = hs - 1 - i; c = j; break;              // flip(rows)
    case 3: r = hs - 1 - j; c = i; break;              // rot90(3)
    case 4: r = i; c = ws - 1 - j; break;              // rot90(2).flip(rows)  = flip(cols)
    case 5: r = j; c = ws - 1 - i; break;              // rot90(1)
    case 6: r = hs - 1 - i; c = ws - 1 - j; break;     // rot90(2)
    default: r = hs - 1 - j; c = ws - 1 - i; break;    // rot90(3).flip(rows)  = anti-transpose
  }
}
__device__ __forceinline__ bool quarter_turn(int m) { return m == 1 || m == 3 || m == 5 || m == 7; }

__global__ void tiles_gather_kernel(const kair_tiles_args a) {
  const long long per = (long long)a.C * a.th * a.tw;     // th x tw: size of one gathered patch (after the transform)
  const long long total = (long long)a.n * per;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(e / per);
    const int rem = (int)(e - (long long)n * per);
    const int ch = rem / (a.th * a.tw);
    const int ij = rem - ch * a.th * a.tw;
    const int i = ij / a.tw, j = ij - i * a.tw;
    const int* o = a.origins + 3 * n;
    const int mode = o[2];
    const int hs = quarter_turn(mode) ? a.tw : a.th, ws = quarter_turn(mode) ? a.th : a.tw;   // source window size
    int r, c;
    dihedral_src(mode, i, j, hs, ws, r, c);
    a.tiles[e] = __ldg(a.img + ((size_t)ch * a.H + o[0] + r) * a.W + o[1] + c);
  }
}

// inverse of mode m as a mode (utils_model.test_x8: 3 <-> 5, the rest are involutions)
__device__ __forceinline__ int inverse_mode(int m) { return m == 3 ? 5 : (m == 5 ? 3 : m); }

__global__ void tiles_accumulate_kernel(const kair_tiles_args a) {
  const long long total = (long long)a.C * a.H * a.W;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int ch = (int)(e / ((long long)a.H * a.W));
    const int yx = (int)(e - (long long)ch * a.H * a.W);
    const int y = yx / a.W, x = yx - y * a.W;
    float acc = 0.f, wacc = 0.f;
    bool hit = false;
    for (int n = 0; n < a.n; ++n) {
      const int* o = a.origins + 3 * n;
      const int mode = o[2];
      // the tile as stored is (th x tw); mapped back it covers (hb x wb) at (o[0], o[1])
      const int hb = quarter_turn(mode) ? a.tw : a.th, wb = quarter_turn(mode) ? a.th : a.tw;
      const int i = y - o[0], j = x - o[1];
      if (i < 0 || i >= hb || j < 0 || j >= wb) continue;
      int r, c;
      dihedral_src(inverse_mode(mode), i, j, a.th, a.tw, r, c);
      acc += a.weight * a.tiles[(((size_t)n * a.C + ch) * a.th + r) * a.tw + c];
      wacc += a.weight;
      hit = true;
    }
    if (hit) {
      a.E[e] += acc;
      if (a.Wsum && ch == 0) a.Wsum[yx] += wacc;
    }
  }
}

}  // namespace

namespace {
__global__ void ffd_pack_kernel(const kair_ffd_pack_args a) {
  const int H2 = (a.h + 1) >> 1, W2 = (a.w + 1) >> 1;
  const int groups = (int)(a.pitch >> 3);
  const long long total = (long long)a.batch * H2 * W2 * groups;
  bf16* out = reinterpret_cast<bf16*>(a.out);
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const long long t = e / groups;
    const int g = (int)(e - t * groups);
    const int x = (int)(t % W2), y = (int)((t / W2) % H2), b = (int)(t / ((long long)W2 * H2));
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int col = g * 8 + i;
      float val = 0.f;
      if (col < 4 * a.C) {
        const int c = col >> 2, dy = (col >> 1) & 1, dx = col & 1;
        const int yy = min(2 * y + dy, a.h - 1), xx = min(2 * x + dx, a.w - 1);   // ReplicationPad2d bottom / right
        val = __ldg(a.img + (((size_t)b * a.C + c) * a.h + yy) * a.w + xx);
      } else if (col == 4 * a.C) {
        val = __ldg(a.sigma + b);
      }
      v[i] = val;
    }
    *reinterpret_cast<uint4*>(out + t * a.pitch + g * 8) =
        make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
}
}  // namespace

int ffd_pack_launch(const kair_ffd_pack_args& a, cudaStream_t st) {
  KB_CHECK(a.img && a.sigma && a.out, "kair_ffd_pack: null pointer");
  KB_CHECK(a.batch > 0 && a.C > 0 && a.h > 0 && a.w > 0 && a.pitch % 8 == 0 && a.pitch >= 4 * a.C + 1,
           "kair_ffd_pack: bad geometry (C=%d, pitch=%lld)", a.C, (long long)a.pitch);
  const long long total = (long long)a.batch * ((a.h + 1) / 2) * ((a.w + 1) / 2) * (a.pitch >> 3);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  ffd_pack_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

static int tiles_check(const kair_tiles_args& a, const char* who) {
  KB_CHECK(a.origins && a.tiles && a.n > 0 && a.C > 0 && a.th > 0 && a.tw > 0 && a.H > 0 && a.W > 0, "%s: bad arguments", who);
  return 0;
}

int tiles_gather_launch(const kair_tiles_args& a, cudaStream_t st) {
  if (tiles_check(a, "kair_tiles_gather")) return -1;
  KB_CHECK(a.img, "kair_tiles_gather: null image");
  const long long total = (long long)a.n * a.C * a.th * a.tw;
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  tiles_gather_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

int tiles_accumulate_launch(const kair_tiles_args& a, cudaStream_t st) {
  if (tiles_check(a, "kair_tiles_accumulate")) return -1;
  KB_CHECK(a.E, "kair_tiles_accumulate: null accumulator");
  const long long total = (long long)a.C * a.H * a.W;
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  tiles_accumulate_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  KB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace kb
