// Host-side helpers shared by the launchers: error reporting and TMA tensor-map encoding.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

namespace kb {

// thread-local last error string (returned through kair_last_error())
void set_error(const char* fmt, ...);
const char* get_error();

#define KB_CHECK(cond, ...)          \
  do {                               \
    if (!(cond)) {                   \
      kb::set_error(__VA_ARGS__);    \
      return -1;                     \
    }                                \
  } while (0)

#define KB_CUDA(expr)                                                                   \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      kb::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return (int)_e;                                                                   \
    }                                                                                   \
  } while (0)

// Encode a tiled tensor map. dims/box are innermost-first; strides_bytes has rank-1 entries (dims 1..rank-1).
// elem_strides may be null (all 1). Returns 0 on success.
// l2_promotion: bytes the L2 fetches from DRAM around every accessed 128-byte line.  256 (default) pays when the
// neighbouring line is consumed soon after; a kernel that comes back for the neighbour much later (window_attn_bwd walks
// the channel dimension in its OUTER loop) must ask for 128, or half of what it pulls from DRAM is thrown away.
int make_tmap(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
              const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* elem_strides,
              CUtensorMapSwizzle swizzle, int l2_promotion = 256);

// Programmatic dependent launch: the grid may become resident (barrier init, TMEM allocation, descriptor prefetch)
// while the previous kernel of the stream drains; the kernel executes pdl_wait() before its first global access.
// KB_PDL=0 in the environment launches with plain stream order.
bool pdl_enabled();
template <typename P>
inline cudaError_t launch_pdl(void (*kern)(const P), dim3 grid, int block, size_t smem, cudaStream_t st, const P& p) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, p);
}

template <typename P>
inline cudaError_t launch_pdl(void (*kern)(const P), int grid, int block, size_t smem, cudaStream_t st, const P& p) {
  return launch_pdl(kern, dim3((unsigned)grid), block, smem, st, p);
}

// Development hooks (ablation masks KB_*_DBG, clock64 timelines KB_*_TS, KB_NT_BM) are compiled only into builds made
// with -DKB_DEVEL (python -m kair_b200.build --devel): a stray environment variable must never change what a
// production launch computes or insert a cudaMalloc / synchronous copy into it.
#ifdef KB_DEVEL
#define KB_DEV_ENV(name) getenv(name)
#else
#define KB_DEV_ENV(name) ((const char*)nullptr)
#endif

int sm_count();
int check_device();  // 0 if current device is sm_100; error otherwise

}  // namespace kb
