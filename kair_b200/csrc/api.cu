// extern "C" entry points declared in include/kair_b200.h
#include "../../include/kair_b200.h"

#include "host.h"

namespace kb {
int gemm_tn_launch(const kair_gemm_tn_args& a, cudaStream_t st);
int gemm_nt_launch(const kair_gemm_nt_args& a, cudaStream_t st);
int window_attn_fwd_launch(const kair_window_attn_args& a, cudaStream_t st);
int window_attn_bwd_launch(const kair_window_attn_args& a, cudaStream_t st);
}

extern "C" {

const char* kair_last_error(void) { return kb::get_error(); }
int kair_version(void) { return 1; }
int kair_check_device(void) { return kb::check_device(); }

int kair_gemm_tn(const kair_gemm_tn_args* args, void* stream) {
  if (!args) {
    kb::set_error("kair_gemm_tn: null args");
    return -1;
  }
  return kb::gemm_tn_launch(*args, (cudaStream_t)stream);
}

int kair_gemm_nt(const kair_gemm_nt_args* args, void* stream) {
  if (!args) {
    kb::set_error("kair_gemm_nt: null args");
    return -1;
  }
  return kb::gemm_nt_launch(*args, (cudaStream_t)stream);
}

int kair_window_attn_fwd(const kair_window_attn_args* args, void* stream) {
  if (!args) {
    kb::set_error("kair_window_attn_fwd: null args");
    return -1;
  }
  return kb::window_attn_fwd_launch(*args, (cudaStream_t)stream);
}

int kair_window_attn_bwd(const kair_window_attn_args* args, void* stream) {
  if (!args) {
    kb::set_error("kair_window_attn_bwd: null args");
    return -1;
  }
  return kb::window_attn_bwd_launch(*args, (cudaStream_t)stream);
}

}  // extern "C"
