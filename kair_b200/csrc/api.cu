// extern "C" entry points declared in include/kair_b200.h
#include "../../include/kair_b200.h"

#include <string.h>

#include "host.h"

namespace kb {
int gemm_tn_launch(const kair_gemm_tn_args& a, cudaStream_t st);
int gemm_nt_launch(const kair_gemm_nt_args& a, cudaStream_t st);
int window_attn_fwd_launch(const kair_window_attn_args& a, cudaStream_t st);
int window_attn_bwd_launch(const kair_window_attn_args& a, cudaStream_t st);
int window_attn_small_fwd_launch(const kair_window_attn_args& a, cudaStream_t st);
int window_attn_small_bwd_launch(const kair_window_attn_args& a, cudaStream_t st);
int pack_linear_launch(const kair_pack_linear_args& a, cudaStream_t st);
int unpack_linear_launch(const kair_unpack_linear_args& a, cudaStream_t st);
int pack_conv_launch(const kair_pack_conv_args& a, cudaStream_t st);
int unpack_conv_launch(const kair_unpack_conv_args& a, cudaStream_t st);
int layernorm_fwd_launch(const kair_layernorm_args& a, cudaStream_t st);
int layernorm_bwd_launch(const kair_layernorm_bwd_args& a, cudaStream_t st);
int conv_in_fwd_launch(const kair_conv_in_args& a, cudaStream_t st);
int conv_in_wgrad_launch(const kair_conv_in_args& a, cudaStream_t st);
int conv_out_fwd_launch(const kair_conv_out_args& a, cudaStream_t st);
int conv_out_dgrad_launch(const kair_conv_out_args& a, cudaStream_t st);
int conv_out_wgrad_launch(const kair_conv_out_args& a, cudaStream_t st);
int im2col3x3_launch(const kair_im2col_args& a, cudaStream_t st);
int conv_out_pack_launch(const kair_conv_out_pack_args& a, cudaStream_t st);
int s2d_launch(const kair_regroup_args& a, cudaStream_t st);
int d2s_launch(const kair_regroup_args& a, cudaStream_t st);
int colsum_launch(const kair_colsum_args& a, cudaStream_t st);
int bn_stats_launch(const kair_bn_args& a, cudaStream_t st);
int bn_finalize_launch(const kair_bn_args& a, cudaStream_t st);
int bn_apply_launch(const kair_bn_args& a, cudaStream_t st);
int bn_bwd_reduce_launch(const kair_bn_args& a, cudaStream_t st);
int bn_bwd_apply_launch(const kair_bn_args& a, cudaStream_t st);
int bn_fold_launch(const kair_bn_args& a, cudaStream_t st);
int pixel_shuffle_launch(const kair_shuffle_args& a, cudaStream_t st);
int upsample_nearest2_launch(const kair_shuffle_args& a, cudaStream_t st);
int pixel_shuffle_out_launch(const kair_shuffle_out_args& a, cudaStream_t st);
int pixel_loss_launch(const kair_pixel_loss_args& a, cudaStream_t st);
int conv_chain_launch(const kair_conv_chain_args& a, cudaStream_t st);
int adam_ema_launch(const kair_adam_ema_args& a, cudaStream_t st);
int pixel_shuffle_out_bwd_launch(const kair_shuffle_out_args& a, cudaStream_t st);
int pixel_shuffle_bwd_launch(const kair_shuffle_args& a, cudaStream_t st);
int upsample_nearest2_bwd_launch(const kair_shuffle_args& a, cudaStream_t st);
int add_inplace_launch(float* y, const float* x, void* yb, long long n, cudaStream_t st);
int split_terms_launch(const kair_split_args& a, cudaStream_t st);
int mlp_fused_launch(const kair_mlp_fused_args& a, cudaStream_t st);
int spect_patches_launch(const kair_spect_patch_args& a, cudaStream_t st);
int spect_normalize_launch(const kair_spect_norm_args& a, cudaStream_t st);
int spect_metrics_launch(const kair_spect_metric_args& a, cudaStream_t st);
int tiles_gather_launch(const kair_tiles_args& a, cudaStream_t st);
int axpby_launch(const kair_axpby_args& a, cudaStream_t st);
int ffd_pack_launch(const kair_ffd_pack_args& a, cudaStream_t st);
int tiles_accumulate_launch(const kair_tiles_args& a, cudaStream_t st);
int ln_bwd_split_launch(const kair_ln_bwd_split_args& a, cudaStream_t st);
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
  // 8x8 windows: tcgen05 / TMEM kernel; smaller windows (7x7 JPEG models): CUDA-core kernel of attn_small.cu
  if (args->window != 8 || args->fp32) return kb::window_attn_small_fwd_launch(*args, (cudaStream_t)stream);
  return kb::window_attn_fwd_launch(*args, (cudaStream_t)stream);
}

int kair_window_attn_bwd(const kair_window_attn_args* args, void* stream) {
  if (!args) {
    kb::set_error("kair_window_attn_bwd: null args");
    return -1;
  }
  if (args->window != 8 || args->fp32) return kb::window_attn_small_bwd_launch(*args, (cudaStream_t)stream);
  return kb::window_attn_bwd_launch(*args, (cudaStream_t)stream);
}

#define KB_WRAP(NAME, TYPE, FN)                                  \
  int NAME(const TYPE* args, void* stream) {                     \
    if (!args) {                                                 \
      kb::set_error(#NAME ": null args");                        \
      return -1;                                                 \
    }                                                            \
    return kb::FN(*args, (cudaStream_t)stream);                  \
  }
KB_WRAP(kair_pack_linear, kair_pack_linear_args, pack_linear_launch)
KB_WRAP(kair_unpack_linear_grad, kair_unpack_linear_args, unpack_linear_launch)
KB_WRAP(kair_pack_conv, kair_pack_conv_args, pack_conv_launch)
KB_WRAP(kair_unpack_conv_grad, kair_unpack_conv_args, unpack_conv_launch)
KB_WRAP(kair_layernorm_fwd, kair_layernorm_args, layernorm_fwd_launch)
KB_WRAP(kair_layernorm_bwd, kair_layernorm_bwd_args, layernorm_bwd_launch)
KB_WRAP(kair_conv_in_fwd, kair_conv_in_args, conv_in_fwd_launch)
KB_WRAP(kair_conv_in_wgrad, kair_conv_in_args, conv_in_wgrad_launch)
KB_WRAP(kair_conv_out_fwd, kair_conv_out_args, conv_out_fwd_launch)
KB_WRAP(kair_conv_out_dgrad, kair_conv_out_args, conv_out_dgrad_launch)
KB_WRAP(kair_conv_out_wgrad, kair_conv_out_args, conv_out_wgrad_launch)
KB_WRAP(kair_im2col3x3, kair_im2col_args, im2col3x3_launch)
KB_WRAP(kair_conv_out_pack, kair_conv_out_pack_args, conv_out_pack_launch)
KB_WRAP(kair_space_to_depth, kair_regroup_args, s2d_launch)
KB_WRAP(kair_depth_to_space, kair_regroup_args, d2s_launch)
KB_WRAP(kair_colsum, kair_colsum_args, colsum_launch)
KB_WRAP(kair_bn_stats, kair_bn_args, bn_stats_launch)
KB_WRAP(kair_bn_finalize, kair_bn_args, bn_finalize_launch)
KB_WRAP(kair_bn_apply, kair_bn_args, bn_apply_launch)
KB_WRAP(kair_bn_bwd_reduce, kair_bn_args, bn_bwd_reduce_launch)
KB_WRAP(kair_bn_bwd_apply, kair_bn_args, bn_bwd_apply_launch)
KB_WRAP(kair_bn_fold, kair_bn_args, bn_fold_launch)
KB_WRAP(kair_pixel_shuffle, kair_shuffle_args, pixel_shuffle_launch)
KB_WRAP(kair_upsample_nearest2, kair_shuffle_args, upsample_nearest2_launch)
KB_WRAP(kair_pixel_shuffle_out, kair_shuffle_out_args, pixel_shuffle_out_launch)
KB_WRAP(kair_pixel_loss, kair_pixel_loss_args, pixel_loss_launch)
KB_WRAP(kair_conv_chain, kair_conv_chain_args, conv_chain_launch)
KB_WRAP(kair_adam_ema, kair_adam_ema_args, adam_ema_launch)
KB_WRAP(kair_pixel_shuffle_out_bwd, kair_shuffle_out_args, pixel_shuffle_out_bwd_launch)
KB_WRAP(kair_pixel_shuffle_bwd, kair_shuffle_args, pixel_shuffle_bwd_launch)
KB_WRAP(kair_upsample_nearest2_bwd, kair_shuffle_args, upsample_nearest2_bwd_launch)
KB_WRAP(kair_split_terms, kair_split_args, split_terms_launch)
KB_WRAP(kair_mlp_fused, kair_mlp_fused_args, mlp_fused_launch)
KB_WRAP(kair_spect_patches, kair_spect_patch_args, spect_patches_launch)
KB_WRAP(kair_spect_normalize, kair_spect_norm_args, spect_normalize_launch)
KB_WRAP(kair_spect_metrics, kair_spect_metric_args, spect_metrics_launch)
KB_WRAP(kair_tiles_gather, kair_tiles_args, tiles_gather_launch)
KB_WRAP(kair_axpby, kair_axpby_args, axpby_launch)
KB_WRAP(kair_ffd_pack, kair_ffd_pack_args, ffd_pack_launch)
KB_WRAP(kair_tiles_accumulate, kair_tiles_args, tiles_accumulate_launch)
KB_WRAP(kair_ln_bwd_split, kair_ln_bwd_split_args, ln_bwd_split_launch)
#undef KB_WRAP

int kair_add_inplace(float* y, const float* x, void* yb, int64_t n, void* stream) {
  return kb::add_inplace_launch(y, x, yb, (long long)n, (cudaStream_t)stream);
}

int kair_add_inplace_s(const kair_add_args* a, void* stream) {
  if (!a) {
    kb::set_error("kair_add_inplace_s: null args");
    return -1;
  }
  return kb::add_inplace_launch(a->y, a->x, a->yb, (long long)a->n, (cudaStream_t)stream);
}

int kair_memcpy_d2d(const kair_memcpy_args* a, void* stream) {
  if (!a || !a->dst || !a->src || a->bytes < 0) {
    kb::set_error("kair_memcpy_d2d: bad arguments");
    return -1;
  }
  cudaError_t e = cudaMemcpyAsync(a->dst, a->src, (size_t)a->bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    kb::set_error("kair_memcpy_d2d: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

// ---- launch lists: one table of (name, type-erased entry point); kair_op_id / kair_run_list index into it
namespace {
typedef int (*kb_entry)(const void*, void*);
struct KbOp {
  const char* name;
  kb_entry fn;
};
#define KB_OP(NAME) {#NAME, reinterpret_cast<kb_entry>(&NAME)}
const KbOp kb_ops[] = {
    KB_OP(kair_gemm_tn), KB_OP(kair_gemm_nt), KB_OP(kair_window_attn_fwd), KB_OP(kair_window_attn_bwd),
    KB_OP(kair_pack_linear), KB_OP(kair_unpack_linear_grad), KB_OP(kair_pack_conv), KB_OP(kair_unpack_conv_grad),
    KB_OP(kair_layernorm_fwd), KB_OP(kair_layernorm_bwd), KB_OP(kair_conv_in_fwd), KB_OP(kair_conv_in_wgrad),
    KB_OP(kair_conv_out_fwd), KB_OP(kair_conv_out_dgrad), KB_OP(kair_conv_out_wgrad), KB_OP(kair_im2col3x3),
    KB_OP(kair_conv_out_pack), KB_OP(kair_space_to_depth), KB_OP(kair_depth_to_space), KB_OP(kair_colsum),
    KB_OP(kair_bn_stats), KB_OP(kair_bn_finalize), KB_OP(kair_bn_apply), KB_OP(kair_bn_bwd_reduce),
    KB_OP(kair_bn_bwd_apply), KB_OP(kair_bn_fold), KB_OP(kair_pixel_shuffle), KB_OP(kair_upsample_nearest2),
    KB_OP(kair_pixel_shuffle_out), KB_OP(kair_pixel_loss), KB_OP(kair_adam_ema), KB_OP(kair_pixel_shuffle_out_bwd),
    KB_OP(kair_pixel_shuffle_bwd), KB_OP(kair_upsample_nearest2_bwd), KB_OP(kair_split_terms),
    KB_OP(kair_ln_bwd_split), KB_OP(kair_add_inplace_s), KB_OP(kair_memcpy_d2d), KB_OP(kair_mlp_fused),
    KB_OP(kair_spect_patches), KB_OP(kair_spect_normalize), KB_OP(kair_spect_metrics),
    KB_OP(kair_tiles_gather), KB_OP(kair_tiles_accumulate), KB_OP(kair_axpby), KB_OP(kair_ffd_pack), KB_OP(kair_conv_chain),
};
#undef KB_OP
constexpr int kb_n_ops = (int)(sizeof(kb_ops) / sizeof(kb_ops[0]));
}  // namespace

int kair_op_id(const char* name) {
  if (!name) return -1;
  for (int i = 0; i < kb_n_ops; ++i)
    if (strcmp(kb_ops[i].name, name) == 0) return i;
  return -1;
}

int kair_run_list(const kair_launch* list, int32_t n, void* stream) {
  if (!list || n < 0) {
    kb::set_error("kair_run_list: null list");
    return -1;
  }
  for (int i = 0; i < n; ++i) {
    const int op = list[i].op;
    if (op < 0 || op >= kb_n_ops) {
      kb::set_error("kair_run_list: launch %d has unknown op %d", i, op);
      return -1;
    }
    const int rc = kb_ops[op].fn(list[i].args, stream);
    if (rc != 0) {
      std::string msg = kb::get_error();
      kb::set_error("kair_run_list: launch %d (%s) failed: %s", i, kb_ops[op].name, msg.c_str());
      return rc;
    }
  }
  return 0;
}

}  // extern "C"
