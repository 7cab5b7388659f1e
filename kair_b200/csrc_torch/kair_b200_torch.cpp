// Thin torch C++ extension over the C ABI (include/kair_b200.h): the binding style the reference uses for its own CUDA
// ops (models/op/fused_act.py:10-72, models/op/upfirdn2d.py: `torch.utils.cpp_extension` modules called from
// autograd.Functions).  It is a SECOND binding of the same extern "C" entry points that kair_b200/_lib.py binds with
// ctypes: tensors in, tensors out, the current CUDA stream of the tensors' device, errors as c10::Error.
//
//   torch.ops.kair_b200.version() -> int
//   torch.ops.kair_b200.window_attn_fwd(qkv, table, batch, H, W, heads, window, shift) -> out
//   torch.ops.kair_b200.window_attn_bwd(qkv, table, dout, batch, H, W, heads, window, shift) -> (dqkv, dtable)
//   torch.ops.kair_b200.pixel_loss(x, y, eps, mode, max_pixel) -> (loss, dx)
//   torch.ops.kair_b200.run_list(anchor, list_address, n) -> ()     replay of a recorded launch list (kair_run_list)
//
// No kernel lives here: everything forwards to libkair_b200.so.
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>

#include <tuple>

#include "../../include/kair_b200.h"

namespace {

void check(int rc, const char* what) {
  if (rc != 0) {
    const char* msg = kair_last_error();
    TORCH_CHECK(false, "kair_b200.", what, " failed (rc=", rc, "): ", msg ? msg : "?");
  }
}

void* stream_of(const at::Tensor& t) { return (void*)at::cuda::getCurrentCUDAStream(t.device().index()).stream(); }

void need_cuda(const at::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), "kair_b200: ", name, " must be a CUDA tensor (there is no CPU fallback)");
}

int64_t version() { return kair_version(); }

kair_window_attn_args attn_args(const at::Tensor& qkv, const at::Tensor& table, int64_t batch, int64_t H, int64_t W,
                                int64_t heads, int64_t window, int64_t shift) {
  need_cuda(qkv, "qkv");
  need_cuda(table, "table");
  TORCH_CHECK(qkv.dim() == 2 && qkv.stride(1) == 1, "kair_b200: qkv must be [tokens, 3*heads*32] with unit column stride");
  TORCH_CHECK(qkv.scalar_type() == at::kBFloat16 || qkv.scalar_type() == at::kFloat, "kair_b200: qkv must be bf16 or fp32");
  TORCH_CHECK(table.scalar_type() == at::kFloat && table.is_contiguous(), "kair_b200: table must be contiguous fp32");
  kair_window_attn_args a = {};
  a.qkv = qkv.data_ptr();
  a.qkv_pitch = qkv.stride(0);
  a.table = table.data_ptr<float>();
  a.batch = (int32_t)batch;
  a.H = (int32_t)H;
  a.W = (int32_t)W;
  a.num_heads = (int32_t)heads;
  a.window = (int32_t)window;
  a.shift = (int32_t)shift;
  a.fp32 = qkv.scalar_type() == at::kFloat ? 1 : 0;
  return a;
}

at::Tensor window_attn_fwd(const at::Tensor& qkv, const at::Tensor& table, int64_t batch, int64_t H, int64_t W,
                           int64_t heads, int64_t window, int64_t shift) {
  need_cuda(qkv, "qkv");
  c10::cuda::CUDAGuard guard(qkv.device());
  kair_window_attn_args a = attn_args(qkv, table, batch, H, W, heads, window, shift);
  at::Tensor out = at::empty({qkv.size(0), qkv.size(1) / 3}, qkv.options());
  a.out = out.data_ptr();
  a.out_pitch = out.stride(0);
  check(kair_window_attn_fwd(&a, stream_of(qkv)), "window_attn_fwd");
  return out;
}

std::tuple<at::Tensor, at::Tensor> window_attn_bwd(const at::Tensor& qkv, const at::Tensor& table, const at::Tensor& dout,
                                                   int64_t batch, int64_t H, int64_t W, int64_t heads, int64_t window,
                                                   int64_t shift) {
  need_cuda(qkv, "qkv");
  c10::cuda::CUDAGuard guard(qkv.device());
  kair_window_attn_args a = attn_args(qkv, table, batch, H, W, heads, window, shift);
  need_cuda(dout, "dout");
  TORCH_CHECK(dout.scalar_type() == qkv.scalar_type() && dout.dim() == 2 && dout.stride(1) == 1,
              "kair_b200: dout must match qkv's dtype and be [tokens, heads*32]");
  at::Tensor dqkv = at::zeros_like(qkv, qkv.options().memory_format(at::MemoryFormat::Contiguous));
  at::Tensor dtable = at::zeros_like(table);
  a.dout = dout.data_ptr();
  a.dout_pitch = dout.stride(0);
  a.dqkv = dqkv.data_ptr();
  a.dqkv_pitch = dqkv.stride(0);
  a.dtable = dtable.data_ptr<float>();
  check(kair_window_attn_bwd(&a, stream_of(qkv)), "window_attn_bwd");
  return {dqkv, dtable};
}

std::tuple<at::Tensor, at::Tensor> pixel_loss(const at::Tensor& x, const at::Tensor& y, double eps, int64_t mode,
                                              double max_pixel) {
  need_cuda(x, "x");
  need_cuda(y, "y");
  TORCH_CHECK(x.scalar_type() == at::kFloat && y.scalar_type() == at::kFloat && x.is_contiguous() && y.is_contiguous() &&
                  x.sizes() == y.sizes(),
              "kair_b200: pixel_loss needs contiguous fp32 x, y of one shape");
  c10::cuda::CUDAGuard guard(x.device());
  at::Tensor dx = at::empty_like(x);
  at::Tensor loss = at::zeros({1}, x.options());
  at::Tensor scratch = at::zeros({2}, x.options().dtype(at::kDouble));
  kair_pixel_loss_args a = {};
  a.x = x.data_ptr<float>();
  a.y = y.data_ptr<float>();
  a.dx = dx.data_ptr<float>();
  a.loss = loss.data_ptr<float>();
  a.scratch = scratch.data_ptr<double>();
  a.n = x.numel();
  a.eps = (float)eps;
  a.mode = (int32_t)mode;
  a.max_pixel = (float)max_pixel;
  check(kair_pixel_loss(&a, stream_of(x)), "pixel_loss");
  return {loss, dx};
}

// `list_address`: address of a host array of kair_launch records whose argument structs the caller keeps alive
// (kair_b200.ops.LaunchList); `anchor`: any tensor on the device the list belongs to (selects device and stream)
void run_list(const at::Tensor& anchor, int64_t list_address, int64_t n) {
  need_cuda(anchor, "anchor");
  c10::cuda::CUDAGuard guard(anchor.device());
  check(kair_run_list(reinterpret_cast<const kair_launch*>(list_address), (int32_t)n, stream_of(anchor)), "run_list");
}

}  // namespace

TORCH_LIBRARY(kair_b200, m) {
  m.def("version() -> int", &version);
  m.def("window_attn_fwd(Tensor qkv, Tensor table, int batch, int H, int W, int heads, int window, int shift) -> Tensor",
        &window_attn_fwd);
  m.def("window_attn_bwd(Tensor qkv, Tensor table, Tensor dout, int batch, int H, int W, int heads, int window, int shift)"
        " -> (Tensor, Tensor)",
        &window_attn_bwd);
  m.def("pixel_loss(Tensor x, Tensor y, float eps, int mode, float max_pixel) -> (Tensor, Tensor)", &pixel_loss);
  m.def("run_list(Tensor anchor, int list_address, int n) -> ()", &run_list);
}
