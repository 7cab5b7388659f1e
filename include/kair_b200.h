/*
 * kair_b200 C ABI -- sm_100a (B200) kernels for KAIR's image-restoration hot path.
 *
 * Every entry point takes plain device pointers, sizes and a CUDA stream handle (cudaStream_t passed as void*),
 * launches asynchronously on that stream, never synchronises, and returns 0 on success.  A non-zero return is
 * either a cudaError_t (> 0) or a kair error (< 0); kair_last_error() gives the message for the calling thread.
 * There is no CPU fallback: calling on a non-sm_100 device is an error.
 *
 * The reference (Owen1B/KAIR) has no native interface for this path -- it calls ATen ops from Python
 * (models/network_swinir.py:114-145, 239-279; models/network_dncnn.py:69-71; models/network_unet.py:59-76).
 * Its own plugin convention for native code is a torch extension of plain functions wrapped in
 * torch.autograd.Function (models/op/fused_act.py:10-72, models/op/fused_bias_act.cpp:7-21); this header is the
 * C-ABI layer under that convention.  Each entry point cites the reference lines whose arithmetic it replaces.
 *
 * Internal activation layout ("token layout"): tokens are NHWC pixels, t = (b*H + h)*W + w, channel-padded to a
 * multiple of 64 (C=180 -> 192), bf16 for GEMM operands and fp32 for the residual stream.  Padded channels are
 * zero, except an optional "ones column" used to obtain bias gradients from the weight-gradient GEMM.
 */
#ifndef KAIR_B200_H
#define KAIR_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* kair_last_error(void);
int kair_version(void);
/* 0 if the current CUDA device can run the kernels (compute capability 10.x). */
int kair_check_device(void);

/* ------------------------------------------------------------------------------------------------------------
 * kair_gemm_tn: D[M,N] = epilogue( A[M,K] * B[N,K]^T ) on tcgen05 tensor cores (bf16 in, fp32 accumulate in TMEM).
 *   conv == 0: A is a row-major bf16 matrix [M][a_pitch]  (nn.Linear: network_swinir.py:25-28, 121, 143)
 *   conv == 1: A is an NHWC bf16 image [B][H][W][a_pitch]; implicit-GEMM 3x3 stride-1 zero-padded convolution,
 *              K = 9 taps x Cin_pad, B packed as [N][9*Cin_pad] with k = tap*Cin_pad + ci
 *              (nn.Conv2d(C,C,3,1,1): network_swinir.py:465, 669, 729; basicblock.py:65)
 * Epilogue modes (what happens to acc = A*B^T, one row of the tile per thread):
 *   KAIR_EPI_LINEAR   v = act(acc + bias [+ resid]);                          outB = bf16(v), outF = v
 *                     with aux (and act == NONE): activation backward, v = (acc + bias) * (aux > 0 ? 1 : act_slope)
 *                     [+ resid], aux = the saved ReLU / LeakyReLU output (basicblock.py:70-77, :211-223)
 *   KAIR_EPI_GELU     u = acc + bias; outB2 = bf16(u); outB = bf16(gelu_erf(u))   (network_swinir.py:25-26)
 *   KAIR_EPI_RESID_LN v = acc + bias + resid; outF = v; outB = bf16((v-mean)*rstd); rowstat = rstd
 *                     (residual add + the next LayerNorm, network_swinir.py:245, 276-277)
 *   KAIR_EPI_DGELU    outB = bf16(acc * gelu_erf'(aux))                        (backward of :26)
 *   KAIR_EPI_LNBWD    LayerNorm backward fused with the residual gradient:
 *                     outF = resid + rstd*(acc - mean(acc) - xhat*mean(acc*xhat)), outB = bf16(outF), xhat = aux
 * ------------------------------------------------------------------------------------------------------------ */
enum { KAIR_EPI_LINEAR = 0, KAIR_EPI_GELU = 1, KAIR_EPI_RESID_LN = 2, KAIR_EPI_DGELU = 3, KAIR_EPI_LNBWD = 4 };
enum { KAIR_ACT_NONE = 0, KAIR_ACT_LRELU = 1 /* slope 0 == ReLU */ };

typedef struct kair_gemm_tn_args {
  const void* A;     /* bf16 */
  int64_t a_pitch;   /* elements per row / per pixel */
  const void* B;     /* bf16 [N][K_total], K_total = taps*K */
  int64_t b_pitch;   /* elements per row of B */
  int32_t M;         /* rows (tokens); for conv: batch*H*W */
  int32_t N;         /* output columns, multiple of bn */
  int32_t K;         /* reduction length per tap, multiple of 64 */
  int32_t bn;        /* N tile: 64, 128, 192 or 256 */
  int32_t mode;      /* KAIR_EPI_* */
  int32_t act;       /* KAIR_ACT_* (LINEAR mode only) */
  float act_slope;
  float ln_eps;
  const float* bias; /* [N] or NULL */
  const float* resid;
  int64_t resid_pitch;
  const void* aux;   /* bf16 */
  int64_t aux_pitch;
  float* rowstat;    /* [M] rstd: written by RESID_LN, read by LNBWD */
  void* outB;        /* bf16 [M][outB_pitch] or NULL */
  int64_t outB_pitch;
  void* outB2;       /* bf16, GELU mode pre-activation, or NULL */
  int64_t outB2_pitch;
  float* outF;       /* fp32 [M][outF_pitch] or NULL */
  int64_t outF_pitch;
  int32_t n_valid;   /* number of real (unpadded) columns, for LayerNorm statistics */
  int32_t ones_col;  /* column of outB forced to 1.0 (bias-gradient trick), -1 for none */
  int32_t conv;      /* 0 dense, 1 conv3x3 */
  int32_t batch, H, W;
  /* DropPath (network_swinir.py:204, 276-277; per-sample Bernoulli / keep_prob): sample = token / tokens_per_sample.
   * acc_scale[sample] multiplies (acc + bias) before the residual is added (LINEAR, RESID_LN);
   * outB_scale[sample] multiplies only the bf16 copy outB (LINEAR, LNBWD) -- the gradient entering the next branch. */
  const float* acc_scale;
  const float* outB_scale;
  int32_t tokens_per_sample;
  /* conv == 1: dilation of the 3x3 taps (tap (dy, dx) reads pixel p + dilation * (dy-1, dx-1), zero padding = dilation;
   * IRCNN's nn.Conv2d(.., padding=d, dilation=d), models/network_dncnn.py:99-111); 0 or 1 = dense taps */
  int32_t dilation;
} kair_gemm_tn_args;

int kair_gemm_tn(const kair_gemm_tn_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * kair_mlp_fused: the token-local MLP half of a Swin block in one launch (Mlp.forward, network_swinir.py:24-30, with the
 * residual add / DropPath of :277 and, optionally, the LayerNorm that follows, :245):
 *     u = X W1^T + b1;  g = gelu_erf(u);  v = (g W2^T + b2) * acc_scale[sample] + resid
 *     ln != 0: outF = v, outB = bf16((v - mean) * rstd) (ones_col forced to 1), rowstat = rstd
 *     ln == 0: exactly one of outF = v / outB = bf16(v) (ones_col forced to 1)
 * The 2C-wide hidden activations stay in shared memory / TMEM (the unfused pair of kair_gemm_tn launches writes u and
 * g to HBM and re-reads g); outU / outG (optional, training plans) receive bf16(u) and bf16(g) once for the backward
 * pass.  X is the un-affine normalised input (norm2's affine folded into W1 / b1 by kair_pack_linear); W1 [F][N] and
 * W2 [N][F] are packed bf16 operands; N = padded channels (64, 192 or 256), F = 2N.  Same rounding points and k-order
 * as the unfused pair: results are bit-identical to it.
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_mlp_fused_args {
  const void* X;        /* bf16 [M][x_pitch] */
  int64_t x_pitch;
  const void* W1;       /* bf16 [F][w1_pitch] */
  int64_t w1_pitch;
  const void* W2;       /* bf16 [N][w2_pitch] */
  int64_t w2_pitch;
  const float* bias1;   /* [F] or NULL */
  const float* bias2;   /* [N] or NULL */
  const float* resid;   /* fp32 [M][resid_pitch] or NULL */
  int64_t resid_pitch;
  float* outF;          /* fp32 [M][outF_pitch] or NULL */
  int64_t outF_pitch;
  void* outB;           /* bf16 [M][outB_pitch] or NULL */
  int64_t outB_pitch;
  void* outU;           /* bf16 [M][outU_pitch] or NULL: pre-activation (training) */
  int64_t outU_pitch;
  void* outG;           /* bf16 [M][outG_pitch] or NULL: GELU output with ones_col_g forced to 1 (training) */
  int64_t outG_pitch;
  float* rowstat;       /* [M] rstd (ln != 0) or NULL */
  const float* acc_scale; /* per-sample DropPath scale or NULL */
  int32_t tokens_per_sample;
  int32_t M, N, F;
  int32_t ln;           /* apply LayerNorm to v */
  int32_t n_valid;      /* real channels (LayerNorm statistics) */
  int32_t ones_col;     /* column of outB forced to 1.0, -1 for none */
  int32_t ones_col_g;   /* column of g forced to 1.0 (fc2 bias gradient of the backward pass), -1 for none */
  float ln_eps;
} kair_mlp_fused_args;
int kair_mlp_fused(const kair_mlp_fused_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * kair_gemm_nt: D[n1][taps*n2] += sum_t A[t][n1] * B[t (+tap shift)][n2]   (fp32 accumulate into D, which the
 * caller zero-fills).  The weight-gradient GEMM of nn.Linear / nn.Conv2d 3x3 (autograd of
 * network_swinir.py:19-21, 105-107, 465, 729); with a ones column in B it also yields the bias gradient.
 *   conv == 0: A, B are [T][pitch] bf16 token matrices.
 *   conv == 1: A, B are NHWC images; column block `tap` of D holds sum_p A[p] * B[p + (dy-1, dx-1)].
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_gemm_nt_args {
  const void* A;    /* bf16 [T][a_pitch] (output gradient) */
  int64_t a_pitch;
  const void* B;    /* bf16 [T][b_pitch] (layer input) */
  int64_t b_pitch;
  float* D;         /* fp32 [n1][d_pitch] */
  int64_t d_pitch;
  int32_t T;
  int32_t n1;       /* rows of D (multiple of 64) */
  int32_t n2;       /* columns of D per tap (multiple of bn) */
  int32_t bn;       /* 64, 128, 192, 256 */
  int32_t splits;   /* token-range splits, 0 = auto */
  int32_t conv;
  int32_t batch, H, W;
  int32_t dilation;  /* conv == 1: tap (dy, dx) pairs A[p] with B[p + dilation * (dy-1, dx-1)]; 0 or 1 = dense taps */
  int32_t swap;      /* conv == 1, n2 a multiple of 64: the same products, written TRANSPOSED -- D is fp32 [9*n2][d_pitch] with
                        D[tap*n2 + i][o] (kair_unpack_conv_grad `transposed`).  The tap-shifted input is then the M
                        operand (9*n2 rows in 128-row tiles) and the n1 output channels the N operand: for layers with 64
                        output channels every MMA row is useful, where the plain order wastes half of a 128-row accumulator */
} kair_gemm_nt_args;

int kair_gemm_nt(const kair_gemm_nt_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * kair_conv_chain -- L consecutive 3x3 convolutions, 64 -> 64 channels, padding 1, + bias, + leaky-ReLU, as ONE
 * persistent cooperative launch (one CTA per SM, weights of the running layer resident in shared memory, a grid-wide
 * barrier between layers).  Replaces the body loop of the reference's plain conv stacks at inference:
 * models/network_dncnn.py:40-71 (`sequential(m_head, *m_body, m_tail)`, body entries 'CR' / 'CBR' with the eval-mode
 * BatchNorm folded by kair_pack_conv), models/network_dncnn.py:103-148 (FDnCNN).  Results are bit-identical to L
 * launches of kair_gemm_tn (conv = 1, bn = 64, bias, act) over the same ping-pong buffers.
 *   layer l reads act(l & 1) and writes act((l + 1) & 1); the input is act0, the result is in act(L & 1).
 * `sync`: two zero-initialised 32-bit words owned by the caller (barrier arrivals, exit arrivals); the kernel leaves
 * them at zero.  Launches that share `sync` (or the activation buffers) must not overlap on the GPU.
 * ------------------------------------------------------------------------------------------------------------ */
#define KAIR_CONV_CHAIN_MAX_LAYERS 32
typedef struct kair_conv_chain_args {
  void* act0;           /* bf16 [batch*H*W][pitch], channels 0..63 used */
  void* act1;
  const void* Wp;       /* bf16 [L][64][9*64]: kair_pack_conv layout, k = tap*64 + input channel */
  const float* bias;    /* [L][64] or NULL */
  const float* slope;   /* [L] device: out = v > 0 ? v : v * slope  (0 = ReLU, 1 = no activation) */
  int32_t* sync;        /* 2 words, zero before the first launch */
  int64_t pitch;        /* elements per pixel row of act0 / act1 (>= 64, multiple of 8) */
  int32_t L, batch, H, W;
} kair_conv_chain_args;
int kair_conv_chain(const kair_conv_chain_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * kair_window_attn_fwd / _bwd: shifted-window multi-head attention core (head_dim padded to 32).  window = 8 runs on
 * tcgen05 / TMEM (shift 0 or 4); window = 2..7 (the 7x7 JPEG-artifact models) on a CUDA-core kernel (any shift < window).
 *   out[t, h*32+d] = sum_j softmax_j(q_i.k_j + table[rel(i,j)][h] + mask(i,j)) v_j
 * qkv holds [q | k | v] per token with columns [3][nH][32]; q is already multiplied by head_dim^-0.5.
 * Window partition / reverse (network_swinir.py:33-62), the cyclic shift (:249-252, 269-272) and the shifted-window
 * mask (:216-237, values {0,-100}) are index arithmetic inside the kernel; `table` is relative_position_bias_table
 * (:89-90, 127-130) in its state_dict layout.  Backward returns d(qkv) and accumulates d(table) (fp32, caller zeroes).
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_window_attn_args {
  const void* qkv;   /* bf16 [B*H*W][qkv_pitch] */
  int64_t qkv_pitch;
  void* out;         /* bf16 [B*H*W][out_pitch] (forward output) */
  int64_t out_pitch;
  const float* table; /* fp32 [(2*window-1)^2][num_heads] */
  const void* dout;  /* bf16, backward only */
  int64_t dout_pitch;
  void* dqkv;        /* bf16, backward only */
  int64_t dqkv_pitch;
  float* dtable;     /* fp32 [(2*window-1)^2][num_heads], backward only, accumulated */
  int32_t batch, H, W;
  int32_t num_heads, window, shift;
  /* fp32-accurate mode (SPLIT PRECISION below): qkv / out / dout / dqkv are fp32 arrays (same column layout and
   * pitches in elements) and every window size runs on the CUDA-core fp32 kernel (expf, no bf16 rounding anywhere) */
  int32_t fp32;
} kair_window_attn_args;

int kair_window_attn_fwd(const kair_window_attn_args* args, void* stream);
int kair_window_attn_bwd(const kair_window_attn_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Weight packing.  fp32 master parameters (state_dict layout) -> padded bf16 GEMM operands, batched over L layers
 * of identical shape (device arrays of L pointers).
 *   Wp[l][r][c] = row_scale[r] * W[l][row_map[r]][col_map[c]] * gamma[l][col_map[c]]      (0 where a map is -1)
 *   bp[l][r]    = row_scale[r] * (b[l][row_map[r]] + sum_k W[l][row_map[r]][k] * beta[l][k]) + bias_const[r]
 * i.e. the LayerNorm affine (network_swinir.py:199,205) that precedes qkv / fc1 and the q scale (:124) are folded
 * into the weights, so the GEMM consumes the un-affine normalised activations directly.  WpT is the transpose used
 * by the input-gradient GEMM.  kair_unpack_linear_grad maps d(Wp) (with d(bp) in column ones_col) back to
 * dW, db, d(gamma), d(beta) in the parameters' own layouts (overwrites).
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_pack_linear_args {
  const void* const* W;      /* [L] -> fp32 (out_f, in_f) */
  const void* const* b;      /* [L] -> fp32 (out_f), or NULL */
  const void* const* gamma;  /* [L] -> fp32 (in_f), or NULL */
  const void* const* beta;   /* [L] -> fp32 (in_f), or NULL */
  const int32_t* row_map;    /* [N_pad] */
  const int32_t* col_map;    /* [K_pad] */
  const float* row_scale;    /* [N_pad] or NULL */
  const float* bias_const;   /* [N_pad] or NULL */
  void* Wp;                  /* bf16 [L][N_pad][K_pad] */
  void* WpT;                 /* bf16 [L][K_pad][N_pad] or NULL */
  float* bp;                 /* fp32 [L][N_pad] or NULL */
  int32_t L, N_pad, K_pad, out_f, in_f;
  /* SPLIT PRECISION: terms = 0/1 packs plain bf16; terms = 3 / 6 packs the weight-side segments of a bf16 split,
   * Wp [L][N_pad][terms*K_pad] and WpT [L][K_pad][terms*N_pad] (see kair_split_terms) */
  int32_t terms;
} kair_pack_linear_args;
int kair_pack_linear(const kair_pack_linear_args* args, void* stream);

typedef struct kair_unpack_linear_args {
  const float* dWp;          /* fp32 [L][N_pad][K_pad] */
  const void* const* W;
  const void* const* gamma;
  const void* const* beta;
  void* const* dW;           /* [L] -> fp32 (out_f, in_f) */
  void* const* db;           /* [L] -> fp32 (out_f) or NULL */
  void* const* dgamma;       /* [L] -> fp32 (in_f) or NULL */
  void* const* dbeta;
  const int32_t* row_map;
  const int32_t* col_map;
  const float* row_scale;
  int32_t L, N_pad, K_pad, out_f, in_f, ones_col;
} kair_unpack_linear_args;
int kair_unpack_linear_grad(const kair_unpack_linear_args* args, void* stream);

/* conv weights (O, I, kh, kw) -> Wp [L][O_pad][taps*I_pad] (k = tap*I_pad + i) and the flipped / channel-swapped
 * WpT [L][I_pad][taps*O_pad] used by the input-gradient convolution; bp is the zero-padded bias. */
typedef struct kair_pack_conv_args {
  const void* const* W;
  const void* const* b;      /* or NULL */
  void* Wp;
  void* WpT;                 /* or NULL */
  float* bp;                 /* fp32 [L][O_pad] or NULL */
  int32_t L, O, I, O_pad, I_pad, taps;
  /* optional per-output-channel affine folded into the packed operands (eval-mode BatchNorm after the conv,
   * basicblock.py:68-69; the arithmetic of utils/utils_bnorm.py:44-58): Wp = W * oscale[o], bp = b * oscale[o] + oshift[o] */
  const void* const* oscale; /* [L] -> fp32 (O) or NULL */
  const void* const* oshift; /* [L] -> fp32 (O) or NULL */
  /* SPLIT PRECISION: terms = 3 / 6: Wp [L][O_pad][taps*terms*I_pad] with k = (tap*terms + seg)*I_pad + i and
   * WpT [L][I_pad][taps*terms*O_pad] likewise (the activation carries terms*I_pad channels per pixel) */
  int32_t terms;
} kair_pack_conv_args;
int kair_pack_conv(const kair_pack_conv_args* args, void* stream);

typedef struct kair_unpack_conv_args {
  const float* dWp;          /* fp32 [L][O_pad][taps*I_pad] */
  void* const* dW;           /* [L] -> fp32 (O, I, kh, kw) */
  void* const* db;           /* [L] -> fp32 (O) or NULL; taken from column ones_col of the centre tap */
  int32_t L, O, I, O_pad, I_pad, taps, ones_col, center_tap;
  int32_t dwp_pitch;         /* row pitch of dWp in elements; 0 = taps*I_pad (transposed: 0 = O_pad) */
  int32_t transposed;        /* dWp is fp32 [L][taps*I_pad][O_pad] (kair_gemm_nt `swap`); db is not available then */
} kair_unpack_conv_args;
int kair_unpack_conv_grad(const kair_unpack_conv_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Standalone LayerNorm over C channels of fp32 token rows (nn.LayerNorm(C, eps): network_swinir.py:683, 725).
 * Outputs are optional: affine result in fp32 / bf16, un-affine xhat in bf16, per-row mean and rstd.
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_layernorm_args {
  const float* x;
  const float* gamma;
  const float* beta;
  float* y_f32;
  void* y_bf16;
  void* xhat_bf16;
  float* mean;
  float* rstd;
  int64_t T;
  int32_t C, pitch, ones_col;
  float eps;
  /* SPLIT PRECISION: terms = 3 / 6: the bf16 outputs are split operands [T][terms*pitch] (kair_split_terms) */
  int32_t terms;
} kair_layernorm_args;
int kair_layernorm_fwd(const kair_layernorm_args* args, void* stream);

typedef struct kair_layernorm_bwd_args {
  const float* dy_f32;   /* either or both; summed */
  const void* dy_bf16;
  const float* x;
  const float* mean;
  const float* rstd;
  const float* gamma;
  const float* dx_add;   /* optional residual gradient added to dx */
  float* dx_f32;
  void* dx_bf16;
  float* dgamma;         /* accumulated (caller zeroes) */
  float* dbeta;
  int64_t T;
  int32_t C, pitch;
} kair_layernorm_bwd_args;
int kair_layernorm_bwd(const kair_layernorm_bwd_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Thin 3x3 convolutions between NCHW fp32 images (<= 4 channels) and token features.
 *   conv_in : feat[t][o] = act(b[o] + sum W[o][c][tap] * ((img[c][p+tap] - mean[c]) * range))
 *             (SwinIR.conv_first with the (x - mean) * img_range of network_swinir.py:809-810, :669;
 *              DnCNN / DRUNet head: basicblock.py:65, network_dncnn.py:63, network_unet.py:25)
 *   conv_out: out[c][p] = (b[c] + sum W[c][k][tap] * feat[p+tap][k] + skip_scale*(skip[c][p]-skip_mean[c]))
 *                          * out_scale + out_mean[c]
 *             (conv_last + "x + ..." + "/ img_range + mean": network_swinir.py:762, :835-837; DnCNN tail + x - n)
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_conv_in_args {
  const float* img;      /* (B, ci, H, W) */
  const float* W;        /* weights (O, ci, 3, 3) */
  const float* b;        /* (O) or NULL */
  const float* mean;     /* (ci) */
  float* out_f32;        /* [T][pitch] or NULL */
  void* out_bf16;        /* [T][pitch] or NULL */
  const float* dy_f32;   /* wgrad only */
  const void* dy_bf16;   /* wgrad only */
  float* dW;             /* wgrad only, accumulated (caller zeroes) */
  float* db;
  float range, act_slope;
  int32_t batch, ci, H, Wd /* width */, O, O_pad, pitch, act, ones_col;
} kair_conv_in_args;
int kair_conv_in_fwd(const kair_conv_in_args* args, void* stream);
int kair_conv_in_wgrad(const kair_conv_in_args* args, void* stream);

typedef struct kair_conv_out_args {
  const void* feat;      /* bf16 [T][pitch] */
  const float* W;        /* weights (co, C, 3, 3) */
  const float* b;        /* (co) or NULL */
  const float* skip;     /* (B, co, H, W) or NULL */
  const float* skip_mean;/* (co) */
  const float* out_mean; /* (co) */
  float* out;            /* (B, co, H, W) */
  const float* dout;     /* backward: d(out) */
  float* dfeat_f32;      /* dgrad outputs [T][pitch] */
  void* dfeat_bf16;
  float* dW;             /* wgrad, accumulated (caller zeroes) */
  float* db;
  float skip_scale, out_scale;
  int32_t batch, co, H, Wd /* width */, C, pitch;
  /* dgrad only, optional: feat = act(u) was produced by ReLU / LeakyReLU (network_dncnn.py:63-65); the gradient is
   * multiplied by act'(u) = (mask > 0 ? 1 : mask_slope), mask = the saved bf16 feat */
  const void* mask;
  float mask_slope;
  /* fp32-accurate mode: features as fp32 [T][pitch] instead of `feat` (forward and wgrad; CUDA-core fp32 kernels) */
  const float* feat_f32;
} kair_conv_out_args;
int kair_conv_out_fwd(const kair_conv_out_args* args, void* stream);
int kair_conv_out_dgrad(const kair_conv_out_args* args, void* stream);
int kair_conv_out_wgrad(const kair_conv_out_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * SPLIT PRECISION -- the fp32-accurate mode (north_star: 1e-4 max-abs forward, 1e-3 relative gradients against the
 * reference's fp32 arithmetic, models/network_swinir.py:114-145, 239-279).  The tensor cores only take bf16 (or
 * narrower) operands, so an fp32 value x is carried as a sum of bf16 components c0 = bf16(x), c1 = bf16(x - c0),
 * c2 = bf16(x - c0 - c1), and a product  X * W  becomes the sum of the significant component products, accumulated
 * in fp32 (TMEM) by ONE ordinary kair_gemm_tn launch over a K axis that is `terms` times longer:
 *     terms = 3:  activation segments [c0 | c1 | c0],            weight segments [C0 | C0 | C1]       (~2^-17)
 *     terms = 6:  activation segments [c0 | c1 | c2 | c0 | c1 | c0], weight [C0 | C0 | C0 | C1 | C1 | C2] (~2^-24)
 * kair_split_terms writes the activation side: fp32 rows -> bf16 rows of terms*n columns, optionally applying the
 * element-wise stage that precedes the next GEMM in the reference (GELU network_swinir.py:26 and its derivative,
 * LeakyReLU masks of the '3conv' / SR heads, DropPath row scales :204).  kair_pack_linear / kair_pack_conv with
 * `terms` write the weight side.  Weight-gradient GEMMs (kair_gemm_nt) are launched once per component pair on
 * column slices of the same buffers.
 * ------------------------------------------------------------------------------------------------------------ */
enum { KAIR_SPLIT_COPY = 0,      /* v = x                                   */
       KAIR_SPLIT_GELU = 1,      /* v = gelu_erf(x)                         */
       KAIR_SPLIT_DGELU = 2,     /* v = x * gelu_erf'(aux)                  */
       KAIR_SPLIT_ACT_BWD = 3,   /* v = x * (aux > 0 ? 1 : slope)           */
       KAIR_SPLIT_LRELU = 4      /* v = x > 0 ? x : x * slope               */ };
typedef struct kair_split_args {
  const float* x;        /* fp32 [T][x_pitch] */
  int64_t x_pitch;
  const float* aux;      /* fp32 [T][aux_pitch] or NULL */
  int64_t aux_pitch;
  void* out;             /* bf16 [T][out_pitch], out_pitch >= terms*n; segment s at columns [s*n, (s+1)*n) */
  int64_t out_pitch;
  float* out_f32;        /* optional fp32 copy of v, [T][out_f32_pitch] (may alias x) */
  int64_t out_f32_pitch;
  const float* row_scale;/* optional per-sample scale of v (DropPath), sample = row / rows_per_sample */
  int64_t T;
  int32_t n;             /* columns per segment (multiple of 8) */
  int32_t terms;         /* 3 or 6 (1: a plain bf16 copy) */
  int32_t op;            /* KAIR_SPLIT_* */
  int32_t ones_col;      /* column of v forced to 1.0, or -1 */
  int32_t rows_per_sample;
  float slope;
} kair_split_args;
int kair_split_terms(const kair_split_args* args, void* stream);

/* LayerNorm backward on split operands, fused with the residual gradient (the fp32-mode counterpart of
 * KAIR_EPI_LNBWD; autograd of network_swinir.py:245, 277 with gamma / beta folded into the next linear layer):
 *   dx = resid + rstd * (g - mean(g) - xhat * mean(g * xhat)),  xhat = the sum of the components stored in `xhat`
 * outF = dx (fp32); out = split(dx * row_scale) for the next GEMMs (optional). */
typedef struct kair_ln_bwd_split_args {
  const float* g;        /* fp32 [T][pitch]: gradient w.r.t. the normalised activations */
  const void* xhat;      /* bf16 split operand [T][xhat_pitch] written by kair_layernorm_fwd with the same terms */
  int64_t xhat_pitch;
  const float* rstd;     /* [T] */
  const float* resid;    /* fp32 [T][pitch] or NULL */
  float* outF;           /* fp32 [T][pitch] */
  void* out;             /* bf16 split [T][out_pitch] or NULL */
  int64_t out_pitch;
  const float* row_scale;
  int64_t T;
  int32_t C, pitch, terms, rows_per_sample;
} kair_ln_bwd_split_args;
int kair_ln_bwd_split(const kair_ln_bwd_split_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Thin (<= 4 channel) convolutions expressed as GEMMs.  kair_im2col3x3 expands an NCHW fp32 image into a bf16 token
 * matrix  out[t][tap*cpad + ch] = (img[ch][p(t) + dir*off(tap)] - mean[ch]) * scale  (0 outside the image, 1.0 in
 * column ones_col, 0 in the other padding columns), off(tap) = (tap/3 - 1, tap%3 - 1).
 *   dir = +1: input patches of conv_first  -> weight gradient = kair_gemm_nt(dY, patches)   (autograd of
 *             network_swinir.py:669, :811; network_dncnn.py:63; network_unet.py:25)
 *   dir = -1: shifted output gradients of conv_last -> input gradient = kair_gemm_tn(shifted, packed W) and weight
 *             gradient = kair_gemm_nt(features, shifted)   (autograd of network_swinir.py:762, :835)
 * kair_conv_out_pack converts between conv_last's (co, C, 3, 3) fp32 layout and those GEMM operands:
 *   mode 0: Wt[k][tap*cpad + c] = W[c][k][tap]  (bf16 [C_pad][pitch])
 *   mode 1: dW[c][k][tap] = D[k][tap*cpad + c], db[c] = D[ones_row][4*cpad + c]  (D fp32 [C_pad][pitch])
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_im2col_args {
  const float* img;      /* (B, c, H, W) */
  const float* mean;     /* (c) or NULL */
  void* out;             /* bf16 [B*H*W][pitch] */
  float scale;
  int32_t batch, c, H, Wd, cpad, pitch, dir, ones_col;
} kair_im2col_args;
int kair_im2col3x3(const kair_im2col_args* args, void* stream);

typedef struct kair_conv_out_pack_args {
  const float* W;        /* mode 0: (co, C, 3, 3) */
  void* Wt;              /* mode 0: bf16 [C_pad][pitch] */
  const float* D;        /* mode 1: fp32 [C_pad][pitch] */
  float* dW;             /* mode 1: (co, C, 3, 3) */
  float* db;             /* mode 1: (co) or NULL */
  int32_t mode, co, C, C_pad, cpad, pitch, ones_row;
} kair_conv_out_pack_args;
int kair_conv_out_pack(const kair_conv_out_pack_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Conv-net helpers (DnCNN: models/network_dncnn.py:40-71; DRUNet / UNetRes: models/network_unet.py:21-76).
 *
 * kair_space_to_depth / kair_depth_to_space: the 2x2 stride-2 pixel regrouping that turns
 *   nn.Conv2d(Cin, Cout, 2, 2, 0) (basicblock.py:495-501) into a dense GEMM over [T/4][4*Cin] rows, and
 *   nn.ConvTranspose2d(Cin, Cout, 2, 2, 0) (basicblock.py:471-477) into a dense GEMM producing [T/4][4*Cout] rows:
 *     coarse[(b*H/2 + h)*W/2 + w][tap*C + c] <-> fine[(b*H + 2h+dy)*W + 2w+dx][c],  tap = dy*2 + dx
 *   (H, W are the FINE resolution; both must be even).  s2d reads bf16 and writes bf16; d2s reads fp32 or bf16 and
 *   writes fp32 and/or bf16.  They are also each other's adjoints, so they serve the backward passes.
 * kair_colsum: out[c] += sum_t x[t][c]  (bias gradients: autograd of basicblock.py:65 with bias=True).
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_regroup_args {
  const void* fine_bf16;   /* s2d input  [B*H*W][fine_pitch] */
  void* coarse_bf16;       /* s2d output [B*H/2*W/2][coarse_pitch] */
  const float* coarse_in_f32;  /* d2s input (either) */
  const void* coarse_in_bf16;
  float* fine_out_f32;     /* d2s outputs (either or both) */
  void* fine_out_bf16;
  int64_t fine_pitch, coarse_pitch;
  int32_t batch, H, W, C;  /* C % 8 == 0 */
} kair_regroup_args;
int kair_space_to_depth(const kair_regroup_args* args, void* stream);
int kair_depth_to_space(const kair_regroup_args* args, void* stream);

typedef struct kair_colsum_args {
  const void* x_bf16;      /* either */
  const float* x_f32;
  float* out;              /* [C], accumulated (caller zeroes) */
  int64_t T, pitch;
  int32_t C;               /* even */
  float scale;
} kair_colsum_args;
int kair_colsum(const kair_colsum_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * BatchNorm2d over NHWC token rows (B.conv mode 'B': nn.BatchNorm2d(C, momentum=0.9, eps=1e-4, affine=True),
 * basicblock.py:68-69), training mode: biased batch statistics over all T = B*H*W tokens.
 *   kair_bn_stats       acc[0][c] += sum_t z, acc[1][c] += sum_t z^2            (double accumulators, caller zeroes once)
 *   kair_bn_finalize    mean, rstd = 1/sqrt(var_biased + eps); running_mean/var <- (1-m)*running + m*batch (unbiased
 *                       variance), num_batches_tracked += 1; acc is reset to zero
 *   kair_bn_apply       y = act(gamma * (z - mean) * rstd + beta)  -> bf16 (and/or fp32)
 *   kair_bn_bwd_reduce  g = dy * act'(y);  dbeta[c] += sum_t g,  dgamma[c] += sum_t g * xhat      (caller zeroes)
 *   kair_bn_bwd_apply   dz = gamma * rstd * (g - dbeta/T - xhat * dgamma/T) -> bf16;  dbias[c] += sum_t dz (optional)
 *   kair_bn_fold        eval mode: scale = gamma / sqrt(running_var + eps), shift = beta - running_mean * scale
 *                       (consumed by kair_pack_conv's oscale / oshift)
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_bn_args {
  const float* z;          /* fp32 [T][pitch] conv output (pre-BN) */
  const void* y;           /* bf16 [T][pitch] activated output (backward: activation mask) */
  const void* dy;          /* bf16 [T][pitch] gradient w.r.t. the activated output (or dy_f32) */
  const float* dy_f32;     /* fp32 alternative: the BN backward subtracts projections of dy, which amplifies bf16 rounding */
  void* out_bf16;          /* apply: y;  bwd_apply: dz */
  float* out_f32;          /* apply only, optional */
  const float* gamma;
  const float* beta;
  float* mean;             /* [C] */
  float* rstd;             /* [C] */
  double* acc;             /* [2][C] */
  float* running_mean;     /* finalize / fold */
  float* running_var;
  int64_t* num_batches_tracked; /* finalize, optional */
  float* dgamma;           /* [C] */
  float* dbeta;            /* [C] */
  float* dbias;            /* [C] optional (bwd_apply) */
  float* scale;            /* fold outputs [C] */
  float* shift;
  int64_t T, pitch;
  int32_t C;               /* C % 4 == 0, C <= 512 */
  int32_t act;             /* KAIR_ACT_* applied after the affine (slope 0 == ReLU) */
  float act_slope, eps, momentum;
} kair_bn_args;
int kair_bn_stats(const kair_bn_args* args, void* stream);
int kair_bn_finalize(const kair_bn_args* args, void* stream);
int kair_bn_apply(const kair_bn_args* args, void* stream);
int kair_bn_bwd_reduce(const kair_bn_args* args, void* stream);
int kair_bn_bwd_apply(const kair_bn_args* args, void* stream);
int kair_bn_fold(const kair_bn_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * SwinIR reconstruction heads (network_swinir.py:572-615, 740-762, 816-831).
 *   kair_pixel_shuffle  nn.PixelShuffle(r) on token features: in [B*H*W][r*r*C] with channel c*r*r + dy*r + dx
 *                       -> out [B*(rH)*(rW)][C] at pixel (r*h+dy, r*w+dx); bf16 in / bf16 out  (r = 2, 3, 4)
 *   kair_upsample_nearest2  torch.nn.functional.interpolate(scale_factor=2, mode='nearest') on token features
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_shuffle_args {
  const void* in;          /* bf16 */
  void* out;               /* bf16 */
  int64_t in_pitch, out_pitch;
  int32_t batch, H, W, C, r; /* H, W = LOW resolution; C = channels at the HIGH resolution */
  /* backward entry points only: the gradient written is multiplied by (mask > 0 ? 1 : mask_slope), mask being the
   * LeakyReLU output that fed the forward op (network_swinir.py:747, 829-830); NULL = no activation */
  const void* mask;        /* bf16, shaped like the backward output, or NULL */
  int64_t mask_pitch;
  float mask_slope;
  /* fp32-accurate mode, forward entry points only: `in` / `out` hold fp32 elements (pitches in elements) */
  int32_t fp32;
} kair_shuffle_args;
int kair_pixel_shuffle(const kair_shuffle_args* args, void* stream);
int kair_upsample_nearest2(const kair_shuffle_args* args, void* stream);
/* adjoints: `in` is the HIGH-resolution gradient [B*rH*rW][C], `out` the LOW-resolution one
 * ([B*H*W][r*r*C] for the shuffle; [B*H*W][C] = sum over the 2x2 footprint for nearest x2) */
int kair_pixel_shuffle_bwd(const kair_shuffle_args* args, void* stream);
int kair_upsample_nearest2_bwd(const kair_shuffle_args* args, void* stream);

/* UpsampleOneStep ('pixelshuffledirect', network_swinir.py:594-615, 819-822) tail: PixelShuffle(r) of the conv output
 * straight into the NCHW fp32 image with the final affine of SwinIR.forward (:837):
 *   out[b][c][r*h+dy][r*w+dx] = in[(b,h,w)][c*r*r + dy*r + dx] * scale + mean[c] */
typedef struct kair_shuffle_out_args {
  const void* in;          /* bf16 [B*H*W][in_pitch] */
  float* out;              /* fp32 (B, co, r*H, r*W) */
  const float* mean;       /* (co) */
  int64_t in_pitch;
  float scale;
  int32_t batch, H, W, co, r;
  int32_t fp32;            /* kair_pixel_shuffle_out only: `in` holds fp32 elements (fp32-accurate mode) */
} kair_shuffle_out_args;
int kair_pixel_shuffle_out(const kair_shuffle_out_args* args, void* stream);
/* adjoint of kair_pixel_shuffle_out: roles swap, `out` (the fp32 NCHW gradient of the image) is READ and `in`
 * (bf16 [B*H*W][in_pitch]) is WRITTEN: in[(b,h,w)][c*r*r + dy*r + dx] = out[b][c][r*h+dy][r*w+dx] * scale; columns
 * >= co*r*r are left untouched (they are the zero padding of the GEMM operand).  mean is ignored. */
int kair_pixel_shuffle_out_bwd(const kair_shuffle_out_args* args, void* stream);

/* y += x (fp32, n elements, n % 4 == 0); yb (bf16, optional) receives the updated y */
int kair_add_inplace(float* y, const float* x, void* yb, int64_t n, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * SPECT input pipeline and validation metrics (the fork's data path around the network: data/dataset_spect.py:57-132,
 * utils/utils_spect.py:3-35, models/model_base.py:355-377, utils/utils_image.py:629-700, :833-865).
 *   kair_spect_patches   one training batch from device-resident frame stacks [n_files][2][FH][FW] (the .dat layout:
 *                        anterior, posterior): per sample (file, top, left, mode) -> posterior view mirrored
 *                        (np.fliplr), P x P crop, augment_img mode 0..7, normalize_spect -> (B, 2, P, P) for L and H.
 *   kair_spect_normalize normalize_spect (inverse = 0) / denormalize_spect (inverse = 1); method 0 log, 1 anscombe,
 *                        2 linear, 3 identity.
 *   kair_spect_metrics   one image (C, h, w) of de-normalised counts E (estimate) and H (reference):
 *                        out[4c + 0] += sum (q(E) - q(H))^2 with q(x) = uint8(clip(x, 0, max_val) / max_val * 255)
 *                        out[4c + 1] += sum over the valid region of the SSIM map of q(E), q(H) (11x11 Gaussian 1.5)
 *                        out[4c + 2] += sum of the Poisson log-pmf of round(H) under mean max(E, 1e-9)
 *                        (fp64 sums, caller zeroes `out` [C][4]; PSNR = 20 log10(255 / sqrt(out0 / (h w))),
 *                         SSIM = out1 / ((h - 10)(w - 10)), pll = sum_c out2 / (C h w)).
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_spect_patch_args {
  const float* frames_L;  /* fp32 [n_files][2][FH][FW] or NULL */
  const float* frames_H;
  const int32_t* sample;  /* [B][4] = (file, top, left, mode) */
  float* out_L;           /* (B, 2, P, P) or NULL */
  float* out_H;
  int32_t B, P, FH, FW, method;
  float max_pixel;
} kair_spect_patch_args;
int kair_spect_patches(const kair_spect_patch_args* args, void* stream);

typedef struct kair_spect_norm_args {
  const float* x;
  float* y;               /* may alias x */
  int64_t n;
  int32_t method, inverse;
  float max_pixel;
} kair_spect_norm_args;
int kair_spect_normalize(const kair_spect_norm_args* args, void* stream);

typedef struct kair_spect_metric_args {
  const float* E;         /* (C, h, w) */
  const float* H;
  double* out;            /* [C][4], accumulated */
  int32_t C, h, w;
  float max_val;
} kair_spect_metric_args;
int kair_spect_metrics(const kair_spect_metric_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Tile / self-ensemble drivers (main_test_swinir.py:256-284 `test` with args.tile; utils/utils_model.py:127-185
 * test_split_fn / test_x8; utils/utils_image.py:408-426 augment_img_tensor4).
 *   kair_tiles_gather      tiles[n] = D_mode(img[:, y:y+h_src, x:x+w_src]), origins[n] = (y, x, mode); a stored tile is
 *                          (C, th, tw), its source window (th, tw) or, for the quarter-turn modes 1/3/5/7, (tw, th)
 *   kair_tiles_accumulate  E[:, y:y+.., x:x+..] += weight * D_mode^-1(tiles[n]); Wsum[y.., x..] += weight (optional):
 *                          one thread per output pixel, tiles visited in index order (deterministic sums)
 * img / E are single images (C, H, W) fp32; the batch dimension of the reference's loops is the tile index.
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_tiles_args {
  const float* img;        /* gather: (C, H, W) source */
  float* E;                /* accumulate: (C, H, W) sums */
  float* Wsum;             /* accumulate: (H, W) weights or NULL */
  float* tiles;            /* (n, C, th, tw): written by gather, read by accumulate */
  const int32_t* origins;  /* [n][3] = (y, x, mode) in the coordinates of img / E */
  int32_t n, C, th, tw, H, W;
  float weight;
} kair_tiles_args;
int kair_tiles_gather(const kair_tiles_args* args, void* stream);
int kair_tiles_accumulate(const kair_tiles_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * kair_axpby: v = a * x + b * y on fp32 token rows, optionally masked by a saved activation, with fp32 and / or bf16
 * outputs -- the scaled residual adds of the dense blocks ("x5 * 0.2 + x", "out * 0.2 + x": models/network_rrdbnet.py:
 * 49-56, 67-72) and the slice-and-mask step of their backward pass:
 *     out_f32[t][c]  = a * x[t][c] + b * y[t][c]                                   c < n
 *     out_bf16[t][c] = bf16(out * (mask ? (mask[t][c] > 0 ? 1 : mask_slope) : 1))   c < n;   0 for n <= c < n_pad
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_axpby_args {
  const float* x;
  int64_t x_pitch;
  const float* y;        /* or NULL */
  int64_t y_pitch;
  float* out_f32;        /* or NULL; may alias x or y */
  int64_t out_f32_pitch;
  void* out_bf16;        /* or NULL */
  int64_t out_bf16_pitch;
  const void* mask;      /* bf16 or NULL */
  int64_t mask_pitch;
  int64_t T;
  int32_t n, n_pad;      /* multiples of 8 */
  float a, b, mask_slope;
} kair_axpby_args;
int kair_axpby(const kair_axpby_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * kair_ffd_pack: the input side of FFDNet.forward (models/network_ffdnet.py:54-64) as one gather: ReplicationPad2d to
 * even H, W -> PixelUnShuffle(2) (models/basicblock.py:100-127: channel c*4 + dy*2 + dx) -> torch.cat with the noise
 * level map sigma.repeat(1, 1, H/2, W/2) -> bf16 token rows for the first convolution (implicit-GEMM, K padded to 64):
 *   out[(b, y, x)][c*4 + dy*2 + dx] = img[b][c][min(2y+dy, h-1)][min(2x+dx, w-1)],  out[..][4C] = sigma[b],  0 beyond
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_ffd_pack_args {
  const float* img;      /* (B, C, h, w) */
  const float* sigma;    /* (B) */
  void* out;             /* bf16 [B * H2 * W2][pitch], H2 = ceil(h / 2), W2 = ceil(w / 2) */
  int64_t pitch;         /* multiple of 8, >= 4C + 1 */
  int32_t batch, C, h, w;
} kair_ffd_pack_args;
int kair_ffd_pack(const kair_ffd_pack_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * Launch lists.  A plan (one forward or backward pass of a network at a fixed shape) is a static sequence of the entry
 * points above with pre-filled argument structs; kair_run_list replays such a sequence in ONE call, so the host cost
 * of a pass is one FFI crossing instead of one per kernel (SwinIR-M training: 558 kernels per step).
 *   op   : kair_op_id("kair_gemm_tn") etc. -- the index of an entry point in the library's own table
 *   args : the argument struct that entry point takes
 * Returns 0, or the failing launch's return code with kair_last_error() naming its index.
 * kair_add_inplace_s / kair_memcpy_d2d are struct-taking forms of the two scalar-argument helpers so that they can
 * appear in lists.
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_launch {
  int32_t op;
  int32_t reserved;
  const void* args;
} kair_launch;
int kair_op_id(const char* name);   /* -1 if `name` is not a listable entry point */
int kair_run_list(const kair_launch* list, int32_t n, void* stream);

typedef struct kair_add_args {
  float* y;          /* y += x */
  const float* x;
  void* yb;          /* bf16 shadow copy of the updated y, or NULL */
  int64_t n;
} kair_add_args;
int kair_add_inplace_s(const kair_add_args* args, void* stream);

typedef struct kair_memcpy_args {
  void* dst;
  const void* src;
  int64_t bytes;
} kair_memcpy_args;
int kair_memcpy_d2d(const kair_memcpy_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * The serial tail of a training step.
 *   kair_pixel_loss  CharbonnierLoss (models/loss.py:208-218; mode 0) or nn.L1Loss (mode 1) and its gradient in one pass:
 *                    loss[0] = mean(sqrt((x-y)^2 + eps)),  dx = d(loss)/dx.  `scratch` is two zero-initialised 8-byte
 *                    words owned by the caller (sum, arrival counter); the kernel re-arms them.
 *   kair_adam_ema    torch.optim.Adam (L2 weight decay, no amsgrad; torch/optim/adam.py `_single_tensor_adam`) for L tensors
 *                    in one launch + the EMA of ModelBase.update_E (models/model_base.py:247-252) on the updated values:
 *                    e = e*ema_decay + p*(1-ema_decay).  step_size = lr / (1-beta1^t), bc2_sqrt = sqrt(1-beta2^t).
 *                    chunk_tensor / chunk_offset list the (tensor, first element) of every 8192-element chunk.
 * ------------------------------------------------------------------------------------------------------------ */
typedef struct kair_pixel_loss_args {
  const float* x;
  const float* y;
  float* dx;         /* or NULL */
  float* loss;       /* 1 element */
  double* scratch;   /* 2 x 8 bytes, zero before the first call */
  int64_t n;
  float eps;
  int32_t mode;      /* 0 Charbonnier, 1 L1, 2/3/4 Poisson log-likelihood with linear / log / anscombe de-normalisation,
                        5 mean squared difference of the uint8 images round(clamp(v,0,1)*255) (tensor2uint + calculate_psnr,
                        utils/utils_image.py:296-300, 629-645; no gradient) */
  float max_pixel;   /* modes 2-4: PoissonLLLoss (models/loss.py:294-345): lambda = max(denorm(x), eps), k = max(denorm(y), 0),
                        loss = -mean(k log(lambda) - lambda); denorm as utils/utils_spect.py:21-35 */
} kair_pixel_loss_args;
int kair_pixel_loss(const kair_pixel_loss_args* args, void* stream);

/* kair_pixel_loss_noalloc -- NO COUNTERPART IN THE REFERENCE; added for callers
 * that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same arguments, modes and results as kair_pixel_loss, but `scratch` must be NULL: the partial sums live in a
 * file-scope __device__ array inside the library (8192 fp64 slots + a counter, 64.25 KiB, held by every GPU that loads
 * the library until exit).  The call itself allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   - limit: n <= KAIR_PIXEL_LOSS_NOALLOC_MAX_N (one slot per 8192 elements); larger n returns an error, use
 *     kair_pixel_loss with caller-owned scratch instead;
 *   - the loss is summed in a fixed order (no floating-point atomics): bit-identical from call to call; against
 *     kair_pixel_loss it differs by the fp64 summation order only (<= 1 ulp of the fp32 result);
 *   - every call on a device shares the array: calls that OVERLAP on the GPU (two streams, concurrently replayed graphs)
 *     race on it -- serialise them.  Nothing has to be cleared between calls (the kernel re-arms its counter);
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress): make it before stream
 *     capture begins. */
#define KAIR_PIXEL_LOSS_NOALLOC_MAX_N 67108864LL   /* 8192 slots x 8192 elements */
int kair_pixel_loss_noalloc(const kair_pixel_loss_args* args, void* stream);

typedef struct kair_adam_ema_args {
  void* const* p;            /* [L] -> fp32 parameters (updated in place) */
  const void* const* g;      /* [L] -> fp32 gradients */
  void* const* m;            /* [L] -> exp_avg */
  void* const* v;            /* [L] -> exp_avg_sq */
  void* const* e;            /* [L] -> EMA copies, or NULL */
  const int64_t* sizes;      /* [L] element counts */
  const int32_t* chunk_tensor;
  const int64_t* chunk_offset;
  int32_t n_chunks, chunk_elems;  /* chunk_elems must be 8192 */
  /* doubles: torch forms 1-beta1, 1-beta2 (and update_E 1-decay) in double before rounding to the fp32 compute type */
  double beta1, beta2, ema_decay;
  float step_size, eps, bc2_sqrt, weight_decay;
} kair_adam_ema_args;
int kair_adam_ema(const kair_adam_ema_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KAIR_B200_H */
