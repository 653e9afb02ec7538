/* vcn_b200.h - C ABI of libvcn_b200.so: B200 (sm_100a) kernels for the StyleGAN2-ADA synthesis
 * hot path of VisCoIN-Next.
 *
 * Drop-in boundary.  The reference binds its two CUDA plugins through pybind11 torch extensions
 * (stylegan2_ada/torch_utils/ops/bias_act.cpp:32,94-97 and upfirdn2d.cpp:16,98-101) and calls
 * ATen/cuDNN for the convolution (conv2d_gradfix.py:38,43).  This library replaces all three with
 * plain-C entry points: raw device pointers, sizes, a dtype code and the CUDA stream; no torch
 * types.  The caller allocates every output and workspace (reference: bias_act.cpp:55 and
 * upfirdn2d.cpp:35 allocate inside C++); a NULL pointer means "absent" (reference: empty tensor,
 * bias_act.py:39).  Every function returns 0 on success and a non-zero vcn_status otherwise, with
 * a thread-local message in vcn_last_error() (reference: TORCH_CHECK exceptions).  Launches are
 * asynchronous on the stream given; nothing here synchronises the device or allocates device
 * memory.  There is no CPU fallback.
 */
#ifndef VCN_B200_H_
#define VCN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VCN_ABI_VERSION 2

typedef enum {
  VCN_OK = 0,
  VCN_ERR_INVALID = 1,      /* bad argument (shape, dtype, alignment, NULL) */
  VCN_ERR_UNSUPPORTED = 2,  /* valid request this build has no kernel for */
  VCN_ERR_CUDA = 3,         /* CUDA runtime / driver error, text in vcn_last_error() */
  VCN_ERR_DEVICE = 4        /* device is not sm_100 */
} vcn_status;

typedef enum { VCN_F32 = 0, VCN_F16 = 1, VCN_BF16 = 2, VCN_F64 = 3 } vcn_dtype;

/* Activation ids = the reference's cuda_idx column (bias_act.py:23-33). */
typedef enum {
  VCN_ACT_LINEAR = 1, VCN_ACT_RELU = 2, VCN_ACT_LRELU = 3, VCN_ACT_TANH = 4, VCN_ACT_SIGMOID = 5,
  VCN_ACT_ELU = 6, VCN_ACT_SELU = 7, VCN_ACT_SOFTPLUS = 8, VCN_ACT_SWISH = 9
} vcn_act;

/* ---- library ---------------------------------------------------------------------------- */
int vcn_abi_version(void);
const char* vcn_last_error(void);
/* Number of kernels this library has launched in this process (bench.py's gpu_launches). */
uint64_t vcn_launch_count(void);
/* Profiling aid: with VCN_CONV_PROFILE=1 in the environment vcn_conv_igemm accumulates per-role cycle counters
 * ([0] epilogue busy, [1] epilogue waiting, [2] MMA waiting for an accumulator, [3] MMA waiting for operands,
 * [4] tiles, [5] TMA producer waiting for a stage, [6] k-blocks); synchronises the device. */
int vcn_conv_profile_counters(uint64_t out[8], int reset);
/* 0 iff `device` exists and is compute capability 10.x; fills sm count / L2 bytes when non-NULL. */
int vcn_device_info(int device, int* sm_count, int64_t* l2_bytes, int* cc_major, int* cc_minor);

/* ---- bias_act ---------------------------------------------------------------------------
 * Replaces `bias_act(x, b, xref, yref, dy, grad, dim, act, alpha, gain, clamp)`
 * (bias_act.cpp:32-90, kernel bias_act.cu:23-147).  x is a dense tensor seen as a flat array of
 * size_x elements; the bias index of flat element i is (i / step_b) % size_b (bias_act.cu:44),
 * step_b = x.stride(dim).
 *   grad 0: y = clamp(act(x + b) * gain)
 *   grad 1: x holds dy;        y = dx   (needs yref, or xref(+b) for swish)
 *   grad 2: x holds d_dx, dy;  y = d_x  (second-order; only acts with a 2nd grad)
 * New vs the reference: when `db` is non-NULL the per-channel sum of the result (the reference's
 * separate `dx.sum(...)`, bias_act.py:173,204) is produced by the same call; it needs a
 * workspace of vcn_bias_act_workspace() bytes.  db has size_b elements of `dtype`.
 */
size_t vcn_bias_act_workspace(int dtype, int64_t size_x, int64_t size_b, int64_t step_b);
int vcn_bias_act(const void* x, const void* b, const void* xref, const void* yref, const void* dy,
                 void* y, void* db, void* workspace, size_t workspace_bytes,
                 int dtype, int grad, int act, float alpha, float gain, float clamp,
                 int64_t size_x, int64_t size_b, int64_t step_b, void* stream);

/* Allocation-free variant of vcn_bias_act.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments and results as vcn_bias_act without the workspace pair: the partial sums of the fused `db`
 * reduction go to a file-scope `static __device__` array inside the library (VCN_BIAS_ACT_STATIC_SCRATCH_BYTES = 1 MiB,
 * held by every GPU that loads the library until the process exits; its device address is resolved once per device with
 * cudaGetSymbolAddress).  The call itself performs no cudaMalloc/cudaFree and calls no allocating routine.
 * A shape whose vcn_bias_act_workspace() exceeds the array returns VCN_ERR_INVALID with a message (1 MiB = 262144 fp32
 * partial sums, one per (sample, channel, row chunk); the largest hot-path call, [16,128,256,256], needs 128 KiB).
 * All calls on a device share the array and it keeps the last call's contents; nothing in it has to start at zero
 * (every slot read is written earlier in the same call).  Two calls whose kernels overlap on the GPU (different
 * streams, no ordering between them) RACE on it: order them, or use vcn_bias_act with separate workspaces. */
#define VCN_BIAS_ACT_STATIC_SCRATCH_BYTES (1u << 20)
size_t vcn_bias_act_static_scratch_bytes(void);
int vcn_bias_act_static(const void* x, const void* b, const void* xref, const void* yref, const void* dy,
                        void* y, void* db,
                        int dtype, int grad, int act, float alpha, float gain, float clamp,
                        int64_t size_x, int64_t size_b, int64_t step_b, void* stream);

/* ---- upfirdn2d ---------------------------------------------------------------------------
 * Replaces `upfirdn2d(x, f, upx, upy, downx, downy, padx0, padx1, pady0, pady1, flip, gain)`
 * (upfirdn2d.cpp:16-94, kernels upfirdn2d.cu:29-200).  x is [n, c, in_h, in_w] with element
 * strides xs[4]; y is [n, c, out_h, out_w] with strides ys[4]; out_w must equal
 * (in_w*upx + padx0 + padx1 - fw + downx) / downx (upfirdn2d.cpp:32-33) and likewise out_h.
 * f is a dense fp32 [fh, fw] filter in device memory.  flip != 0 = correlation with f as given,
 * flip == 0 = true convolution (upfirdn2d.cu:125-127).
 * Optional fused epilogue (new; all NULL/0 = plain reference semantics):
 *   y = upfirdn(x) * scale[n*c_idx] + noise[n?, h, w]   then   bias_act(b, act, ...)
 * used by the up=2 modulated convolution (networks.py:69-71 fma + :302-304 bias_act).
 */
typedef struct {
  const float* scale;        /* fp32 [n, c] per-sample per-channel multiplier (demod coefs) or NULL */
  const float* noise;        /* fp32 [noise_n, out_h, out_w] added after scaling, or NULL */
  int64_t noise_stride_n;    /* elements between samples; 0 = one plane broadcast over n */
  const float* bias;         /* fp32 [c] or NULL */
  int act;                   /* vcn_act; 0 = no bias/activation stage at all */
  float alpha, act_gain, clamp;  /* as vcn_bias_act; clamp < 0 disables */
} vcn_upfirdn2d_epilogue;

int vcn_upfirdn2d(const void* x, const float* f, void* y, int dtype,
                  int n, int c, int in_h, int in_w, const int64_t xs[4],
                  int out_h, int out_w, const int64_t ys[4],
                  int fh, int fw, int upx, int upy, int downx, int downy,
                  int padx0, int padx1, int pady0, int pady1, int flip, float gain,
                  const vcn_upfirdn2d_epilogue* epi, void* stream);

/* ---- modulated convolution -----------------------------------------------------------------
 * Replaces the convolution under `modulated_conv2d` (training/networks.py:27-84), i.e. the ATen/cuDNN calls of
 * conv2d_gradfix.py:38,43 reached through conv2d_resample.py:59-154, with a TMA-fed tcgen05/TMEM implicit GEMM.
 * Formulation = the reference's non-fused branch (networks.py:63-72), proven equal to the fused one (SURVEY
 * App. B):  xhat = x * s[n,ci]  ->  shared-weight convolution  ->  * d[n,co] + noise  ->  bias_act.
 *
 *   vcn_modulate_pack   x (any strides, f32/f16/bf16) * styles[n,ci] -> xhat, bf16 NHWC [n,h,w,cin_pad]
 *                       (cin_pad = cin rounded up to 8; pad channels are written as 0)
 *   vcn_pack_weights    W [cout,cin,kh,kw] (f32/bf16, dense) -> wt bf16 [kh*kw][rows][cols_pad] K-major per tap
 *                       (rows,cols) = (cout,cin), or (cin,cout) with transpose_io (input-gradient GEMM); optional
 *                       wsq[cout,cin] = sum_k W^2 (fp32) for the demodulation coefficients
 *   vcn_demod_coefs     d[n,co] = rsqrt(sum_ci (s[n,ci])^2 * wsq[co,ci] + 1e-8)       (networks.py:55-58)
 *   vcn_conv_igemm      the implicit GEMM; accumulates in fp32 (TMEM), operands bf16
 */
typedef enum {
  VCN_CONV_FPROP = 0,   /* stride-1 correlation, padding (kh/2, kw/2): y[oy,ox] = sum x[oy+ky-ph, ox+kx-pw] W[ky,kx] */
  VCN_CONV_TCONV2 = 1,  /* 3x3 stride-2 transposed conv, no padding: y[2i+ky, 2j+kx] += x[i,j] W[ky,kx]; out (2h+1, 2w+1)
                           (conv2d_resample.py:125-138 with flip_weight=False) */
  VCN_CONV_DGRAD = 2,   /* stride-1 correlation with the spatially flipped taps (input gradient of FPROP when wt was
                           packed with transpose_io) */
  VCN_CONV_SCONV2 = 3   /* 3x3 stride-2 correlation, no padding: y[i,j] = sum x[2i+ky, 2j+kx] W[ky,kx]; (2h+1, 2w+1) -> (h, w)
                           (input gradient of TCONV2 when wt was packed with transpose_io) */
} vcn_conv_kind;

typedef struct {
  int kind;                 /* vcn_conv_kind */
  int n, h, w;              /* xhat is bf16 NHWC [n, h, w, cin_pad] */
  int cin_pad;              /* multiple of 8 */
  int cout;                 /* rows of wt per tap */
  int kh, kw;               /* 1x1 or 3x3 (TCONV2: 3x3) */
  int out_h, out_w;         /* must match the kind: (h, w), (2h+1, 2w+1) or ((h-1)/2, (w-1)/2) */
  int y_dtype;              /* VCN_F32 / VCN_BF16 / VCN_F16 */
  int64_t ys[4];            /* element strides of y for (n, c, h, w): NCHW, channels_last or padded pitch */
  int w_per_sample;         /* wt has a leading sample dimension [n][tap][cout][cin_pad] (per-sample modulated weights) */
  int tf32;                 /* 0: xhat / wt are bf16 (tcgen05 kind::f16); 1: both are fp32 and the GEMM runs kind::tf32 - the
                               fp32-fidelity mode: 10-bit-mantissa operands like the reference's cuDNN TF32 convolutions
                               (torch default), fp32 accumulation, half the bf16 tensor rate */
} vcn_conv_desc;

int vcn_modulate_pack(const void* x, int x_dtype, const int64_t xs[4], const float* styles, void* xhat,
                      int n, int cin, int cin_pad, int h, int w, void* stream);
int vcn_pack_weights(const void* w, int w_dtype, int cout, int cin, int kh, int kw, int transpose_io,
                     void* wt, int cols_pad, float* wsq, void* stream);
/* _ex variants: the GEMM operand dtype is explicit - VCN_BF16 (what the plain entry points produce) or VCN_F32 (operands
 * of the kind::tf32 GEMM, desc.tf32 = 1). */
int vcn_modulate_pack_ex(const void* x, int x_dtype, const int64_t xs[4], const float* styles, void* xhat, int xhat_dtype,
                         int n, int cin, int cin_pad, int h, int w, void* stream);
int vcn_pack_weights_ex(const void* w, int w_dtype, int cout, int cin, int kh, int kw, int transpose_io, void* wt, int wt_dtype,
                        int cols_pad, float* wsq, void* stream);
int vcn_demod_coefs(const float* styles, const float* wsq, float* d, int n, int cin, int cout, void* stream);
/* Per-sample modulated (and, with dcoef, demodulated) weights of the reference's fused branch (networks.py:54-60):
 * wn bf16 [n][kh*kw][cout][cin_pad] = W[co,ci,tap] * styles[n,ci] * dcoef[n,co]; consumed with desc.w_per_sample = 1. */
int vcn_pack_weights_mod(const void* w, int w_dtype, const float* styles, const float* dcoef, void* wn,
                         int n, int cout, int cin, int cin_pad, int kh, int kw, void* stream);
/* epi (may be NULL) reuses vcn_upfirdn2d_epilogue: acc * scale[n,co] + noise[n?,oy,ox], then bias_act. */
/* Leave `sms` SMs out of vcn_conv_igemm's persistent grid (default 0).  For data-parallel runs whose gradient all-reduce
 * overlaps the convolutions: the CTAs of a collective kernel own their SMs, and a persistent CTA that cannot become resident
 * would hold its statically assigned tiles back until the collective ends.  Process-wide; no reference counterpart (the
 * reference has no overlapped collective, stylegan2_ada/training/training_loop.py:300-307). */
int vcn_set_conv_sm_reserve(int sms);
int vcn_conv_igemm(const void* xhat, const void* wt, void* y, const vcn_conv_desc* desc,
                   const vcn_upfirdn2d_epilogue* epi, void* stream);

/* Backward of the (fused) layer  y = bias_act(conv(x*s, W) * d + noise, b)  w.r.t. x and s (the reference gets these
 * from autograd over networks.py:63-72 + bias_act.py:160-175; closed form verified in SURVEY App. B):
 *   vcn_modconv_bwd_pack    dz = bias_act'(dy; y) (epi->act: 0 none, linear, or lrelu), gsum[n,c] += sum_p dz*(y0*d),
 *                           db[c] += sum_p dz, dy0 = dz * d[n,c] -> bf16 NHWC [n,h,w,c_pad] (operand of the DGRAD /
 *                           SCONV2 GEMM).  epi->scale = d (NULL: no demodulation), epi->noise/bias as in forward.
 *                           gsum and db accumulate with atomics: zero them first.
 *   vcn_modconv_bwd_finish  dx = dxhat * s[n,c] in place (fp32), dsum[n,c] += sum_p dxhat * x
 *   vcn_demod_grad          ds[n,i] += s[n,i] * sum_o (-gsum[n,o] d[n,o]^2) wsq[o,i]
 */
int vcn_modconv_bwd_pack(const void* dy, const void* y, int dtype, const int64_t dys[4], const int64_t ys[4],
                         const vcn_upfirdn2d_epilogue* epi, void* dy0, float* gsum, float* db,
                         int n, int c, int c_pad, int h, int w, void* stream);
int vcn_modconv_bwd_pack_ex(const void* dy, const void* y, int dtype, const int64_t dys[4], const int64_t ys[4],
                            const vcn_upfirdn2d_epilogue* epi, void* dy0, int dy0_dtype, float* gsum, float* db,
                            int n, int c, int c_pad, int h, int w, void* stream);
int vcn_modconv_bwd_finish(float* dxhat, const int64_t ds[4], const void* x, int x_dtype, const int64_t xs[4],
                           const float* styles, float* dsum, int n, int c, int h, int w, void* stream);
/* Resident 16-bit channels-last pipeline: dxhat is a dense bf16 [n,h,w,c_pad] buffer (the DGRAD / SCONV2 GEMM output),
 * x the layer input (bf16, channel stride 1, strides xs); dx = dxhat * styles[n,c] in place, dsum[n,c] += sum_p dxhat*x.
 * vcn_modconv_bwd_pack takes the same route on its own when dy / y are 16-bit with channel stride 1. */
int vcn_modconv_bwd_finish_cl(void* dxhat, const void* x, int dtype, const int64_t xs[4], const float* styles,
                              float* dsum, int n, int c, int c_pad, int h, int w, void* stream);
/* Weight gradient of the convolution under modulated_conv2d (the ATen/cuDNN wgrad reached from conv2d_gradfix.py:38,43):
 *   kind 0: dw[co][ci][ky][kx] += sum_{n,p} dy0[n,p,co] * xhat[n, p + (ky-kh/2, kx-kw/2), ci]          (VCN_CONV_FPROP)
 *   kind 1: dw[co][ci][ky][kx] += sum_{n,i,j} dy0[n, 2i+ky, 2j+kx, co] * xhat[n,i,j,ci]               (VCN_CONV_TCONV2)
 * dy0 bf16 NHWC [n,dy_h,dy_w,cout_pad], xhat bf16 NHWC [n,h,w,cin_pad] (the operands of the forward / input-gradient
 * GEMMs, as they are); dw fp32 [cout,cin,kh,kw], ACCUMULATED (zero it first).  tcgen05 GEMM with both operands MN-major,
 * split over pixels across CTAs. */
/* Fused backward of a ToRGB layer (networks.py:307-323: 1x1 modulated convolution to cout <= 4 channels, no demodulation) in
 * the resident bf16 channels-last pipeline: dxhat[n,p,c] = sum_o wgt[o,c] * dy0[n,p,o] (dy0 = the [n,h,w,8] operand rows of
 * vcn_modconv_bwd_pack), dx = dxhat * styles[n,c], dsum[n,c] += sum_p dxhat * x - one pass instead of vcn_conv_igemm +
 * vcn_modconv_bwd_finish_cl.  x, dx: dense bf16 [n,h,w,c], c a power of two >= 64; wgt fp32 [cout,c]; dsum may be NULL;
 * acc (may be NULL): a second gradient of the same activation, dense bf16 [n,h,w,c], added to dx in the same pass (the skip
 * architecture reads x twice, networks.py:411-423). */
int vcn_torgb_bwd_fused(const void* x, const void* dy0, const float* wgt, const float* styles, const void* acc, void* dx,
                        float* dsum, int dtype, int n, int c, int h, int w, int cout, void* stream);
int vcn_conv_wgrad(const void* dy0, const void* xhat, float* dw, int kind, int n, int h, int w, int cin_pad, int cin,
                   int cout_pad, int cout, int kh, int kw, int dy_h, int dy_w, void* stream);
int vcn_demod_grad(float* ds, const float* styles, const float* gsum, const float* dcoef, const float* wsq,
                   int n, int cin, int cout, void* stream);

/* ---- LPIPS feature tap (the reconstruction-loss reductions next to the hot path) -----------------------
 * One tap of lpips v0.1.4 (`normalize_tensor` -> squared difference -> 1x1 `lin` -> spatial mean), as called from
 * viscoin/training/losses.py:113-123:
 *   val[n] += mean_p sum_c lin[c] * (a[n,p,c]/(|a[n,p,:]|+1e-10) - b[n,p,c]/(|b[n,p,:]|+1e-10))^2
 * fa, fb: [n,c,h,w] feature maps with channel stride 1 (channels_last), f32 / f16 / bf16, strides as/bs in elements;
 * c a multiple of 8, c/8 a power of two or a multiple of 32, c <= 512.  val must be zeroed by the caller.
 * bwd: da (dense channels-last [n,h,w,c], dtype of fa) = gval[n] * d val[n] / d fa. */
int vcn_lpips_tap_fwd(const void* fa, const void* fb, int dtype, const int64_t as[4], const int64_t bs[4],
                      const float* lin, float* val, int n, int c, int h, int w, void* stream);
int vcn_lpips_tap_bwd(const void* fa, const void* fb, int dtype, const int64_t as[4], const int64_t bs[4],
                      const float* lin, const float* gval, void* da, int n, int c, int h, int w, void* stream);

/* ---- frozen-network passes (LPIPS VGG-16 and the classifier f around the hot path) ------------------------
 * 16-bit (bf16 / f16) dense channels-last tensors [n,h,w,c], c a multiple of 8, even h and w.  These replace ATen's
 * max_pool2d / max_pool2d_with_indices_backward / add / threshold_backward passes of
 * viscoin/training/losses.py:111-123 (lpips VGG-16) and of the ResNet-50 behind viscoin/models/classifiers.py:83-92:
 *   vcn_maxpool2x2_fwd       y[n,h/2,w/2,c] = max of each 2x2 window (stride 2)
 *   vcn_relu_mask_bwd        dz = (g + g2) * (y > 0); g2 may be NULL; flat arrays of numel elements
 *   vcn_maxpool2x2_relu_bwd  dz[n,h,w,c] = (gtap + U(gpool)) * (y > 0), U = max-pool backward (gradient to the first
 *                            maximum of the window in row-major scan order); gtap may be NULL */
int vcn_maxpool2x2_fwd(const void* x, void* y, int dtype, int n, int h, int w, int c, void* stream);
/* 3x3 / stride 2 / padding 1 max-pool of the ResNet stem (HF ResNetEmbeddings.pooler = torch.nn.MaxPool2d(3, 2, 1); the
 * reference reaches it through transformers' ResNetModel, viscoin/models/classifiers.py:83-92): x [n,h,w,c] -> y [n,oh,ow,c],
 * oh = (h-1)/2+1.  idx (may be NULL when no backward follows) = one byte per output element: the winning window position
 * dy*3+dx, first maximum in scan order (ATen's max_pool2d_with_indices).  _bwd: gx [n,h,w,c] from gy [n,oh,ow,c] and idx. */
/* Fused elementwise passes over the [n,3,h,w] fp32 images of the iteration (csrc/image_ops.cu).
 *   vcn_image_prep_fwd  y[n,h,w,3] bf16 = (x - shift[c]) * inv_scale[c]  (NULL shift / inv_scale = 0 / 1): the input conversion of
 *                       the frozen f and LPIPS's ScalingLayer (lpips v0.1.4; viscoin/training/losses.py:111-123)
 *   vcn_image_prep_bwd  gx[n,3,h,w] fp32 = gy[n,h,w,3] * inv_scale[c]
 *   vcn_l1l2_fwd        out2[0] += sum |a-b|, out2[1] += sum (a-b)^2  (out2 zeroed by the caller; losses.py:125-157 L1 + L2 terms)
 *   vcn_l1l2_bwd        ga = g2[0] * sign(a-b) + 2 * g2[1] * (a-b), g2 = device pointer to the two upstream scale factors */
int vcn_image_prep_fwd(const float* x, void* y, const float* shift, const float* inv_scale, int n, int h, int w, void* stream);
int vcn_image_prep_bwd(const void* gy, float* gx, const float* inv_scale, int n, int h, int w, void* stream);
int vcn_l1l2_fwd(const float* a, const float* b, float* out2, int64_t numel, void* stream);
int vcn_l1l2_bwd(const float* a, const float* b, const float* g2, float* ga, int64_t numel, void* stream);

/* First convolution of a frozen image network: 3 -> 64 channels, 3x3/stride 1/pad 1 (VGG-16 conv1_1, lpips v0.1.4 through
 * viscoin/training/losses.py:111-123) or 7x7/stride 2/pad 3 (the ResNet stem, viscoin/models/classifiers.py:83-92), bias
 * (+ ReLU), bf16 channels-last: x [n,h,w,3], y [n,oh,ow,64].  wp = the weight re-laid out as [KP][64] bf16 with
 * k = (ky*ks + kx)*3 + c and KP = ks*ks*3 rounded up to 16 (zero rows); bias fp32 [64] or NULL.  Replaces the cuDNN call the
 * reference reaches through torch (conv2d on a 3-channel image). */
int vcn_conv_first3(const void* x, const void* wp, const float* bias, void* y, int dtype, int ks, int stride, int pad, int relu,
                    int n, int h, int w, void* stream);
int vcn_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int dtype, int n, int h, int w, int c, void* stream);
int vcn_maxpool3x3s2_bwd(const void* gy, const void* idx, void* gx, int dtype, int n, int h, int w, int c, void* stream);
int vcn_relu_mask_bwd(const void* y, const void* g, const void* g2, void* dz, int dtype, int64_t numel, void* stream);
int vcn_maxpool2x2_relu_bwd(const void* y, const void* gpool, const void* gtap, void* dz, int dtype, int n, int h, int w, int c,
                            void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VCN_B200_H_ */
