// act_math.cuh - per-element arithmetic of bias_act (forward, first and second derivative), shared by the
// bias_act kernels and by the fused epilogues of upfirdn2d / modulated conv.
// Semantics follow the reference table: stylegan2_ada/torch_utils/ops/bias_act.py:23-33 (specs) and
// bias_act.cu:51-142 (formulae, exp-range guards, strict clamp inequalities).
#pragma once
#include "common.cuh"

namespace vcn {

template <typename S> struct ActConst;
template <> struct ActConst<float> {
  static __device__ __forceinline__ float e(float v) { return expf(v); }
  static __device__ __forceinline__ float l(float v) { return logf(v); }
};
template <> struct ActConst<double> {
  static __device__ __forceinline__ double e(double v) { return exp(v); }
  static __device__ __forceinline__ double l(double v) { return log(v); }
};

struct ActParams {
  int act;      // vcn_act
  int grad;     // 0, 1, 2
  float alpha, gain, clamp;
};

// One element.  `x` is the streamed operand (input for grad 0, incoming gradient for grad 1/2) with the bias already
// added where the contract says so; `xref`/`yref` are the saved forward tensors (xref with bias added); `dy` = 1 when
// absent.  ACT is a compile-time activation id, or 0 for "dispatch on p.act at run time" (uniform branch).
template <typename S, int ACT>
__device__ __forceinline__ S act_eval(S x, S xref, S yref, S dy, int act_rt, int grad, S alpha, S gain, S clamp) {
  const int act = ACT ? ACT : act_rt;
  const S one = S(1), two = S(2), zero = S(0);
  const S kRange = S(80), kHalfRange = S(40);
  const S kSeluScale = S(1.0507009873554804934193349852946);
  const S kSeluAlpha = S(1.6732632423543772848170429916717);
  const S yy = (gain != zero) ? yref / gain : zero;
  S y = zero;
  switch (act) {
    case VCN_ACT_LINEAR:
      y = (grad <= 1) ? x : zero;
      break;
    case VCN_ACT_RELU:
      if (grad == 0) y = x > zero ? x : zero;
      else if (grad == 1) y = yy > zero ? x : zero;
      break;
    case VCN_ACT_LRELU:
      if (grad == 0) y = x > zero ? x : x * alpha;
      else if (grad == 1) y = yy > zero ? x : x * alpha;
      break;
    case VCN_ACT_TANH:
      if (grad == 0) {
        const S c = ActConst<S>::e(x), d = one / c;
        y = x < -kRange ? -one : (x > kRange ? one : (c - d) / (c + d));
      } else {
        y = x * (one - yy * yy);
        if (grad == 2) y *= -two * yy;
      }
      break;
    case VCN_ACT_SIGMOID:
      if (grad == 0) y = x < -kRange ? zero : one / (ActConst<S>::e(-x) + one);
      else {
        y = x * yy * (one - yy);
        if (grad == 2) y *= one - two * yy;
      }
      break;
    case VCN_ACT_ELU:
      if (grad == 0) y = x >= zero ? x : ActConst<S>::e(x) - one;
      else if (grad == 1) y = yy >= zero ? x : x * (yy + one);
      else y = yy >= zero ? zero : x * (yy + one);
      break;
    case VCN_ACT_SELU:
      if (grad == 0) y = x >= zero ? kSeluScale * x : (kSeluScale * kSeluAlpha) * (ActConst<S>::e(x) - one);
      else if (grad == 1) y = yy >= zero ? x * kSeluScale : x * (yy + kSeluScale * kSeluAlpha);
      else y = yy >= zero ? zero : x * (yy + kSeluScale * kSeluAlpha);
      break;
    case VCN_ACT_SOFTPLUS:
      if (grad == 0) y = x > kRange ? x : ActConst<S>::l(ActConst<S>::e(x) + one);
      else {
        const S c = ActConst<S>::e(-yy);
        y = grad == 1 ? x * (one - c) : x * c * (one - c);
      }
      break;
    case VCN_ACT_SWISH:
      if (grad == 0) y = x < -kRange ? zero : x / (ActConst<S>::e(-x) + one);
      else {
        const S c = ActConst<S>::e(xref), d = c + one;
        if (grad == 1) y = xref > kHalfRange ? x : x * c * (xref + d) / (d * d);
        else y = xref > kHalfRange ? zero : x * c * (xref * (two - d) + two * d) / (d * d * d);
        yref = xref < -kRange ? zero : xref / (ActConst<S>::e(-xref) + one) * gain;  // clamp mask uses recomputed y
      }
      break;
    default: break;
  }
  y *= gain * dy;
  if (clamp >= zero) {
    if (grad == 0) y = (y > -clamp && y < clamp) ? y : (y >= zero ? clamp : -clamp);
    else y = (yref > -clamp && yref < clamp) ? y : zero;
  }
  return y;
}

}  // namespace vcn
