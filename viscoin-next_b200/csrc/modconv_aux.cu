// modconv_aux.cu - operand preparation for the tcgen05 implicit GEMM of modconv.cu (all memory-bound).
//   modulate_pack : x[n,ci,h,w] (any strides, f32/f16/bf16) * styles[n,ci] -> xhat bf16 NHWC [n,h,w,cin_pad]
//                   (networks.py:64 `x = x * styles.reshape(N,-1,1,1)` fused with the layout/dtype change the tensor
//                   cores need; a 64-channel x 32-pixel tile is transposed through shared memory so that both the
//                   NCHW reads and the NHWC writes are full 128-byte lines)
//   pack_weights  : W[cout,cin,kh,kw] -> wt bf16 [tap][rows][cols_pad] (+ wsq[cout,cin] = sum_k W^2)
//   demod_coefs   : d[n,co] = rsqrt(sum_ci s^2 wsq + 1e-8)  (networks.py:55-58 without materialising w*s)
#include <string.h>

#include "common.cuh"

namespace vcn {
namespace {

// GEMM operand element: bf16 (tcgen05 kind::f16) or fp32 (kind::tf32: the tensor core reads the upper 19 bits)
template <typename TO> __device__ __forceinline__ TO to_op(float v);
template <> __device__ __forceinline__ __nv_bfloat16 to_op<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ float to_op<float>(float v) { return v; }

template <typename TO> __device__ __forceinline__ void store_pair(TO* p, float a, float b);
template <> __device__ __forceinline__ void store_pair<__nv_bfloat16>(__nv_bfloat16* p, float a, float b) {
  *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}
template <> __device__ __forceinline__ void store_pair<float>(float* p, float a, float b) {
  *reinterpret_cast<float2*>(p) = make_float2(a, b);
}

template <typename TO> __device__ __forceinline__ void store8(TO* p, const float (&v)[8]);
template <> __device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[8]) {
  Vec16<__nv_bfloat16> o;
#pragma unroll
  for (int k = 0; k < 8; ++k) o.v[k] = __float2bfloat16_rn(v[k]);
  *reinterpret_cast<Vec16<__nv_bfloat16>*>(p) = o;
}
template <> __device__ __forceinline__ void store8<float>(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

template <typename T, typename TO>
__global__ void __launch_bounds__(256) modulate_pack_kernel(const T* __restrict__ x, int64_t xs_n, int64_t xs_c,
                                                            int64_t xs_h, int64_t xs_w, const float* __restrict__ styles,
                                                            TO* __restrict__ xhat, int cin, int cin_pad, int h,
                                                            int w) {
  // grid: (ceil(h*w/32), ceil(cin_pad/64), n); tile = 64 channels x 32 pixels
  __shared__ float tile[64][33];
  const int n = blockIdx.z, c0 = blockIdx.y * 64, p0 = blockIdx.x * 32;
  const int hw = h * w;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* xn = x + (int64_t)n * xs_n;
  const int p = p0 + lane;
  const int py = p / w, px = p - py * w;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = c0 + warp * 8 + i;
    float v = 0.f;
    if (c < cin && p < hw) {
      v = to_acc<T>(xn[(int64_t)c * xs_c + (int64_t)py * xs_h + (int64_t)px * xs_w]);
      if (styles) v *= styles[(int64_t)n * cin + c];
    }
    tile[warp * 8 + i][lane] = v;
  }
  __syncthreads();
  // each warp writes 4 pixels; a lane owns channels (2*lane, 2*lane+1): 128 contiguous bytes per pixel
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int pp = warp * 4 + i;
    const int gp = p0 + pp;
    const int c = c0 + 2 * lane;
    if (gp < hw && c < cin_pad) store_pair<TO>(xhat + ((int64_t)n * hw + gp) * cin_pad + c, tile[2 * lane][pp], tile[2 * lane + 1][pp]);
  }
}

// channels-fastest input (xs_c == 1): elementwise pass, 8 channels (16 bytes of bf16 output) per thread
template <typename T, typename TO>
__global__ void __launch_bounds__(256) modulate_pack_cl_kernel(const T* __restrict__ x, int64_t xs_n, int64_t xs_h,
                                                               int64_t xs_w, const float* __restrict__ styles,
                                                               TO* __restrict__ xhat, int n_img, int cin,
                                                               int cin_pad, int h, int w, int vec_in) {
  const int groups = cin_pad / 8;
  const int64_t total = (int64_t)n_img * h * w * groups;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % groups) * 8;
    int64_t r = i / groups;
    const int px = (int)(r % w); r /= w;
    const int py = (int)(r % h);
    const int n = (int)(r / h);
    const T* src = x + n * xs_n + py * xs_h + px * xs_w + c;
    float v[8];
    if (vec_in && c + 8 <= cin) {
      if constexpr (sizeof(T) == 2) {
        const Vec16<T> in = *reinterpret_cast<const Vec16<T>*>(src);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = to_acc<T>(in.v[k]);
      } else {
#pragma unroll
        for (int h4 = 0; h4 < 2; ++h4) {
          const Vec16<T> in = *reinterpret_cast<const Vec16<T>*>(src + h4 * 4);
#pragma unroll
          for (int k = 0; k < 4; ++k) v[h4 * 4 + k] = to_acc<T>(in.v[k]);
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = (c + k < cin) ? to_acc<T>(src[k]) : 0.f;
    }
    if (styles) {
      const float* sp = styles + (int64_t)n * cin + c;
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] *= (c + k < cin) ? sp[k] : 0.f;
    }
    store8<TO>(xhat + (((int64_t)n * h + py) * w + px) * cin_pad + c, v);
  }
}

// Dense rows (stride_h == w * stride_w), 16-bit input, cin == cin_pad a power of two <= 2048, vector-aligned: a thread owns
// ONE channel group for the whole launch (its 8 style factors stay in registers), walks pixels with a stride of
// 256 / groups, four 16-byte loads in flight - no 64-bit divisions and no per-vector style loads (the generic kernel above
// spends ~3 int64 div/mod and 8 scalar loads per 16-byte vector).  grid = (pixel chunks, n).
template <typename T>
__global__ void __launch_bounds__(256) modulate_pack_cl_rows_kernel(const T* __restrict__ x, int64_t xs_n, int64_t xs_w,
                                                                    const float* __restrict__ styles,
                                                                    __nv_bfloat16* __restrict__ xhat, int cin, int hw, int chunk) {
  const int groups = cin >> 3, rows = 256 / groups;
  const int g = threadIdx.x % groups, row = threadIdx.x / groups;
  const int n = blockIdx.y, c = g * 8;
  float sc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) sc[k] = styles ? styles[(int64_t)n * cin + c + k] : 1.f;
  const T* src = x + n * xs_n + c;
  __nv_bfloat16* dst = xhat + (int64_t)n * hw * cin + c;
  const int p_end = min(hw, (int)(blockIdx.x + 1) * chunk);
  constexpr int kBatch = 4;
  for (int p = blockIdx.x * chunk + row; p < p_end; p += kBatch * rows) {
    Vec16<T> in[kBatch];
#pragma unroll
    for (int b = 0; b < kBatch; ++b)
      if (p + b * rows < p_end) in[b] = ld_stream<T>(src + (int64_t)(p + b * rows) * xs_w);
#pragma unroll
    for (int b = 0; b < kBatch; ++b) {
      if (p + b * rows < p_end) {
        Vec16<__nv_bfloat16> o;
#pragma unroll
        for (int k = 0; k < 8; ++k) o.v[k] = __float2bfloat16_rn(to_acc<T>(in[b].v[k]) * sc[k]);
        *reinterpret_cast<Vec16<__nv_bfloat16>*>(dst + (int64_t)(p + b * rows) * cin) = o;
      }
    }
  }
}

template <typename T, typename TO>
__global__ void __launch_bounds__(256) pack_weights_kernel(const T* __restrict__ w, int cout, int cin, int taps,
                                                           int transpose_io, TO* __restrict__ wt, int rows,
                                                           int cols, int cols_pad) {
  // one thread per (tap, row, col_pad) element of wt
  const int64_t total = (int64_t)taps * rows * cols_pad;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int col = (int)(i % cols_pad);
    int64_t r = i / cols_pad;
    const int row = (int)(r % rows);
    const int tap = (int)(r / rows);
    float v = 0.f;
    if (col < cols) {
      const int co = transpose_io ? col : row, ci = transpose_io ? row : col;
      v = to_acc<T>(w[((int64_t)co * cin + ci) * taps + tap]);
    }
    wt[i] = to_op<TO>(v);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) weight_sq_kernel(const T* __restrict__ w, int64_t pairs, int taps,
                                                        float* __restrict__ wsq) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= pairs) return;
  float acc = 0.f;
  for (int t = 0; t < taps; ++t) {
    const float v = to_acc<T>(w[i * taps + t]);
    acc = fmaf(v, v, acc);
  }
  wsq[i] = acc;
}

// one warp per (n, co): d = rsqrt(sum_ci s[n,ci]^2 * wsq[co,ci] + 1e-8)
__global__ void __launch_bounds__(256) demod_coefs_kernel(const float* __restrict__ styles, const float* __restrict__ wsq,
                                                          float* __restrict__ d, int n, int cin, int cout) {
  const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (gw >= n * cout) return;
  const int in = gw / cout, co = gw - in * cout;
  const float* s = styles + (int64_t)in * cin;
  const float* q = wsq + (int64_t)co * cin;
  float acc = 0.f;
  for (int c = lane; c < cin; c += 32) acc = fmaf(s[c] * s[c], q[c], acc);
  acc = warp_sum(acc);
  if (lane == 0) d[gw] = rsqrtf(acc + 1e-8f);
}


// ---------------------------------------------------------------------------------------------------------
// backward helpers
// ---------------------------------------------------------------------------------------------------------
struct BwdPackArgs {
  const void* dy; const void* y;
  int64_t dys[4], ys[4];
  const float* noise; int64_t noise_stride_n;
  const float* bias; const float* dcoef;
  void* dy0; float* gsum; float* db;
  int n, c, c_pad, h, w;
  int act; float alpha, gain, clamp;
  int subs;  // 32-pixel sub-tiles per CTA
};

// dz = bias_act'(dy; y), r = act^-1(y) - bias - noise (= y0 * d), gsum[n,c] += sum_p dz*r, db[c] += sum_p dz,
// dy0 = dz * d[n,c] -> bf16 NHWC.  Tile = 64 channels x 128 pixels; lanes run along pixels for the NCHW reads and
// along channels for the NHWC writes (transposed through shared memory).
template <typename T, typename TO>
__global__ void __launch_bounds__(256) modconv_bwd_pack_kernel(const BwdPackArgs a) {
  __shared__ float tile[64][33];
  const int n = blockIdx.z, c0 = blockIdx.y * 64;
  const int hw = a.h * a.w;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* dyn = static_cast<const T*>(a.dy) + (int64_t)n * a.dys[0];
  const T* yn = a.y ? static_cast<const T*>(a.y) + (int64_t)n * a.ys[0] : nullptr;
  float gacc[8], bacc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) gacc[i] = bacc[i] = 0.f;
  for (int sub = 0; sub < a.subs; ++sub) {
    const int p0 = (blockIdx.x * a.subs + sub) * 32;
    if (p0 >= hw) break;  // block-uniform
    const int p = p0 + lane;
    const int py = p / a.w, px = p - py * a.w;
    float nz = 0.f;
    if (a.noise && p < hw) nz = a.noise[(int64_t)n * a.noise_stride_n + p];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = c0 + warp * 8 + i;
      float out = 0.f;
      if (c < a.c && p < hw) {
        const float dyv = to_acc<T>(dyn[(int64_t)c * a.dys[1] + (int64_t)py * a.dys[2] + (int64_t)px * a.dys[3]]);
        const float yv = yn ? to_acc<T>(yn[(int64_t)c * a.ys[1] + (int64_t)py * a.ys[2] + (int64_t)px * a.ys[3]]) : 0.f;
        float dz = dyv, r = yv;
        if (a.act) {
          // first derivative of bias_act keyed on the stored output (bias_act.cu:48,72,141)
          const float yy = yv / a.gain;
          dz = (a.act == VCN_ACT_LRELU && !(yy > 0.f)) ? dyv * a.alpha : dyv;
          dz *= a.gain;
          if (a.clamp >= 0.f && !(yv > -a.clamp && yv < a.clamp)) dz = 0.f;
          r = (a.act == VCN_ACT_LRELU && !(yy > 0.f)) ? yy / a.alpha : yy;
          if (a.bias) r -= a.bias[c];
        }
        r -= nz;
        gacc[i] = fmaf(dz, r, gacc[i]);
        bacc[i] += dz;
        out = a.dcoef ? dz * a.dcoef[(int64_t)n * a.c + c] : dz;
      }
      tile[warp * 8 + i][lane] = out;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int pp = warp * 4 + i;
      const int gp = p0 + pp;
      const int c = c0 + 2 * lane;
      if (gp < hw && c < a.c_pad)
        store_pair<TO>(static_cast<TO*>(a.dy0) + ((int64_t)n * hw + gp) * a.c_pad + c, tile[2 * lane][pp], tile[2 * lane + 1][pp]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = c0 + warp * 8 + i;
    const float g = warp_sum(gacc[i]), b = warp_sum(bacc[i]);
    if (lane == 0 && c < a.c) {
      if (a.gsum) atomicAdd(a.gsum + (int64_t)n * a.c + c, g);
      if (a.db) atomicAdd(a.db + c, b);
    }
  }
}

// Few channels (c <= 8, c_pad == 8: the ToRGB layers, whose dy is the fp32 NCHW image gradient): the 64-channel tile above
// keeps one warp in eight busy (230 us for [16,3,256,256]).  Here a thread owns a PIXEL: the per-channel reads are
// coalesced along the pixels of a plane, the 8-channel NHWC operand row is one 16-byte (bf16) or two 16-byte (fp32) stores,
// and the channel sums stay in registers until one block reduction.  grid (pixel chunks, n).
template <typename T, typename TO>
__global__ void __launch_bounds__(256) modconv_bwd_pack_smallc_kernel(const BwdPackArgs a, const int chunk) {
  __shared__ float red[8][16];
  const int n = blockIdx.y;
  const int hw = a.h * a.w;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* dyn = static_cast<const T*>(a.dy) + (int64_t)n * a.dys[0];
  const T* yn = a.y ? static_cast<const T*>(a.y) + (int64_t)n * a.ys[0] : nullptr;
  float gacc[8], bacc[8], dc[8], bi[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    gacc[i] = bacc[i] = 0.f;
    dc[i] = (a.dcoef && i < a.c) ? a.dcoef[(int64_t)n * a.c + i] : 1.f;
    bi[i] = (a.bias && a.act && i < a.c) ? a.bias[i] : 0.f;
  }
  const int p_end = min(hw, (int)(blockIdx.x + 1) * chunk);
  const bool dense_a = a.dys[3] == 1 && a.dys[2] == a.w, dense_y = a.ys[3] == 1 && a.ys[2] == a.w;
  for (int p = blockIdx.x * chunk + threadIdx.x; p < p_end; p += 256) {
    const int py = p / a.w, px = p - py * a.w;
    const int64_t oa = dense_a ? (int64_t)p : (int64_t)py * a.dys[2] + (int64_t)px * a.dys[3];
    const int64_t oy = dense_y ? (int64_t)p : (int64_t)py * a.ys[2] + (int64_t)px * a.ys[3];
    const float nz = a.noise ? a.noise[(int64_t)n * a.noise_stride_n + p] : 0.f;
    float out[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      out[i] = 0.f;
      if (i < a.c) {
        const float dyv = to_acc<T>(dyn[(int64_t)i * a.dys[1] + oa]);
        const float yv = yn ? to_acc<T>(yn[(int64_t)i * a.ys[1] + oy]) : 0.f;
        float dz = dyv, r = yv;
        if (a.act) {  // same arithmetic as modconv_bwd_pack_kernel
          const float yy = yv / a.gain;
          dz = (a.act == VCN_ACT_LRELU && !(yy > 0.f)) ? dyv * a.alpha : dyv;
          dz *= a.gain;
          if (a.clamp >= 0.f && !(yv > -a.clamp && yv < a.clamp)) dz = 0.f;
          r = (a.act == VCN_ACT_LRELU && !(yy > 0.f)) ? yy / a.alpha : yy;
          r -= bi[i];
        }
        r -= nz;
        gacc[i] = fmaf(dz, r, gacc[i]);
        bacc[i] += dz;
        out[i] = dz * dc[i];
      }
    }
    TO* dst = static_cast<TO*>(a.dy0) + ((int64_t)n * hw + p) * 8;
    if constexpr (sizeof(TO) == 2) {
      Vec16<TO> o;
#pragma unroll
      for (int i = 0; i < 8; ++i) o.v[i] = from_acc<TO>(out[i]);
      *reinterpret_cast<Vec16<TO>*>(dst) = o;
    } else {
      *reinterpret_cast<float4*>(dst) = make_float4(out[0], out[1], out[2], out[3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(out[4], out[5], out[6], out[7]);
    }
  }
  if (!a.gsum && !a.db) return;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float g = warp_sum(gacc[i]), b = warp_sum(bacc[i]);
    if (lane == 0) { red[warp][i] = g; red[warp][8 + i] = b; }
  }
  __syncthreads();
  if (threadIdx.x < 16) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
    const int i = threadIdx.x & 7;
    if (i < a.c) {
      if (threadIdx.x < 8) { if (a.gsum) atomicAdd(a.gsum + (int64_t)n * a.c + i, t); }
      else if (a.db) atomicAdd(a.db + i, t);
    }
  }
}

// dx = dxhat * s[n,c] (in place), ds1[n,c] += sum_p dxhat * x.  grid (chunks, c, n); dxhat fp32, any strides.
template <typename T>
__global__ void __launch_bounds__(256) modconv_bwd_finish_kernel(float* __restrict__ dxhat, const int64_t ds0, const int64_t ds1_,
                                                                 const int64_t ds2, const int64_t ds3,
                                                                 const T* __restrict__ x, const int64_t xs0, const int64_t xs1,
                                                                 const int64_t xs2, const int64_t xs3,
                                                                 const float* __restrict__ styles, float* __restrict__ dsum,
                                                                 int c_total, int h, int w, int chunk) {
  const int n = blockIdx.z, c = blockIdx.y;
  const int hw = h * w;
  const int p_begin = blockIdx.x * chunk;
  const int p_end = min(hw, p_begin + chunk);
  const float s = styles ? styles[(int64_t)n * c_total + c] : 1.f;
  float* dp = dxhat + n * ds0 + c * ds1_;
  const T* xp = x ? x + n * xs0 + c * xs1 : nullptr;
  float acc = 0.f;
  for (int p = p_begin + threadIdx.x; p < p_end; p += blockDim.x) {
    const int py = p / w, px = p - py * w;
    float* d = dp + py * ds2 + px * ds3;
    const float v = *d;
    if (xp) acc = fmaf(v, to_acc<T>(xp[py * xs2 + px * xs3]), acc);
    *d = v * s;
  }
  if (dsum) {
    __shared__ float red[8];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < 8; ++i) t += red[i];
      atomicAdd(dsum + (int64_t)n * c_total + c, t);
    }
  }
}

// ds[n,i] += s[n,i] * sum_o (-gsum[n,o] * d[n,o]^2) * wsq[o,i]
// block = 32 input channels x 8 slices of the output-channel sum (wsq rows are read coalesced), reduced through smem.
__global__ void __launch_bounds__(256) demod_grad_kernel(float* __restrict__ ds, const float* __restrict__ styles,
                                                         const float* __restrict__ gsum, const float* __restrict__ dcoef,
                                                         const float* __restrict__ wsq, int n, int cin, int cout) {
  __shared__ float red[8][33];
  const int lane = threadIdx.x & 31, slice = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane, in = blockIdx.y;
  float acc = 0.f;
  if (i < cin) {
    // eight independent (dcoef, gsum, wsq) load triples in flight per thread: the dependent one-at-a-time loop was pure
    // latency (64 round trips = 25 us for 4 M multiply-adds)
    for (int o0 = slice; o0 < cout; o0 += 64) {
      float d[8], g[8], w[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int o = o0 + 8 * u;
        const bool ok = o < cout;
        d[u] = ok ? dcoef[(int64_t)in * cout + o] : 0.f;
        g[u] = ok ? gsum[(int64_t)in * cout + o] : 0.f;
        w[u] = ok ? wsq[(int64_t)o * cin + i] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) acc = fmaf(-g[u] * d[u] * d[u], w[u], acc);
    }
  }
  red[slice][lane] = acc;
  __syncthreads();
  if (slice == 0 && i < cin) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][lane];
    ds[(int64_t)in * cin + i] += styles[(int64_t)in * cin + i] * t;
  }
}


// Per-sample modulated weights (the reference's fused branch, networks.py:54-60): wn[n][tap][co][ci_pad] =
// bf16( W[co,ci,tap] * s[n,ci] * d[n,co] ).  One CTA = one output channel x 128 input channels: the fp32 taps are read
// once (coalesced) into shared memory and reused for every sample.
template <typename T>
__global__ void __launch_bounds__(128) pack_weights_mod_kernel(const T* __restrict__ w, const float* __restrict__ styles,
                                                               const float* __restrict__ dcoef, __nv_bfloat16* __restrict__ wn,
                                                               int n, int cout, int cin, int cin_pad, int taps) {
  extern __shared__ float sw[];  // [128][taps]
  const int co = blockIdx.y, ci0 = blockIdx.x * 128;
  const int cnt = min(128, cin - ci0);
  const T* src = w + ((int64_t)co * cin + ci0) * taps;
  for (int i = threadIdx.x; i < cnt * taps; i += 128) sw[i] = to_acc<T>(src[i]);
  __syncthreads();
  const int ci = ci0 + threadIdx.x;
  if (ci >= cin_pad) return;
  for (int in = 0; in < n; ++in) {
    float m = 0.f;
    if (ci < cin) {
      m = styles[(int64_t)in * cin + ci];
      if (dcoef) m *= dcoef[(int64_t)in * cout + co];
    }
    for (int t = 0; t < taps; ++t) {
      const float v = ci < cin ? sw[threadIdx.x * taps + t] * m : 0.f;
      wn[(((int64_t)in * taps + t) * cout + co) * cin_pad + ci] = __float2bfloat16_rn(v);
    }
  }
}

template <typename T, typename TO>
int run_modulate(const void* x, const int64_t xs[4], const float* styles, void* xhat, int n, int cin, int cin_pad, int h,
                 int w, cudaStream_t st) {
  const T* px = static_cast<const T*>(x);
  TO* po = static_cast<TO*>(xhat);
  if (xs[1] == 1 && cin > 1) {
    const int64_t total = (int64_t)n * h * w * (cin_pad / 8);
    int64_t grid = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (grid > cap) grid = cap;
    const int64_t vq = 16 / (int64_t)sizeof(T);
    const int vec_in = aligned16(x) && xs[0] % vq == 0 && xs[2] % vq == 0 && xs[3] % vq == 0;
    if constexpr (sizeof(T) == 2 && sizeof(TO) == 2) {
      const int groups = cin_pad / 8;
      const bool pow2 = (groups & (groups - 1)) == 0 && groups <= 256;
      if (vec_in && cin == cin_pad && pow2 && xs[2] == (int64_t)w * xs[3] && (int64_t)h * w <= 0x7fffffff) {
        const int hw = h * w, rows = 256 / groups;
        int chunk = hw;  // about 16 CTAs per SM in total, at least 16 pixel iterations each
        while (chunk > 16 * rows && (int64_t)n * ((hw + chunk - 1) / chunk) < (int64_t)sm_count() * 16) chunk = (chunk + 1) / 2;
        dim3 g2((unsigned)((hw + chunk - 1) / chunk), (unsigned)n);
        modulate_pack_cl_rows_kernel<T><<<g2, 256, 0, st>>>(px, xs[0], xs[3], styles, po, cin, hw, chunk);
        VCN_LAUNCH_CHECK("modulate_pack_cl_rows_kernel");
        return VCN_OK;
      }
    }
    modulate_pack_cl_kernel<T, TO><<<(unsigned)grid, 256, 0, st>>>(px, xs[0], xs[2], xs[3], styles, po, n, cin, cin_pad, h, w, vec_in);
    VCN_LAUNCH_CHECK("modulate_pack_cl_kernel");
    return VCN_OK;
  }
  dim3 grid((unsigned)((h * w + 31) / 32), (unsigned)((cin_pad + 63) / 64), (unsigned)n);
  modulate_pack_kernel<T, TO><<<grid, 256, 0, st>>>(px, xs[0], xs[1], xs[2], xs[3], styles, po, cin, cin_pad, h, w);
  VCN_LAUNCH_CHECK("modulate_pack_kernel");
  return VCN_OK;
}

template <typename T, typename TO>
int run_pack(const void* w, int cout, int cin, int taps, int transpose_io, void* wt, int cols_pad, float* wsq,
             cudaStream_t st) {
  const int rows = transpose_io ? cin : cout, cols = transpose_io ? cout : cin;
  if (wt) {
    const int64_t total = (int64_t)taps * rows * cols_pad;
    int64_t grid = (total + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (grid > cap) grid = cap;
    pack_weights_kernel<T, TO><<<(unsigned)grid, 256, 0, st>>>(static_cast<const T*>(w), cout, cin, taps, transpose_io,
                                                              static_cast<TO*>(wt), rows, cols, cols_pad);
    VCN_LAUNCH_CHECK("pack_weights_kernel");
  }
  if (wsq) {
    const int64_t pairs = (int64_t)cout * cin;
    weight_sq_kernel<T><<<(unsigned)((pairs + 255) / 256), 256, 0, st>>>(static_cast<const T*>(w), pairs, taps, wsq);
    VCN_LAUNCH_CHECK("weight_sq_kernel");
  }
  return VCN_OK;
}

}  // namespace
}  // namespace vcn

namespace vcn {
int bwd_pack_cl_try(const void* dy, const void* y, int dtype, const int64_t dys[4], const int64_t ys[4],
                    const vcn_upfirdn2d_epilogue* epi, void* dy0, float* gsum, float* db, int n, int c, int c_pad, int h,
                    int w, int act, float alpha, float gain, float clamp, cudaStream_t st);  // modconv_bwd_cl.cu
}

extern "C" {

int vcn_modulate_pack(const void* x, int x_dtype, const int64_t xs[4], const float* styles, void* xhat, int n, int cin,
                      int cin_pad, int h, int w, void* stream) {
  return vcn_modulate_pack_ex(x, x_dtype, xs, styles, xhat, VCN_BF16, n, cin, cin_pad, h, w, stream);
}

int vcn_modulate_pack_ex(const void* x, int x_dtype, const int64_t xs[4], const float* styles, void* xhat, int xhat_dtype,
                         int n, int cin, int cin_pad, int h, int w, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(xhat_dtype == VCN_BF16 || xhat_dtype == VCN_F32, "modulate_pack: operand dtype must be bf16 or f32 (tf32), got %d", xhat_dtype);
  VCN_REQUIRE(n >= 0 && cin >= 1 && h >= 1 && w >= 1, "modulate_pack: bad shape [%d,%d,%d,%d]", n, cin, h, w);
  VCN_REQUIRE(cin_pad >= cin && cin_pad % 8 == 0, "modulate_pack: cin_pad (%d) must be a multiple of 8 and >= cin (%d)",
              cin_pad, cin);
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(x && xhat && xs, "modulate_pack: x, xs and xhat must not be NULL");
  VCN_REQUIRE((int64_t)n * h * w * cin_pad <= 0x7fffffffLL * 4, "modulate_pack: tensor too large");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (xhat_dtype == VCN_F32) {
    switch (x_dtype) {
      case VCN_F32: return run_modulate<float, float>(x, xs, styles, xhat, n, cin, cin_pad, h, w, st);
      case VCN_F16: return run_modulate<__half, float>(x, xs, styles, xhat, n, cin, cin_pad, h, w, st);
      case VCN_BF16: return run_modulate<__nv_bfloat16, float>(x, xs, styles, xhat, n, cin, cin_pad, h, w, st);
    }
  } else {
    switch (x_dtype) {
      case VCN_F32: return run_modulate<float, __nv_bfloat16>(x, xs, styles, xhat, n, cin, cin_pad, h, w, st);
      case VCN_F16: return run_modulate<__half, __nv_bfloat16>(x, xs, styles, xhat, n, cin, cin_pad, h, w, st);
      case VCN_BF16: return run_modulate<__nv_bfloat16, __nv_bfloat16>(x, xs, styles, xhat, n, cin, cin_pad, h, w, st);
    }
  }
  set_error("modulate_pack: unsupported dtype %d", x_dtype);
  return VCN_ERR_INVALID;
}

int vcn_pack_weights(const void* w, int w_dtype, int cout, int cin, int kh, int kw, int transpose_io, void* wt,
                     int cols_pad, float* wsq, void* stream) {
  return vcn_pack_weights_ex(w, w_dtype, cout, cin, kh, kw, transpose_io, wt, VCN_BF16, cols_pad, wsq, stream);
}

int vcn_pack_weights_ex(const void* w, int w_dtype, int cout, int cin, int kh, int kw, int transpose_io, void* wt, int wt_dtype,
                        int cols_pad, float* wsq, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(wt_dtype == VCN_BF16 || wt_dtype == VCN_F32, "pack_weights: operand dtype must be bf16 or f32 (tf32), got %d", wt_dtype);
  VCN_REQUIRE(w != nullptr, "pack_weights: w must not be NULL");
  VCN_REQUIRE(cout >= 1 && cin >= 1 && kh >= 1 && kw >= 1, "pack_weights: bad shape [%d,%d,%d,%d]", cout, cin, kh, kw);
  const int cols = transpose_io ? cout : cin;
  if (wt) VCN_REQUIRE(cols_pad >= cols && cols_pad % 8 == 0, "pack_weights: cols_pad (%d) must be a multiple of 8 >= %d", cols_pad, cols);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (wt_dtype == VCN_F32) {
    switch (w_dtype) {
      case VCN_F32: return run_pack<float, float>(w, cout, cin, kh * kw, transpose_io, wt, cols_pad, wsq, st);
      case VCN_F16: return run_pack<__half, float>(w, cout, cin, kh * kw, transpose_io, wt, cols_pad, wsq, st);
      case VCN_BF16: return run_pack<__nv_bfloat16, float>(w, cout, cin, kh * kw, transpose_io, wt, cols_pad, wsq, st);
    }
  } else {
    switch (w_dtype) {
      case VCN_F32: return run_pack<float, __nv_bfloat16>(w, cout, cin, kh * kw, transpose_io, wt, cols_pad, wsq, st);
      case VCN_F16: return run_pack<__half, __nv_bfloat16>(w, cout, cin, kh * kw, transpose_io, wt, cols_pad, wsq, st);
      case VCN_BF16: return run_pack<__nv_bfloat16, __nv_bfloat16>(w, cout, cin, kh * kw, transpose_io, wt, cols_pad, wsq, st);
    }
  }
  set_error("pack_weights: unsupported dtype %d", w_dtype);
  return VCN_ERR_INVALID;
}

int vcn_demod_coefs(const float* styles, const float* wsq, float* d, int n, int cin, int cout, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(n >= 0 && cin >= 1 && cout >= 1, "demod_coefs: bad shape");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(styles && wsq && d, "demod_coefs: NULL pointer");
  const int64_t threads = (int64_t)n * cout * 32;
  demod_coefs_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(styles, wsq, d, n, cin, cout);
  VCN_LAUNCH_CHECK("demod_coefs_kernel");
  return VCN_OK;
}


int vcn_modconv_bwd_pack(const void* dy, const void* y, int dtype, const int64_t dys[4], const int64_t ys[4],
                         const vcn_upfirdn2d_epilogue* epi, void* dy0, float* gsum, float* db, int n, int c, int c_pad,
                         int h, int w, void* stream) {
  return vcn_modconv_bwd_pack_ex(dy, y, dtype, dys, ys, epi, dy0, VCN_BF16, gsum, db, n, c, c_pad, h, w, stream);
}

int vcn_modconv_bwd_pack_ex(const void* dy, const void* y, int dtype, const int64_t dys[4], const int64_t ys[4],
                            const vcn_upfirdn2d_epilogue* epi, void* dy0, int dy0_dtype, float* gsum, float* db, int n, int c,
                            int c_pad, int h, int w, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dy0_dtype == VCN_BF16 || dy0_dtype == VCN_F32, "modconv_bwd_pack: operand dtype must be bf16 or f32 (tf32), got %d", dy0_dtype);
  VCN_REQUIRE(n >= 0 && c >= 1 && h >= 1 && w >= 1, "modconv_bwd_pack: bad shape");
  VCN_REQUIRE(c_pad >= c && c_pad % 8 == 0, "modconv_bwd_pack: c_pad must be a multiple of 8 >= c");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(dy && dy0 && dys, "modconv_bwd_pack: dy, dys, dy0 must not be NULL");
  BwdPackArgs a;
  memset(&a, 0, sizeof(a));
  a.dy = dy; a.y = y; a.dy0 = dy0; a.gsum = gsum; a.db = db;
  for (int i = 0; i < 4; ++i) { a.dys[i] = dys[i]; a.ys[i] = ys ? ys[i] : dys[i]; }
  a.n = n; a.c = c; a.c_pad = c_pad; a.h = h; a.w = w;
  a.gain = 1.f; a.clamp = -1.f;
  if (epi) {
    a.noise = epi->noise; a.noise_stride_n = epi->noise_stride_n; a.bias = epi->bias; a.dcoef = epi->scale;
    a.act = epi->act; a.alpha = epi->alpha; a.gain = epi->act_gain; a.clamp = epi->clamp;
  }
  VCN_REQUIRE(a.act == 0 || a.act == VCN_ACT_LINEAR || (a.act == VCN_ACT_LRELU && a.alpha != 0.f),
              "modconv_bwd_pack: fused backward supports no activation, linear, or lrelu with alpha != 0 (got act %d)", a.act);
  VCN_REQUIRE(a.act == 0 || a.gain != 0.f, "modconv_bwd_pack: gain must be non-zero");
  VCN_REQUIRE((a.act == 0 && !gsum) || y, "modconv_bwd_pack: y is required with an activation or when gsum is requested");
  // 128 pixels per CTA; few channels (ToRGB: 3): every CTA ends in one atomicAdd per channel on the SAME c addresses of db
  // (8192 CTAs on 3 addresses took 180 us at 256^2) - longer pixel runs per CTA while the machine stays covered
  a.subs = 4;
  while (a.subs < 64 && c <= 16 && (int64_t)n * ((h * w + 32 * a.subs - 1) / (32 * a.subs)) > (int64_t)sm_count() * 4) a.subs *= 2;
  dim3 grid((unsigned)((h * w + 32 * a.subs - 1) / (32 * a.subs)), (unsigned)((c_pad + 63) / 64), (unsigned)n);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dy0_dtype == VCN_BF16) {  // 16-bit channels-last operands: streaming kernel of modconv_bwd_cl.cu
    const int rc = bwd_pack_cl_try(dy, y, dtype, a.dys, a.ys, epi, dy0, gsum, db, n, c, c_pad, h, w, a.act, a.alpha, a.gain,
                                   a.clamp, st);
    if (rc >= 0) return rc;
  }
  static const bool allow_smallc = [] { const char* e = getenv("VCN_BWD_PACK_SMALLC"); return !e || atoi(e) != 0; }();
  if (allow_smallc && c_pad == 8 && (reinterpret_cast<uintptr_t>(dy0) & 15u) == 0) {
    // few channels (ToRGB): one thread per pixel; about 4 CTAs per SM, at least 1024 pixels each
    const int hw = h * w;
    int chunk = hw;
    while (chunk > 1024 && (int64_t)n * ((hw + chunk - 1) / chunk) < (int64_t)sm_count() * 4) chunk = (chunk + 1) / 2;
    chunk = (chunk + 255) / 256 * 256;
    dim3 sgrid((unsigned)((hw + chunk - 1) / chunk), (unsigned)n);
    bool done = true;
    if (dy0_dtype == VCN_F32) {
      switch (dtype) {
        case VCN_F32: modconv_bwd_pack_smallc_kernel<float, float><<<sgrid, 256, 0, st>>>(a, chunk); break;
        case VCN_F16: modconv_bwd_pack_smallc_kernel<__half, float><<<sgrid, 256, 0, st>>>(a, chunk); break;
        case VCN_BF16: modconv_bwd_pack_smallc_kernel<__nv_bfloat16, float><<<sgrid, 256, 0, st>>>(a, chunk); break;
        default: done = false;
      }
    } else {
      switch (dtype) {
        case VCN_F32: modconv_bwd_pack_smallc_kernel<float, __nv_bfloat16><<<sgrid, 256, 0, st>>>(a, chunk); break;
        case VCN_F16: modconv_bwd_pack_smallc_kernel<__half, __nv_bfloat16><<<sgrid, 256, 0, st>>>(a, chunk); break;
        case VCN_BF16: modconv_bwd_pack_smallc_kernel<__nv_bfloat16, __nv_bfloat16><<<sgrid, 256, 0, st>>>(a, chunk); break;
        default: done = false;
      }
    }
    if (done) {
      VCN_LAUNCH_CHECK("modconv_bwd_pack_smallc_kernel");
      return VCN_OK;
    }
  }
  if (dy0_dtype == VCN_F32) {
    switch (dtype) {
      case VCN_F32: modconv_bwd_pack_kernel<float, float><<<grid, 256, 0, st>>>(a); break;
      case VCN_F16: modconv_bwd_pack_kernel<__half, float><<<grid, 256, 0, st>>>(a); break;
      case VCN_BF16: modconv_bwd_pack_kernel<__nv_bfloat16, float><<<grid, 256, 0, st>>>(a); break;
      default: set_error("modconv_bwd_pack: unsupported dtype %d", dtype); return VCN_ERR_INVALID;
    }
  } else {
    switch (dtype) {
      case VCN_F32: modconv_bwd_pack_kernel<float, __nv_bfloat16><<<grid, 256, 0, st>>>(a); break;
      case VCN_F16: modconv_bwd_pack_kernel<__half, __nv_bfloat16><<<grid, 256, 0, st>>>(a); break;
      case VCN_BF16: modconv_bwd_pack_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, 256, 0, st>>>(a); break;
      default: set_error("modconv_bwd_pack: unsupported dtype %d", dtype); return VCN_ERR_INVALID;
    }
  }
  VCN_LAUNCH_CHECK("modconv_bwd_pack_kernel");
  return VCN_OK;
}

int vcn_modconv_bwd_finish(float* dxhat, const int64_t ds[4], const void* x, int x_dtype, const int64_t xs[4],
                           const float* styles, float* dsum, int n, int c, int h, int w, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(n >= 0 && c >= 1 && h >= 1 && w >= 1, "modconv_bwd_finish: bad shape");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(dxhat && ds, "modconv_bwd_finish: dxhat must not be NULL");
  VCN_REQUIRE(!dsum || (x && xs), "modconv_bwd_finish: dsum needs x");
  const int hw = h * w;
  int chunk = hw;
  // enough CTAs to fill the machine, at least 2048 pixels each
  while (chunk > 4096 && (int64_t)n * c * ((hw + chunk - 1) / chunk) < (int64_t)sm_count() * 8) chunk = (chunk + 1) / 2;
  dim3 grid((unsigned)((hw + chunk - 1) / chunk), (unsigned)c, (unsigned)n);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t z[4] = {0, 0, 0, 0};
  if (!xs) xs = z;
  switch (x ? x_dtype : VCN_F32) {
    case VCN_F32: modconv_bwd_finish_kernel<float><<<grid, 256, 0, st>>>(dxhat, ds[0], ds[1], ds[2], ds[3], static_cast<const float*>(x), xs[0], xs[1], xs[2], xs[3], styles, dsum, c, h, w, chunk); break;
    case VCN_F16: modconv_bwd_finish_kernel<__half><<<grid, 256, 0, st>>>(dxhat, ds[0], ds[1], ds[2], ds[3], static_cast<const __half*>(x), xs[0], xs[1], xs[2], xs[3], styles, dsum, c, h, w, chunk); break;
    case VCN_BF16: modconv_bwd_finish_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(dxhat, ds[0], ds[1], ds[2], ds[3], static_cast<const __nv_bfloat16*>(x), xs[0], xs[1], xs[2], xs[3], styles, dsum, c, h, w, chunk); break;
    default: set_error("modconv_bwd_finish: unsupported dtype %d", x_dtype); return VCN_ERR_INVALID;
  }
  VCN_LAUNCH_CHECK("modconv_bwd_finish_kernel");
  return VCN_OK;
}

int vcn_demod_grad(float* ds, const float* styles, const float* gsum, const float* dcoef, const float* wsq, int n, int cin,
                   int cout, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(n >= 0 && cin >= 1 && cout >= 1, "demod_grad: bad shape");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(ds && styles && gsum && dcoef && wsq, "demod_grad: NULL pointer");
  dim3 grid((unsigned)((cin + 31) / 32), (unsigned)n);
  demod_grad_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(ds, styles, gsum, dcoef, wsq, n, cin, cout);
  VCN_LAUNCH_CHECK("demod_grad_kernel");
  return VCN_OK;
}


int vcn_pack_weights_mod(const void* w, int w_dtype, const float* styles, const float* dcoef, void* wn, int n, int cout,
                         int cin, int cin_pad, int kh, int kw, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(n >= 0 && cout >= 1 && cin >= 1 && kh >= 1 && kw >= 1, "pack_weights_mod: bad shape");
  VCN_REQUIRE(cin_pad >= cin && cin_pad % 8 == 0, "pack_weights_mod: cin_pad must be a multiple of 8 >= cin");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(w && styles && wn, "pack_weights_mod: NULL pointer");
  const int taps = kh * kw;
  dim3 grid((unsigned)((cin_pad + 127) / 128), (unsigned)cout);
  const size_t smem = (size_t)128 * taps * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (w_dtype) {
    case VCN_F32: pack_weights_mod_kernel<float><<<grid, 128, smem, st>>>(static_cast<const float*>(w), styles, dcoef, static_cast<__nv_bfloat16*>(wn), n, cout, cin, cin_pad, taps); break;
    case VCN_F16: pack_weights_mod_kernel<__half><<<grid, 128, smem, st>>>(static_cast<const __half*>(w), styles, dcoef, static_cast<__nv_bfloat16*>(wn), n, cout, cin, cin_pad, taps); break;
    case VCN_BF16: pack_weights_mod_kernel<__nv_bfloat16><<<grid, 128, smem, st>>>(static_cast<const __nv_bfloat16*>(w), styles, dcoef, static_cast<__nv_bfloat16*>(wn), n, cout, cin, cin_pad, taps); break;
    default: set_error("pack_weights_mod: unsupported dtype %d", w_dtype); return VCN_ERR_INVALID;
  }
  VCN_LAUNCH_CHECK("pack_weights_mod_kernel");
  return VCN_OK;
}

}  // extern "C"
