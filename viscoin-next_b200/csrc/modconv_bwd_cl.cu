// modconv_bwd_cl.cu - backward helpers of the fused synthesis layer for RESIDENT 16-bit channels-last activations
// (the training-mode pipeline: activations stay bf16 NHWC between layers, forward and backward).
//
//   bwd_pack_cl   : dz = bias_act'(dy; y),  gsum[n,c] += sum_p dz * (y0*d),  db[c] += sum_p dz,  dy0 = dz * d[n,c]
//                   dy, y : 16-bit, channels fastest (stride 1 on c);  dy0 : bf16 NHWC [n,h,w,c_pad]
//                   (autograd of networks.py:69-72 fma + :302-304 bias_act, bias_act.py:160-175)
//   bwd_finish_cl : dx = dxhat * s[n,c] in place (bf16 NHWC), dsum[n,c] += sum_p dxhat * x   (autograd of networks.py:64)
//
// Both are one streaming pass: a thread owns 8 consecutive channels (one 16-byte vector) and walks a chunk of pixels;
// threads of a CTA that share a channel group combine their partial sums through shared memory, one atomicAdd per
// (CTA, channel).  HBM-bound: 6 bytes per element each (two 16-bit reads + one 16-bit write).
#include <stdlib.h>
#include <string.h>

#include <type_traits>

#include "common.cuh"

namespace vcn {
namespace {

struct BwdClArgs {
  const void* dy; const void* y;       // pack: incoming gradient / stored output.  finish: dy = dxhat (in/out), y = x
  int64_t as[4], bs[4];                // element strides (n, c, h, w) of dy and of y; c-stride is 1
  const float* noise; int64_t noise_stride_n;
  const float* bias; const float* scale;  // pack: scale = d[n,c].  finish: scale = styles[n,c]
  __nv_bfloat16* out;                  // pack: dy0 [n,h,w,c_pad]
  float* sum_nc; float* sum_c;         // pack: gsum, db.  finish: dsum, unused
  int n, c, c_pad, h, w, chunk;
  int act; float alpha, gain, clamp, inv_gain, inv_alpha;
  // homogeneous form (act none / linear / lrelu with gain > 0, alpha > 0): with pos = (y > 0)
  //   dz = dy * (pos ? g : ag),  y0*d = y * (pos ? cr_pos : cr_neg) - (b + noise)
  int fast_act; float g, ag, cr_pos, cr_neg;
};

__device__ __forceinline__ bool aligned16_dev(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

template <typename T> __device__ __forceinline__ void load8(const T* p, bool vec, int valid, float (&v)[8]) {
  if (vec) {
    const Vec16<T> in = *reinterpret_cast<const Vec16<T>*>(p);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = to_acc<T>(in.v[k]);
  } else {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = k < valid ? to_acc<T>(p[k]) : 0.f;
  }
}

// Reduce per-thread partials over the threads that share a channel group and add them to dst[c0 + k].
__device__ __forceinline__ void group_reduce_add(float (*red)[8], const float (&acc)[8], int tid, int groups, int rows,
                                                 float* dst, int c_valid) {
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 8; ++k) red[tid][k] = acc[k];
  __syncthreads();
  for (int ch = tid; ch < groups * 8; ch += 256) {  // up to 2048 channels on 256 threads
    const int g = ch >> 3, k = ch & 7;
    float t = 0.f;
    for (int j = 0; j < rows; ++j) t += red[j * groups + g][k];
    if (ch < c_valid) atomicAdd(dst + ch, t);
  }
}

// Occupancy is what these passes run on (ncu: long_scoreboard stalls at 16 warps per SM; the same loop without
// accumulators - modulate_pack_cl_rows_kernel, 8 CTAs per SM - streams at the HBM peak): small batches, 3-4 CTAs per SM.
template <typename T, bool FINISH, bool FAST>
__global__ void __launch_bounds__(256, FINISH ? 4 : (FAST ? 3 : 2)) modconv_bwd_cl_kernel(const BwdClArgs a) {
  __shared__ float red[256][8];
  const int tid = threadIdx.x, n = blockIdx.y;
  const int groups = a.c_pad >> 3;        // <= 256
  const int rows = 256 / groups;          // pixels in flight per CTA iteration
  const int g = tid % groups, row = tid / groups;
  const bool active = row < rows;
  const int c = g * 8;
  const int valid = min(8, a.c - c);      // channels of this group that exist (<= 0: padding group)
  const int hw = a.h * a.w;
  const int p_end = min(hw, (int)(blockIdx.x + 1) * a.chunk);
  float sc[8], bi[8], acc0[8], acc1[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const bool ok = k < valid;
    sc[k] = (a.scale && ok) ? a.scale[(int64_t)n * a.c + c + k] : (FINISH ? 0.f : 1.f);
    bi[k] = (!FINISH && a.bias && a.act && ok) ? a.bias[c + k] : 0.f;
    acc0[k] = acc1[k] = 0.f;
  }
  if (FINISH && !a.scale) {
#pragma unroll
    for (int k = 0; k < 8; ++k) sc[k] = 1.f;
  }
  const T* pa = static_cast<const T*>(a.dy) + (int64_t)n * a.as[0] + c;
  const T* pb = a.y ? static_cast<const T*>(a.y) + (int64_t)n * a.bs[0] + c : nullptr;
  const bool va = valid == 8 && ((a.as[0] | a.as[2] | a.as[3]) & 7) == 0 && aligned16_dev(a.dy);
  const bool vb = valid == 8 && ((a.bs[0] | a.bs[2] | a.bs[3]) & 7) == 0 && aligned16_dev(a.y);
  // one pixel of this thread's 8 channels: u = dy (or dxhat), v = y (or x); off_a = element offset of the pixel in dy
  auto pixel = [&](const float (&u)[8], const float (&v)[8], float nz, int64_t off_a, int p) {
    Vec16<__nv_bfloat16> o;
    if (FINISH) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        acc0[k] = fmaf(u[k], v[k], acc0[k]);
        o.v[k] = __float2bfloat16_rn(u[k] * sc[k]);
      }
      // in place: dxhat is a dense [n,h,w,c_pad] bf16 buffer
      *reinterpret_cast<Vec16<__nv_bfloat16>*>(const_cast<T*>(pa) + off_a) = o;
    } else if (FAST) {
      // 9 operations per element instead of ~25 (the generic form below made the pass issue-bound: ncu 35 instructions
      // per element at 16 warps per SM)
      const bool has_clamp = a.clamp >= 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const bool pos = v[k] > 0.f;
        float dz = u[k] * (pos ? a.g : a.ag);
        if (has_clamp) dz = fabsf(v[k]) < a.clamp ? dz : 0.f;
        const float r = fmaf(v[k], pos ? a.cr_pos : a.cr_neg, -(bi[k] + nz));
        acc0[k] = fmaf(dz, r, acc0[k]);
        acc1[k] += dz;
        o.v[k] = __float2bfloat16_rn(dz * sc[k]);
      }
      *reinterpret_cast<Vec16<__nv_bfloat16>*>(a.out + ((int64_t)n * hw + p) * a.c_pad + c) = o;
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float dz = u[k], r = v[k];
        if (a.act) {  // first derivative of bias_act keyed on the stored output (bias_act.cu:48,72,141)
          const float yy = v[k] * a.inv_gain;
          const bool neg = a.act == VCN_ACT_LRELU && !(yy > 0.f);
          dz = (neg ? u[k] * a.alpha : u[k]) * a.gain;
          if (a.clamp >= 0.f && !(v[k] > -a.clamp && v[k] < a.clamp)) dz = 0.f;
          r = (neg ? yy * a.inv_alpha : yy) - bi[k];
        }
        r -= nz;
        acc0[k] = fmaf(dz, r, acc0[k]);
        acc1[k] += dz;
        o.v[k] = __float2bfloat16_rn(k < valid ? dz * sc[k] : 0.f);
      }
      *reinterpret_cast<Vec16<__nv_bfloat16>*>(a.out + ((int64_t)n * hw + p) * a.c_pad + c) = o;
    }
  };
  const float* noise_n = (!FINISH && a.noise) ? a.noise + (int64_t)n * a.noise_stride_n : nullptr;
  // dense rows (stride_h == w * stride_w) of full, aligned channel groups: the pixel offset is p * stride_w and the
  // loads of kBatch pixels are issued before the first one is consumed (the pass is latency-bound otherwise: at 77
  // registers three CTAs per SM keep only 24 KB in flight, HBM needs ~50 KB per SM)
  constexpr int kBatch = 2;
  const bool fast = va && (!pb || vb) && a.as[2] == (int64_t)a.w * a.as[3] && (!pb || a.bs[2] == (int64_t)a.w * a.bs[3]);
  if (active && fast) {
    const int64_t sa = a.as[3], sb = a.bs[3];
    for (int p = blockIdx.x * a.chunk + row; p < p_end; p += kBatch * rows) {
      Vec16<T> ru[kBatch], rv[kBatch];
      float nz[kBatch];
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        const int pp = p + b * rows;
        if (pp < p_end) {
          // finish rewrites dxhat in place: a plain load there, the non-coherent streaming path for read-only data
          ru[b] = FINISH ? *reinterpret_cast<const Vec16<T>*>(pa + pp * sa) : ld_stream<T>(pa + pp * sa);
          if (pb) rv[b] = ld_stream<T>(pb + pp * sb);
          nz[b] = noise_n ? __ldg(noise_n + pp) : 0.f;
        }
      }
#pragma unroll
      for (int b = 0; b < kBatch; ++b) {
        const int pp = p + b * rows;
        if (pp < p_end) {
          float u[8], v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) { u[k] = to_acc<T>(ru[b].v[k]); v[k] = pb ? to_acc<T>(rv[b].v[k]) : 0.f; }
          pixel(u, v, nz[b], pp * sa, pp);
        }
      }
    }
  } else if (active) {
    for (int p = blockIdx.x * a.chunk + row; p < p_end; p += rows) {
      const int py = p / a.w, px = p - py * a.w;
      float u[8], v[8];
      load8<T>(pa + py * a.as[2] + px * a.as[3], va, valid, u);
      if (pb) load8<T>(pb + py * a.bs[2] + px * a.bs[3], vb, valid, v);
      else {
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = 0.f;
      }
      pixel(u, v, noise_n ? noise_n[p] : 0.f, py * a.as[2] + px * a.as[3], p);
    }
  }
  if (!active) {
#pragma unroll
    for (int k = 0; k < 8; ++k) acc0[k] = acc1[k] = 0.f;
  }
  if (a.sum_nc) group_reduce_add(red, acc0, tid, groups, rows, a.sum_nc + (int64_t)n * a.c, a.c);
  if (!FINISH && a.sum_c) group_reduce_add(red, acc1, tid, groups, rows, a.sum_c, a.c);
}

// ---- bulk-copy staged variant (dense NHWC, full channel groups, homogeneous activation) ----------------------------------
// The register-staged kernel above keeps 3 CTAs x 256 threads x 2 x 2 16-byte loads = 49 KB in flight per SM - about what HBM3e
// needs at its loaded latency, hence 0.55 (pack) / 0.73 (finish) of the copy peak.  Here the TMA engine streams both input
// tensors into a shared-memory ring (cp.async.bulk, 16 KB per stage, kStagesBulk stages per CTA, 3 CTAs per SM = 192 KB in
// flight per SM, no registers spent on it); the threads read their 16-byte vectors from the ring, do the same arithmetic and
// write the result straight to global memory.  A stage = 2 * rows consecutive pixels of one sample (contiguous in NHWC).
constexpr int kStagesBulk = 4;
constexpr uint32_t kStageTensorBytes = 8192;  // per tensor and stage: 2 pixels per thread x 256 threads x 16 bytes

template <typename T, bool FINISH>
__global__ void __launch_bounds__(256, 3) modconv_bwd_cl_bulk_kernel(const BwdClArgs a) {
  extern __shared__ __align__(128) unsigned char ring[];  // [kStagesBulk][2][kStageTensorBytes]
  __shared__ float red[256][8];
  __shared__ __align__(8) uint64_t full_bar[kStagesBulk];
  const int tid = threadIdx.x, n = blockIdx.y;
  const int groups = a.c_pad >> 3;  // power of two <= 256 (host-checked)
  const int rows = 256 / groups;
  const int g = tid % groups, row = tid / groups;
  const int c = g * 8;
  const int hw = a.h * a.w;
  const int p0 = blockIdx.x * a.chunk, p_end = min(hw, p0 + a.chunk);
  const int sp = 2 * rows;  // pixels per stage
  const int nit = (p_end - p0 + sp - 1) / sp;
  const uint32_t row_bytes = (uint32_t)a.c_pad * 2u;
  const bool two = a.y != nullptr;
  const unsigned char* ga = static_cast<const unsigned char*>(a.dy) + ((int64_t)n * hw + p0) * row_bytes;
  const unsigned char* gb = two ? static_cast<const unsigned char*>(a.y) + ((int64_t)n * hw + p0) * row_bytes : nullptr;

  auto fill = [&](int it) {  // one elected thread: arm the barrier, start the copies of stage `it`
    const int s = it % kStagesBulk;
    const int px = min(sp, p_end - p0 - it * sp);
    const uint32_t bytes = (uint32_t)px * row_bytes;
    unsigned char* dst = ring + (size_t)s * 2 * kStageTensorBytes;
    mbar_expect_tx(full_bar + s, two ? 2u * bytes : bytes);
    bulk_load_1d(dst, ga + (int64_t)it * sp * row_bytes, bytes, full_bar + s);
    if (two) bulk_load_1d(dst + kStageTensorBytes, gb + (int64_t)it * sp * row_bytes, bytes, full_bar + s);
  };
  if (tid == 0) {
    for (int s = 0; s < kStagesBulk; ++s) mbar_init(full_bar + s, 1);
    fence_mbar_init();
    for (int it = 0; it < min(nit, kStagesBulk); ++it) fill(it);
  }
  float sc[8], bi[8], acc0[8], acc1[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    sc[k] = a.scale ? a.scale[(int64_t)n * a.c + c + k] : 1.f;
    bi[k] = (!FINISH && a.bias && a.act) ? a.bias[c + k] : 0.f;
    acc0[k] = acc1[k] = 0.f;
  }
  const float* noise_n = (!FINISH && a.noise) ? a.noise + (int64_t)n * a.noise_stride_n : nullptr;
  const bool has_clamp = !FINISH && a.clamp >= 0.f;
  __syncthreads();  // barrier inits visible
  for (int it = 0; it < nit; ++it) {
    const int s = it % kStagesBulk;
    mbar_wait(full_bar + s, (uint32_t)(it / kStagesBulk) & 1u);
    const unsigned char* sa = ring + (size_t)s * 2 * kStageTensorBytes;
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int pl = row + b * rows;         // pixel inside the stage
      const int p = p0 + it * sp + pl;       // pixel inside the sample
      if (p < p_end) {
        const Vec16<T> ru = *reinterpret_cast<const Vec16<T>*>(sa + (size_t)pl * row_bytes + g * 16);
        Vec16<T> rv;
        if (two) rv = *reinterpret_cast<const Vec16<T>*>(sa + kStageTensorBytes + (size_t)pl * row_bytes + g * 16);
        Vec16<__nv_bfloat16> o;
        if (FINISH) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float u = to_acc<T>(ru.v[k]);
            if (two) acc0[k] = fmaf(u, to_acc<T>(rv.v[k]), acc0[k]);
            o.v[k] = __float2bfloat16_rn(u * sc[k]);
          }
          // in place: this pixel of dxhat has already been copied into the ring
          *reinterpret_cast<Vec16<__nv_bfloat16>*>(const_cast<unsigned char*>(ga) + ((int64_t)it * sp + pl) * row_bytes + g * 16) = o;
        } else {
          const float nz = noise_n ? __ldg(noise_n + p) : 0.f;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float u = to_acc<T>(ru.v[k]), v = two ? to_acc<T>(rv.v[k]) : 0.f;
            const bool pos = v > 0.f;
            float dz = u * (pos ? a.g : a.ag);
            if (has_clamp) dz = fabsf(v) < a.clamp ? dz : 0.f;
            const float r = fmaf(v, pos ? a.cr_pos : a.cr_neg, -(bi[k] + nz));
            acc0[k] = fmaf(dz, r, acc0[k]);
            acc1[k] += dz;
            o.v[k] = __float2bfloat16_rn(dz * sc[k]);
          }
          st_stream<__nv_bfloat16>(a.out + ((int64_t)n * hw + p) * a.c_pad + c, o);
        }
      }
    }
    __syncthreads();  // every thread has read stage s
    if (tid == 0 && it + kStagesBulk < nit) {
      fence_proxy_async();  // the generic-proxy reads above are ordered before the async-proxy refill
      fill(it + kStagesBulk);
    }
  }
  if (a.sum_nc) group_reduce_add(red, acc0, tid, groups, rows, a.sum_nc + (int64_t)n * a.c, a.c);
  if (!FINISH && a.sum_c) group_reduce_add(red, acc1, tid, groups, rows, a.sum_c, a.c);
}

// ---- fused backward of a ToRGB layer (1x1 modulated convolution to <= 4 channels, no demodulation; networks.py:307-323) ----
//   dxhat[n,p,c] = sum_o W[o,c] * dy0[n,p,o]      (the input-gradient "GEMM" is a rank-<=4 expansion: memory-bound)
//   dx[n,p,c]    = dxhat * s[n,c]                  dsum[n,c] = sum_p dxhat * x[n,p,c]
// One pass over x (read) and dx (write) instead of GEMM (write dxhat) + finish (read dxhat, x; write dx).  Same ring as above:
// a stage = 2 * rows pixels of x (8 KB) plus their 16-byte dy0 rows.
// ACC: dx = acc + dxhat * s - `acc` is the gradient that reaches the same activation through the next block (the skip
// architecture reads x twice: ToRGB and the next block's first convolution); folding it in here replaces a torch add pass.
template <typename T, bool ACC>
__global__ void __launch_bounds__(256, 3) torgb_bwd_bulk_kernel(const T* __restrict__ x, const __nv_bfloat16* __restrict__ dy0,
                                                                const float* __restrict__ wgt, const float* __restrict__ styles,
                                                                const T* __restrict__ acc_in, T* __restrict__ dx,
                                                                float* __restrict__ dsum, int n_, int c, int hw, int cout, int chunk) {
  constexpr int kSt = ACC ? 3 : kStagesBulk;              // three stages with the third stream: 3 CTAs per SM still fit
  extern __shared__ __align__(128) unsigned char ring[];  // [kSt][(1 + ACC) * kStageTensorBytes + 1024]
  __shared__ float red[256][8];
  __shared__ __align__(8) uint64_t full_bar[kStagesBulk];
  constexpr uint32_t kAccOff = kStageTensorBytes + 1024;  // x rows + up to 64 dy0 rows of 16 bytes (+ acc rows)
  constexpr uint32_t kStage = kAccOff + (ACC ? kStageTensorBytes : 0);
  const int tid = threadIdx.x, n = blockIdx.y;
  const int groups = c >> 3, rows = 256 / groups;
  const int g = tid % groups, row = tid / groups;
  const int c0 = g * 8;
  const int p0 = blockIdx.x * chunk, p_end = min(hw, p0 + chunk);
  const int sp = 2 * rows;
  const int nit = (p_end - p0 + sp - 1) / sp;
  const uint32_t row_bytes = (uint32_t)c * 2u;
  const unsigned char* gx = reinterpret_cast<const unsigned char*>(x) + ((int64_t)n * hw + p0) * row_bytes;
  const unsigned char* gd = reinterpret_cast<const unsigned char*>(dy0) + ((int64_t)n * hw + p0) * 16;
  const unsigned char* gacc = ACC ? reinterpret_cast<const unsigned char*>(acc_in) + ((int64_t)n * hw + p0) * row_bytes : nullptr;
  auto fill = [&](int it) {
    const int s = it % kSt;
    const int px = min(sp, p_end - p0 - it * sp);
    unsigned char* dst = ring + (size_t)s * kStage;
    mbar_expect_tx(full_bar + s, (uint32_t)px * ((ACC ? 2u : 1u) * row_bytes + 16u));
    bulk_load_1d(dst, gx + (int64_t)it * sp * row_bytes, (uint32_t)px * row_bytes, full_bar + s);
    bulk_load_1d(dst + kStageTensorBytes, gd + (int64_t)it * sp * 16, (uint32_t)px * 16u, full_bar + s);
    if (ACC) bulk_load_1d(dst + kAccOff, gacc + (int64_t)it * sp * row_bytes, (uint32_t)px * row_bytes, full_bar + s);
  };
  if (tid == 0) {
    for (int s = 0; s < kSt; ++s) mbar_init(full_bar + s, 1);
    fence_mbar_init();
    for (int it = 0; it < min(nit, kSt); ++it) fill(it);
  }
  float wv[4][8], sc[8], acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    sc[k] = styles ? styles[(int64_t)n * c + c0 + k] : 1.f;
    acc[k] = 0.f;
#pragma unroll
    for (int o = 0; o < 4; ++o) wv[o][k] = o < cout ? wgt[(int64_t)o * c + c0 + k] : 0.f;
  }
  __syncthreads();
  for (int it = 0; it < nit; ++it) {
    const int s = it % kSt;
    mbar_wait(full_bar + s, (uint32_t)(it / kSt) & 1u);
    const unsigned char* sa = ring + (size_t)s * kStage;
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int pl = row + b * rows;
      const int p = p0 + it * sp + pl;
      if (p < p_end) {
        const Vec16<T> xv = *reinterpret_cast<const Vec16<T>*>(sa + (size_t)pl * row_bytes + g * 16);
        const uint2 dv = *reinterpret_cast<const uint2*>(sa + kStageTensorBytes + (size_t)pl * 16);  // channels 0..3 of dy0
        const float d0 = __uint_as_float(dv.x << 16), d1 = __uint_as_float(dv.x & 0xffff0000u);
        const float d2 = __uint_as_float(dv.y << 16), d3 = __uint_as_float(dv.y & 0xffff0000u);
        Vec16<T> o, av;
        if (ACC) av = *reinterpret_cast<const Vec16<T>*>(sa + kAccOff + (size_t)pl * row_bytes + g * 16);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float t = fmaf(wv[0][k], d0, fmaf(wv[1][k], d1, fmaf(wv[2][k], d2, wv[3][k] * d3)));
          acc[k] = fmaf(t, to_acc<T>(xv.v[k]), acc[k]);
          o.v[k] = from_acc<T>(ACC ? fmaf(t, sc[k], to_acc<T>(av.v[k])) : t * sc[k]);
        }
        st_stream<T>(dx + ((int64_t)n * hw + p) * c + c0, o);
      }
    }
    __syncthreads();
    if (tid == 0 && it + kSt < nit) {
      fence_proxy_async();
      fill(it + kSt);
    }
  }
  if (dsum) group_reduce_add(red, acc, tid, groups, rows, dsum + (int64_t)n * c, c);
}

// dense [n,h,w,c_pad] 16-bit tensor with c == c_pad, 16-byte aligned
inline bool dense_nhwc(const void* p, const int64_t s[4], int c, int c_pad, int h, int w) {
  return c == c_pad && s[1] == 1 && s[3] == c_pad && s[2] == (int64_t)w * c_pad && s[0] == (int64_t)h * w * c_pad &&
         (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

template <bool FINISH>
int launch_bulk(BwdClArgs& a, int dtype, cudaStream_t st, const char* name) {
  const int hw = a.h * a.w;
  const int rows = 256 / (a.c_pad >> 3);
  const int sp = 2 * rows;
  // one wave of 3 CTAs per SM, every CTA a whole number of stages
  int per_sample = (sm_count() * 3) / a.n;
  if (per_sample < 1) per_sample = 1;
  int chunk = (hw + per_sample - 1) / per_sample;
  chunk = (chunk + sp - 1) / sp * sp;
  if (chunk < 4 * sp) chunk = 4 * sp;
  a.chunk = chunk;
  dim3 grid((unsigned)((hw + chunk - 1) / chunk), (unsigned)a.n);
  const size_t smem = (size_t)kStagesBulk * 2 * kStageTensorBytes;
  static bool attr_set[2][2][64] = {};
  int dev = 0;
  VCN_CUDA(cudaGetDevice(&dev));
  const int ti = dtype == VCN_BF16 ? 0 : 1;
  auto kern = dtype == VCN_BF16 ? modconv_bwd_cl_bulk_kernel<__nv_bfloat16, FINISH> : modconv_bwd_cl_bulk_kernel<__half, FINISH>;
  if (dev < 0 || dev >= 64 || !attr_set[FINISH ? 1 : 0][ti][dev]) {
    VCN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (dev >= 0 && dev < 64) attr_set[FINISH ? 1 : 0][ti][dev] = true;
  }
  kern<<<grid, 256, smem, st>>>(a);
  VCN_LAUNCH_CHECK(name);
  return VCN_OK;
}

template <bool FINISH>
int launch(BwdClArgs& a, int dtype, cudaStream_t st, const char* name) {
  const int hw = a.h * a.w;
  {
    // bulk-copy staged kernel: dense tensors, power-of-two channel count, the homogeneous activation form, enough pixels
    static const bool allow = [] { const char* e = getenv("VCN_BWD_CL_BULK"); return !e || atoi(e) != 0; }();
    const int groups = a.c_pad >> 3;
    const bool pow2 = groups >= 1 && groups <= 256 && (groups & (groups - 1)) == 0;
    if (allow && pow2 && (FINISH || a.fast_act) && dense_nhwc(a.dy, a.as, a.c, a.c_pad, a.h, a.w) &&
        (!a.y || dense_nhwc(a.y, a.bs, a.c, a.c_pad, a.h, a.w)) && (int64_t)a.n * hw * a.c_pad >= (1 << 21))
      return launch_bulk<FINISH>(a, dtype, st, name);
  }
  const int rows = 256 / (a.c_pad >> 3);
  // about 16 CTAs per SM in total, at least 32 pixel iterations each
  int chunk = hw;
  while (chunk > 32 * rows && (int64_t)a.n * ((hw + chunk - 1) / chunk) < (int64_t)sm_count() * 16) chunk = (chunk + 1) / 2;
  // low resolutions (4^2..32^2 at 512 channels): 16-32 CTAs walking 64 dependent iterations each took 25-36 us for a
  // few MB - shorter chunks until every SM has two CTAs
  while (chunk > 8 * rows && (int64_t)a.n * ((hw + chunk - 1) / chunk) < (int64_t)sm_count() * 2) chunk = (chunk + 1) / 2;
  a.chunk = chunk;
  dim3 grid((unsigned)((hw + chunk - 1) / chunk), (unsigned)a.n);
  if (!FINISH && a.fast_act) {
    if (dtype == VCN_BF16) modconv_bwd_cl_kernel<__nv_bfloat16, FINISH, true><<<grid, 256, 0, st>>>(a);
    else modconv_bwd_cl_kernel<__half, FINISH, true><<<grid, 256, 0, st>>>(a);
  } else {
    if (dtype == VCN_BF16) modconv_bwd_cl_kernel<__nv_bfloat16, FINISH, false><<<grid, 256, 0, st>>>(a);
    else modconv_bwd_cl_kernel<__half, FINISH, false><<<grid, 256, 0, st>>>(a);
  }
  VCN_LAUNCH_CHECK(name);
  return VCN_OK;
}

}  // namespace

// Entry used by vcn_modconv_bwd_pack when dy / y are 16-bit with channel stride 1.  Returns -1 when not applicable.
int bwd_pack_cl_try(const void* dy, const void* y, int dtype, const int64_t dys[4], const int64_t ys[4],
                    const vcn_upfirdn2d_epilogue* epi, void* dy0, float* gsum, float* db, int n, int c, int c_pad, int h,
                    int w, int act, float alpha, float gain, float clamp, cudaStream_t st) {
  if (!(dtype == VCN_BF16 || dtype == VCN_F16)) return -1;
  if (dys[1] != 1 || (y && ys[1] != 1) || c_pad > 2048) return -1;
  BwdClArgs a;
  memset(&a, 0, sizeof(a));
  a.dy = dy; a.y = y;
  for (int i = 0; i < 4; ++i) { a.as[i] = dys[i]; a.bs[i] = y ? ys[i] : 0; }
  if (epi) { a.noise = epi->noise; a.noise_stride_n = epi->noise_stride_n; a.bias = epi->bias; a.scale = epi->scale; }
  a.out = static_cast<__nv_bfloat16*>(dy0); a.sum_nc = gsum; a.sum_c = db;
  a.n = n; a.c = c; a.c_pad = c_pad; a.h = h; a.w = w;
  a.act = act; a.alpha = alpha; a.gain = gain; a.clamp = clamp;
  a.inv_gain = gain != 0.f ? 1.f / gain : 0.f; a.inv_alpha = alpha != 0.f ? 1.f / alpha : 0.f;
  static const bool allow_fast = [] { const char* e = getenv("VCN_BWD_CL_FAST"); return !e || atoi(e) != 0; }();
  if (act == 0) {
    a.fast_act = 1; a.g = a.ag = a.cr_pos = a.cr_neg = 1.f; a.clamp = -1.f;
  } else if (act == VCN_ACT_LINEAR && gain > 0.f) {
    a.fast_act = 1; a.g = a.ag = gain; a.cr_pos = a.cr_neg = 1.f / gain;
  } else if (act == VCN_ACT_LRELU && gain > 0.f && alpha > 0.f) {
    a.fast_act = 1; a.g = gain; a.ag = alpha * gain; a.cr_pos = 1.f / gain; a.cr_neg = 1.f / (alpha * gain);
  }
  if (!allow_fast) a.fast_act = 0;
  return launch<false>(a, dtype, st, "modconv_bwd_pack_cl_kernel");
}

}  // namespace vcn

extern "C" int vcn_modconv_bwd_finish_cl(void* dxhat, const void* x, int dtype, const int64_t xs[4], const float* styles,
                                         float* dsum, int n, int c, int c_pad, int h, int w, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(n >= 0 && c >= 1 && h >= 1 && w >= 1, "modconv_bwd_finish_cl: bad shape");
  VCN_REQUIRE(c_pad >= c && c_pad % 8 == 0 && c_pad <= 2048, "modconv_bwd_finish_cl: c_pad must be a multiple of 8 in [c, 2048]");
  VCN_REQUIRE(dtype == VCN_BF16, "modconv_bwd_finish_cl: dxhat / x must be bf16");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(dxhat != nullptr, "modconv_bwd_finish_cl: dxhat must not be NULL");
  VCN_REQUIRE(!dsum || (x && xs && xs[1] == 1), "modconv_bwd_finish_cl: dsum needs a channels-last x");
  BwdClArgs a;
  memset(&a, 0, sizeof(a));
  a.dy = dxhat; a.y = dsum ? x : nullptr;
  a.as[0] = (int64_t)h * w * c_pad; a.as[1] = 1; a.as[2] = (int64_t)w * c_pad; a.as[3] = c_pad;
  for (int i = 0; i < 4; ++i) a.bs[i] = (dsum && xs) ? xs[i] : 0;
  a.scale = styles; a.sum_nc = dsum;
  a.n = n; a.c = c; a.c_pad = c_pad; a.h = h; a.w = w;
  return launch<true>(a, dtype, static_cast<cudaStream_t>(stream), "modconv_bwd_finish_cl_kernel");
}

extern "C" int vcn_torgb_bwd_fused(const void* x, const void* dy0, const float* wgt, const float* styles, const void* acc,
                                   void* dx, float* dsum, int dtype, int n, int c, int h, int w, int cout, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dtype == VCN_BF16, "torgb_bwd_fused: x / dx must be bf16 channels-last");
  VCN_REQUIRE(n >= 0 && h >= 1 && w >= 1 && cout >= 1 && cout <= 4, "torgb_bwd_fused: 1..4 output channels (got %d)", cout);
  const int groups = c >> 3;
  VCN_REQUIRE(c >= 64 && c % 8 == 0 && groups <= 256 && (groups & (groups - 1)) == 0,
              "torgb_bwd_fused: channels (%d) must be a power of two in [64, 2048]", c);
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(x && dy0 && wgt && dx && aligned16(x) && aligned16(dy0) && aligned16(dx), "torgb_bwd_fused: NULL / misaligned pointer");
  const int hw = h * w;
  const int rows = 256 / groups, sp = 2 * rows;
  int per_sample = (sm_count() * 3) / n;
  if (per_sample < 1) per_sample = 1;
  int chunk = (hw + per_sample - 1) / per_sample;
  chunk = (chunk + sp - 1) / sp * sp;
  if (chunk < 4 * sp) chunk = 4 * sp;
  dim3 grid((unsigned)((hw + chunk - 1) / chunk), (unsigned)n);
  VCN_REQUIRE(!acc || aligned16(acc), "torgb_bwd_fused: acc must be 16-byte aligned");
  const size_t smem = acc ? (size_t)3 * (2 * kStageTensorBytes + 1024) : (size_t)kStagesBulk * (kStageTensorBytes + 1024);
  auto kern = acc ? torgb_bwd_bulk_kernel<__nv_bfloat16, true> : torgb_bwd_bulk_kernel<__nv_bfloat16, false>;
  static bool attr_set[2][64] = {};
  int dev = 0;
  VCN_CUDA(cudaGetDevice(&dev));
  const int ai = acc ? 1 : 0;
  if (dev < 0 || dev >= 64 || !attr_set[ai][dev]) {
    VCN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (dev >= 0 && dev < 64) attr_set[ai][dev] = true;
  }
  kern<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy0),
                                                               wgt, styles, static_cast<const __nv_bfloat16*>(acc),
                                                               static_cast<__nv_bfloat16*>(dx), dsum, n, c, hw, cout, chunk);
  VCN_LAUNCH_CHECK("torgb_bwd_bulk_kernel");
  return VCN_OK;
}
