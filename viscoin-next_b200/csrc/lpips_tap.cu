// lpips_tap.cu - one feature tap of the LPIPS perceptual distance, fused (the reconstruction-loss
// reductions of viscoin/training/losses.py:113-157; algorithm of the third-party `lpips` v0.1.4 package):
//
//   d[n] = mean_p  sum_c lin[c] * ( a[n,p,c] / (|a[n,p,:]| + eps)  -  b[n,p,c] / (|b[n,p,:]| + eps) )^2,   eps = 1e-10
//
// a, b: channels-fastest feature maps (VGG-16 relu taps, 64..512 channels, 16-bit under autocast or fp32).  The torch
// formulation materialises ~10 full-size fp32 temporaries forward and as many backward; here the forward reads a and b
// once (a lane group keeps a pixel's channel vector in registers: two phases - norms, then the weighted squared
// difference - without re-reading) and the backward reads them once more and writes da.  HBM-bound.
//   backward:  da_c = g[n]/HW * ( 2 lin_c u_c / na  -  2 a_c T / (na^2 |a|) ),  u = a/na - b/nb,  T = sum_c lin_c u_c a_c,
//              na = |a| + eps (the second term is 0 where |a| = 0).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace vcn {
namespace {

constexpr int kMaxVec = 2;  // 16-byte... 8-channel vectors per lane: C <= 8 * 32 * kMaxVec = 512

struct LpipsArgs {
  const void* a; const void* b;
  int64_t as[4], bs[4];   // (n, c, h, w) element strides, c-stride 1
  const float* lin;
  float* val;             // [n], accumulated (forward)
  const float* gval;      // [n] (backward)
  void* da;               // dense channels-last [n,h,w,c], dtype of a (backward)
  int n, c, h, w;
  float scale;            // 1 / (h*w)
};

template <typename T> __device__ __forceinline__ void ld8(const T* p, float (&v)[8]) {
  if constexpr (sizeof(T) == 2) {
    const Vec16<T> in = *reinterpret_cast<const Vec16<T>*>(p);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = to_acc<T>(in.v[k]);
  } else {
    const float4 lo = *reinterpret_cast<const float4*>(p), hi = *reinterpret_cast<const float4*>(p + 4);
    v[0] = lo.x; v[1] = lo.y; v[2] = lo.z; v[3] = lo.w; v[4] = hi.x; v[5] = hi.y; v[6] = hi.z; v[7] = hi.w;
  }
}

__device__ __forceinline__ float group_sum(float v, int width) {
  for (int o = width >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// G lanes per pixel (power of two <= 32), NV vectors of 8 channels per lane: C = 8 * G * NV.
template <typename T, bool BWD>
__global__ void __launch_bounds__(256) lpips_tap_kernel(const LpipsArgs a, const int G, const int NV, const int chunk) {
  const int n = blockIdx.y;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / G, gl = lane % G;         // pixel slot inside the warp, lane inside the group
  const int ppw = 32 / G;                          // pixels per warp per iteration
  const int hw = a.h * a.w;
  const int p_end = min(hw, (int)(blockIdx.x + 1) * chunk);
  float lin[kMaxVec][8];
#pragma unroll
  for (int v = 0; v < kMaxVec; ++v)
#pragma unroll
    for (int k = 0; k < 8; ++k) lin[v][k] = v < NV ? a.lin[(v * G + gl) * 8 + k] : 0.f;
  const T* pa = static_cast<const T*>(a.a) + (int64_t)n * a.as[0];
  const T* pb = static_cast<const T*>(a.b) + (int64_t)n * a.bs[0];
  const float g = BWD ? a.gval[n] * a.scale : 0.f;
  float acc = 0.f;
  for (int p0 = blockIdx.x * chunk + warp * ppw; p0 < p_end; p0 += 8 * ppw) {
    const int p = p0 + sub;
    const bool ok = p < p_end;
    const int py = ok ? p / a.w : 0, px = ok ? p - py * a.w : 0;
    float x[kMaxVec][8], y[kMaxVec][8];
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int v = 0; v < kMaxVec; ++v) {
      if (v < NV && ok) {
        const int c = (v * G + gl) * 8;
        ld8<T>(pa + py * a.as[2] + px * a.as[3] + c, x[v]);
        ld8<T>(pb + py * a.bs[2] + px * a.bs[3] + c, y[v]);
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) x[v][k] = y[v][k] = 0.f;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) { s0 = fmaf(x[v][k], x[v][k], s0); s1 = fmaf(y[v][k], y[v][k], s1); }
    }
    s0 = group_sum(s0, G); s1 = group_sum(s1, G);
    const float n0 = sqrtf(s0), ia = 1.f / (n0 + 1e-10f), ib = 1.f / (sqrtf(s1) + 1e-10f);
    float d = 0.f;  // forward: sum lin u^2.  backward: T = sum lin u a
#pragma unroll
    for (int v = 0; v < kMaxVec; ++v)
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float u = x[v][k] * ia - y[v][k] * ib;
        d = fmaf(lin[v][k] * u, BWD ? x[v][k] : u, d);
      }
    d = group_sum(d, G);
    if (!BWD) {
      if (ok && gl == 0) acc += d;
    } else if (ok) {
      const float k2 = n0 > 0.f ? 2.f * g * d * ia * ia / n0 : 0.f, k1 = 2.f * g * ia;
#pragma unroll
      for (int v = 0; v < kMaxVec; ++v) {
        if (v < NV) {
          const int c = (v * G + gl) * 8;
          T* dst = static_cast<T*>(a.da) + ((int64_t)n * hw + p) * a.c + c;
          float o[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] = k1 * lin[v][k] * (x[v][k] * ia - y[v][k] * ib) - k2 * x[v][k];
          if constexpr (sizeof(T) == 2) {
            Vec16<T> ov;
#pragma unroll
            for (int k = 0; k < 8; ++k) ov.v[k] = from_acc<T>(o[k]);
            *reinterpret_cast<Vec16<T>*>(dst) = ov;
          } else {
            *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<float4*>(dst + 4) = make_float4(o[4], o[5], o[6], o[7]);
          }
        }
      }
    }
  }
  if (!BWD) {
    __shared__ float red[8];
    acc = warp_sum(acc);
    if (lane == 0) red[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < 8; ++i) t += red[i];
      atomicAdd(a.val + n, t * a.scale);
    }
  }
}

// Bulk-copy staged variant (dense 16-bit channels-last maps): the register-staged kernel above keeps ~25 KB in flight per
// SM (0.31 of the HBM peak on the [16,64,256,256] tap).  Here the TMA engine streams both maps into a shared-memory ring
// (cp.async.bulk, 8 KB per map and stage, 4 stages, 3 CTAs per SM) and the lane groups read their channel vectors from it.
constexpr int kLpStages = 4;
constexpr uint32_t kLpStageBytes = 8192;  // per map and stage

// G (lanes per pixel) and NV (8-channel vectors per lane) are compile-time here: the shuffle trees and the per-vector loops
// unroll exactly (the run-time form computes both kMaxVec vectors and loops over the shuffle distance).
template <int G> __device__ __forceinline__ float group_sum_c(float v) {
#pragma unroll
  for (int o = G >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <typename T, bool BWD, int G, int NV>
__global__ void __launch_bounds__(256, 3) lpips_tap_bulk_kernel(const LpipsArgs a, const int chunk) {
  extern __shared__ __align__(128) unsigned char ring[];  // [kLpStages][2][kLpStageBytes]
  __shared__ __align__(8) uint64_t full_bar[kLpStages];
  __shared__ float red[8];
  const int n = blockIdx.y, tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int sub = lane / G, gl = lane % G;
  const int ppw = 32 / G;
  const int hw = a.h * a.w;
  const int p0 = blockIdx.x * chunk, p_end = min(hw, p0 + chunk);
  const uint32_t row_bytes = (uint32_t)a.c * 2u;
  const int sp = (int)(kLpStageBytes / row_bytes);  // pixels per stage, a multiple of 8 * ppw
  const int nit = (p_end - p0 + sp - 1) / sp;
  const unsigned char* ga = static_cast<const unsigned char*>(a.a) + ((int64_t)n * a.as[0] + (int64_t)p0 * a.c) * 2;
  const unsigned char* gb = static_cast<const unsigned char*>(a.b) + ((int64_t)n * a.bs[0] + (int64_t)p0 * a.c) * 2;
  auto fill = [&](int it) {
    const int s = it % kLpStages;
    const int px = min(sp, p_end - p0 - it * sp);
    const uint32_t bytes = (uint32_t)px * row_bytes;
    unsigned char* dst = ring + (size_t)s * 2 * kLpStageBytes;
    mbar_expect_tx(full_bar + s, 2u * bytes);
    bulk_load_1d(dst, ga + (int64_t)it * sp * row_bytes, bytes, full_bar + s);
    bulk_load_1d(dst + kLpStageBytes, gb + (int64_t)it * sp * row_bytes, bytes, full_bar + s);
  };
  if (tid == 0) {
    for (int s = 0; s < kLpStages; ++s) mbar_init(full_bar + s, 1);
    fence_mbar_init();
    for (int it = 0; it < min(nit, kLpStages); ++it) fill(it);
  }
  float lin[NV][8];
#pragma unroll
  for (int v = 0; v < NV; ++v)
#pragma unroll
    for (int k = 0; k < 8; ++k) lin[v][k] = a.lin[(v * G + gl) * 8 + k];
  const float g = BWD ? a.gval[n] * a.scale : 0.f;
  float acc = 0.f;
  __syncthreads();
  for (int it = 0; it < nit; ++it) {
    const int s = it % kLpStages;
    mbar_wait(full_bar + s, (uint32_t)(it / kLpStages) & 1u);
    const unsigned char* sa = ring + (size_t)s * 2 * kLpStageBytes;
    for (int pl0 = warp * ppw; pl0 < sp; pl0 += 8 * ppw) {
      const int pl = pl0 + sub;
      const int p = p0 + it * sp + pl;
      const bool ok = p < p_end;
      float x[NV][8], y[NV][8];
      float s0 = 0.f, s1 = 0.f;
#pragma unroll
      for (int v = 0; v < NV; ++v) {
        if (ok) {
          const size_t off = (size_t)pl * row_bytes + (size_t)(v * G + gl) * 16;
          const Vec16<T> va = *reinterpret_cast<const Vec16<T>*>(sa + off);
          const Vec16<T> vb = *reinterpret_cast<const Vec16<T>*>(sa + kLpStageBytes + off);
#pragma unroll
          for (int k = 0; k < 8; ++k) { x[v][k] = to_acc<T>(va.v[k]); y[v][k] = to_acc<T>(vb.v[k]); }
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) x[v][k] = y[v][k] = 0.f;
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) { s0 = fmaf(x[v][k], x[v][k], s0); s1 = fmaf(y[v][k], y[v][k], s1); }
      }
      s0 = group_sum_c<G>(s0); s1 = group_sum_c<G>(s1);
      const float n0 = sqrtf(s0), ia = 1.f / (n0 + 1e-10f), ib = 1.f / (sqrtf(s1) + 1e-10f);
      float d = 0.f;
#pragma unroll
      for (int v = 0; v < NV; ++v)
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float u = x[v][k] * ia - y[v][k] * ib;
          d = fmaf(lin[v][k] * u, BWD ? x[v][k] : u, d);
        }
      d = group_sum_c<G>(d);
      if (!BWD) {
        if (ok && gl == 0) acc += d;
      } else if (ok) {
        const float k2 = n0 > 0.f ? 2.f * g * d * ia * ia / n0 : 0.f, k1 = 2.f * g * ia;
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          {
            const int c = (v * G + gl) * 8;
            Vec16<T> ov;
#pragma unroll
            for (int k = 0; k < 8; ++k)
              ov.v[k] = from_acc<T>(k1 * lin[v][k] * (x[v][k] * ia - y[v][k] * ib) - k2 * x[v][k]);
            st_stream<T>(static_cast<T*>(a.da) + ((int64_t)n * hw + p) * a.c + c, ov);
          }
        }
      }
    }
    __syncthreads();
    if (tid == 0 && it + kLpStages < nit) {
      fence_proxy_async();
      fill(it + kLpStages);
    }
  }
  if (!BWD) {
    acc = warp_sum(acc);
    if (lane == 0) red[warp] = acc;
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
      for (int i = 0; i < 8; ++i) t += red[i];
      atomicAdd(a.val + n, t * a.scale);
    }
  }
}

template <typename T, bool BWD, int G, int NV>
int launch_bulk(const LpipsArgs& a, cudaStream_t st, const char* name) {
  const int hw = a.h * a.w;
  const int sp = (int)(kLpStageBytes / ((uint32_t)a.c * 2u));
  int per_sample = (sm_count() * 3) / a.n;
  if (per_sample < 1) per_sample = 1;
  int chunk = (hw + per_sample - 1) / per_sample;
  chunk = (chunk + sp - 1) / sp * sp;
  if (chunk < 4 * sp) chunk = 4 * sp;
  dim3 grid((unsigned)((hw + chunk - 1) / chunk), (unsigned)a.n);
  const size_t smem = (size_t)kLpStages * 2 * kLpStageBytes;
  auto kern = lpips_tap_bulk_kernel<T, BWD, G, NV>;
  static bool attr_set[64] = {};
  int dev = 0;
  VCN_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    VCN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (dev >= 0 && dev < 64) attr_set[dev] = true;
  }
  kern<<<grid, 256, smem, st>>>(a, chunk);
  VCN_LAUNCH_CHECK(name);
  return VCN_OK;
}

template <bool BWD>
int run(const void* fa, const void* fb, int dtype, const int64_t* as, const int64_t* bs, const float* lin, float* val,
        const float* gval, void* da, int n, int c, int h, int w, cudaStream_t st, const char* name) {
  VCN_REQUIRE(n >= 0 && c >= 8 && h >= 1 && w >= 1, "%s: bad shape", name);
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(fa && fb && lin && as && bs, "%s: NULL pointer", name);
  VCN_REQUIRE(as[1] == 1 && bs[1] == 1, "%s: feature maps must be channels-last (channel stride 1)", name);
  const int vecs = c / 8;
  VCN_REQUIRE(c % 8 == 0 && vecs <= 32 * kMaxVec, "%s: channels (%d) must be a multiple of 8, at most %d", name, c, 256 * kMaxVec);
  int G = 1;
  while (G < 32 && G < vecs) G <<= 1;
  const int NV = (vecs + G - 1) / G;
  VCN_REQUIRE(G * NV == vecs, "%s: channels/8 (%d) must be a power of two or a multiple of 32", name, vecs);
  const int64_t align = dtype == VCN_F32 ? 4 : 8;
  VCN_REQUIRE(aligned16(fa) && aligned16(fb) && (as[0] | as[2] | as[3]) % align == 0 && (bs[0] | bs[2] | bs[3]) % align == 0,
              "%s: feature maps must be 16-byte aligned per pixel", name);
  LpipsArgs a;
  memset(&a, 0, sizeof(a));
  a.a = fa; a.b = fb; a.lin = lin; a.val = val; a.gval = gval; a.da = da;
  for (int i = 0; i < 4; ++i) { a.as[i] = as[i]; a.bs[i] = bs[i]; }
  a.n = n; a.c = c; a.h = h; a.w = w; a.scale = 1.f / ((float)h * (float)w);
  const int hw = h * w;
  {
    // dense 16-bit channels-last maps of at least 4 MB: the bulk-copy staged kernel
    static const bool allow = [] { const char* e = getenv("VCN_LPIPS_BULK"); return !e || atoi(e) != 0; }();
    auto dense = [&](const int64_t* s) { return s[3] == c && s[2] == (int64_t)w * c && (s[0] * 2) % 16 == 0; };
    if (allow && dtype != VCN_F32 && dense(as) && dense(bs) && (int64_t)n * hw * c >= (1 << 21) && 8192 % (c * 2) == 0) {
#define VCN_LP(GG, VV)                                                                         \
  if (G == GG && NV == VV)                                                                    \
    return dtype == VCN_BF16 ? launch_bulk<__nv_bfloat16, BWD, GG, VV>(a, st, name) : launch_bulk<__half, BWD, GG, VV>(a, st, name)
      VCN_LP(8, 1); VCN_LP(16, 1); VCN_LP(32, 1); VCN_LP(32, 2);
#undef VCN_LP
    }
  }
  int chunk = hw;
  const int per_iter = 8 * (32 / G);
  while (chunk > 8 * per_iter && (int64_t)n * ((hw + chunk - 1) / chunk) < (int64_t)sm_count() * 16) chunk = (chunk + 1) / 2;
  chunk = (chunk + per_iter - 1) / per_iter * per_iter;
  dim3 grid((unsigned)((hw + chunk - 1) / chunk), (unsigned)n);
  switch (dtype) {
    case VCN_F32: lpips_tap_kernel<float, BWD><<<grid, 256, 0, st>>>(a, G, NV, chunk); break;
    case VCN_F16: lpips_tap_kernel<__half, BWD><<<grid, 256, 0, st>>>(a, G, NV, chunk); break;
    case VCN_BF16: lpips_tap_kernel<__nv_bfloat16, BWD><<<grid, 256, 0, st>>>(a, G, NV, chunk); break;
    default: set_error("%s: unsupported dtype %d", name, dtype); return VCN_ERR_INVALID;
  }
  VCN_LAUNCH_CHECK(name);
  return VCN_OK;
}

}  // namespace
}  // namespace vcn

extern "C" int vcn_lpips_tap_fwd(const void* fa, const void* fb, int dtype, const int64_t as[4], const int64_t bs[4],
                                 const float* lin, float* val, int n, int c, int h, int w, void* stream) {
  VCN_REQUIRE(val != nullptr || n == 0, "lpips_tap_fwd: val must not be NULL");
  return vcn::run<false>(fa, fb, dtype, as, bs, lin, val, nullptr, nullptr, n, c, h, w, static_cast<cudaStream_t>(stream),
                         "lpips_tap_fwd");
}

extern "C" int vcn_lpips_tap_bwd(const void* fa, const void* fb, int dtype, const int64_t as[4], const int64_t bs[4],
                                 const float* lin, const float* gval, void* da, int n, int c, int h, int w, void* stream) {
  VCN_REQUIRE((gval != nullptr && da != nullptr) || n == 0, "lpips_tap_bwd: gval and da must not be NULL");
  return vcn::run<true>(fa, fb, dtype, as, bs, lin, nullptr, gval, da, n, c, h, w, static_cast<cudaStream_t>(stream),
                        "lpips_tap_bwd");
}
