// frozen_ops.cu - the memory-bound passes around the FROZEN networks of the VisCoIN step (LPIPS VGG-16, the classifier
// f), 16-bit channels-last (NHWC) activations:
//   vcn_maxpool2x2_fwd       y = max over 2x2 windows, stride 2 (VGG-16's pools)
//   vcn_relu_mask_bwd        dz = (g + g2) * (y > 0): the backward of a fused conv + bias + ReLU output, with the second
//                            incoming gradient (LPIPS tap / residual branch) added in the same pass
//   vcn_maxpool2x2_relu_bwd  dz = (gtap + unpool(gpool)) * (y > 0): max-pool backward (gradient to the FIRST maximum of the
//                            window in scan order, like ATen's max_pool2d_with_indices), tap-gradient add and ReLU mask in
//                            ONE pass - ATen runs fill + scatter + add + threshold (4 passes, its NHWC pool backward alone
//                            reaches ~1 TB/s on this shape)
// All three are streaming kernels: one thread = 8 channels (16 bytes) of one pixel / window, 128-bit
// ld.global.nc.L1::no_allocate / st.global.L1::no_allocate, grid-stride, channel groups fastest (coalesced rows).
#include "common.cuh"

namespace vcn {
namespace {

constexpr int kThreads = 256;

template <typename T> __device__ __forceinline__ float f(T v) { return to_acc<T>(v); }

template <typename T>
__global__ void __launch_bounds__(kThreads) maxpool2x2_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t total,
                                                                 int cg, int ow, int64_t in_row /* = w*c elements */, int c) {
  // total = n * oh * ow * cg output vectors; decode (n*oh + oy, ox, g)
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kThreads) {
    const int g = (int)(i % cg);
    const int64_t p = i / cg;
    const int ox = (int)(p % ow);
    const int64_t noy = p / ow;  // n * oh + oy  -> input row pair starts at 2 * noy
    const T* src = x + (noy * 2) * in_row + (int64_t)(2 * ox) * c + g * 8;
    const Vec16<T> a = ld_stream(src), b = ld_stream(src + c), cc = ld_stream(src + in_row), d = ld_stream(src + in_row + c);
    Vec16<T> o;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      T m = a.v[k];
      if (f(b.v[k]) > f(m)) m = b.v[k];
      if (f(cc.v[k]) > f(m)) m = cc.v[k];
      if (f(d.v[k]) > f(m)) m = d.v[k];
      o.v[k] = m;
    }
    st_stream(y + i * 8, o);
  }
}

template <typename T, bool HAS2>
__global__ void __launch_bounds__(kThreads) relu_mask_bwd_kernel(const T* __restrict__ y, const T* __restrict__ g,
                                                                const T* __restrict__ g2, T* __restrict__ dz, int64_t nvec) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * kThreads) {
    const Vec16<T> yy = ld_stream(y + i * 8), gg = ld_stream(g + i * 8);
    Vec16<T> hh;
    if (HAS2) hh = ld_stream(g2 + i * 8);
    Vec16<T> o;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float v = f(gg.v[k]);
      if (HAS2) v += f(hh.v[k]);
      o.v[k] = from_acc<T>(f(yy.v[k]) > 0.f ? v : 0.f);
    }
    st_stream(dz + i * 8, o);
  }
}

template <typename T, bool HAS_TAP>
__global__ void __launch_bounds__(kThreads) maxpool2x2_relu_bwd_kernel(const T* __restrict__ y, const T* __restrict__ gpool,
                                                                      const T* __restrict__ gtap, T* __restrict__ dz,
                                                                      int64_t total, int cg, int ow, int64_t in_row, int c) {
  // one thread = one pooled pixel x 8 channels: reads the 2x2 window of y (and gtap), writes the 2x2 window of dz
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kThreads) {
    const int g = (int)(i % cg);
    const int64_t p = i / cg;
    const int ox = (int)(p % ow);
    const int64_t noy = p / ow;
    const int64_t off = (noy * 2) * in_row + (int64_t)(2 * ox) * c + g * 8;
    const int64_t o4[4] = {off, off + c, off + in_row, off + in_row + c};
    Vec16<T> w[4], t[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      w[q] = ld_stream(y + o4[q]);
      if (HAS_TAP) t[q] = ld_stream(gtap + o4[q]);
    }
    const Vec16<T> gp = ld_stream(gpool + i * 8);
    Vec16<T> out[4];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      int best = 0;
      float m = f(w[0].v[k]);
#pragma unroll
      for (int q = 1; q < 4; ++q) {
        const float v = f(w[q].v[k]);
        if (v > m) { m = v; best = q; }  // strict: the first maximum keeps the gradient (ATen's scan order)
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float v = (q == best) ? f(gp.v[k]) : 0.f;
        if (HAS_TAP) v += f(t[q].v[k]);
        out[q].v[k] = from_acc<T>(f(w[q].v[k]) > 0.f ? v : 0.f);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) st_stream(dz + o4[q], out[q]);
  }
}

// 3x3 / stride 2 / padding 1 max-pool (the ResNet stem, HF ResNetEmbeddings.pooler): one thread = one OUTPUT pixel x 8
// channels.  idx (optional) records which of the 9 window positions (dy*3+dx, first maximum in scan order like ATen's
// max_pool2d_with_indices) won, one byte per element, for the backward pass.  ATen's NHWC kernels run this shape at
// ~0.5 TB/s forward and ~0.4 TB/s backward.
template <typename T>
__global__ void __launch_bounds__(kThreads) maxpool3x3s2_fwd_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                                   unsigned char* __restrict__ idx, int64_t total, int cg,
                                                                   int h, int w, int oh, int ow, int c) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kThreads) {
    const int g = (int)(i % cg);
    int64_t p = i / cg;
    const int ox = (int)(p % ow); p /= ow;
    const int oy = (int)(p % oh);
    const int64_t n = p / oh;
    const T* xn = x + n * (int64_t)h * w * c + g * 8;
    float best[8];
    T bestv[8];
    unsigned char bi[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { best[k] = -INFINITY; bestv[k] = from_acc<T>(-INFINITY); bi[k] = 4; }
#pragma unroll
    for (int dy = 0; dy < 3; ++dy) {
      const int iy = 2 * oy - 1 + dy;
      if (iy < 0 || iy >= h) continue;
#pragma unroll
      for (int dx = 0; dx < 3; ++dx) {
        const int ix = 2 * ox - 1 + dx;
        if (ix < 0 || ix >= w) continue;
        const Vec16<T> v = *reinterpret_cast<const Vec16<T>*>(xn + ((int64_t)iy * w + ix) * c);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float fv = f(v.v[k]);
          if (fv > best[k] || fv != fv) { best[k] = fv; bestv[k] = v.v[k]; bi[k] = (unsigned char)(dy * 3 + dx); }
        }
      }
    }
    Vec16<T> o;
#pragma unroll
    for (int k = 0; k < 8; ++k) o.v[k] = bestv[k];
    st_stream(y + i * 8, o);
    if (idx) {
      uint2 packed;
      packed.x = bi[0] | (bi[1] << 8) | (bi[2] << 16) | ((uint32_t)bi[3] << 24);
      packed.y = bi[4] | (bi[5] << 8) | (bi[6] << 16) | ((uint32_t)bi[7] << 24);
      *reinterpret_cast<uint2*>(idx + i * 8) = packed;
    }
  }
}

// one thread = one INPUT pixel x 8 channels: sums the gradients of the (at most 4) windows that cover it and elected it
template <typename T>
__global__ void __launch_bounds__(kThreads) maxpool3x3s2_bwd_kernel(const T* __restrict__ gy, const unsigned char* __restrict__ idx,
                                                                   T* __restrict__ gx, int64_t total, int cg, int h, int w,
                                                                   int oh, int ow, int c) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < total; i += (int64_t)gridDim.x * kThreads) {
    const int g = (int)(i % cg);
    int64_t p = i / cg;
    const int ix = (int)(p % w); p /= w;
    const int iy = (int)(p % h);
    const int64_t n = p / h;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    // windows: oy = (iy + 1) / 2 with dy = iy + 1 - 2 oy (0 or 1), and, when dy == 0, also oy - 1 with dy = 2
    const int oy0 = (iy + 1) >> 1, dy0 = iy + 1 - 2 * oy0;
    const int ox0 = (ix + 1) >> 1, dx0 = ix + 1 - 2 * ox0;
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const int oy = oy0 - a, dy = dy0 + 2 * a;
      if (a == 1 && dy0 != 0) continue;
      if (oy < 0 || oy >= oh) continue;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int ox = ox0 - b, dx = dx0 + 2 * b;
        if (b == 1 && dx0 != 0) continue;
        if (ox < 0 || ox >= ow) continue;
        const int64_t o = (((n * oh + oy) * ow + ox) * cg + g) * 8;
        const Vec16<T> gv = *reinterpret_cast<const Vec16<T>*>(gy + o);
        const uint2 pk = *reinterpret_cast<const uint2*>(idx + o);
        const unsigned code = (unsigned)(dy * 3 + dx);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const unsigned b8 = ((k < 4 ? pk.x : pk.y) >> (8 * (k & 3))) & 0xffu;
          if (b8 == code) acc[k] += f(gv.v[k]);
        }
      }
    }
    Vec16<T> o;
#pragma unroll
    for (int k = 0; k < 8; ++k) o.v[k] = from_acc<T>(acc[k]);
    st_stream(gx + i * 8, o);
  }
}

inline int grid_for(int64_t work) {
  int64_t blocks = (work + kThreads - 1) / kThreads;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace
}  // namespace vcn

extern "C" int vcn_maxpool2x2_fwd(const void* x, void* y, int dtype, int n, int h, int w, int c, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dtype == VCN_BF16 || dtype == VCN_F16, "maxpool2x2_fwd: 16-bit channels-last tensors only (dtype %d)", dtype);
  VCN_REQUIRE(n >= 0 && h >= 2 && w >= 2 && h % 2 == 0 && w % 2 == 0 && c >= 8 && c % 8 == 0,
              "maxpool2x2_fwd: need even h, w and c %% 8 == 0 (got %dx%dx%d)", h, w, c);
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(x && y && aligned16(x) && aligned16(y), "maxpool2x2_fwd: x and y must be 16-byte aligned, not NULL");
  const int cg = c / 8, ow = w / 2, oh = h / 2;
  const int64_t total = (int64_t)n * oh * ow * cg;
  const int64_t in_row = (int64_t)w * c;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCN_BF16)
    maxpool2x2_fwd_kernel<__nv_bfloat16><<<grid_for(total), kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), total, cg, ow, in_row, c);
  else
    maxpool2x2_fwd_kernel<__half><<<grid_for(total), kThreads, 0, st>>>(static_cast<const __half*>(x), static_cast<__half*>(y), total, cg, ow, in_row, c);
  VCN_LAUNCH_CHECK("maxpool2x2_fwd_kernel");
  return VCN_OK;
}

extern "C" int vcn_relu_mask_bwd(const void* y, const void* g, const void* g2, void* dz, int dtype, int64_t numel, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dtype == VCN_BF16 || dtype == VCN_F16, "relu_mask_bwd: 16-bit tensors only (dtype %d)", dtype);
  VCN_REQUIRE(numel >= 0 && numel % 8 == 0, "relu_mask_bwd: numel (%lld) must be a multiple of 8", (long long)numel);
  if (numel == 0) return VCN_OK;
  VCN_REQUIRE(y && g && dz && aligned16(y) && aligned16(g) && aligned16(dz) && (!g2 || aligned16(g2)),
              "relu_mask_bwd: pointers must be 16-byte aligned, y / g / dz not NULL");
  const int64_t nvec = numel / 8;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define VCN_RM(T, H) relu_mask_bwd_kernel<T, H><<<grid_for(nvec), kThreads, 0, st>>>(static_cast<const T*>(y), static_cast<const T*>(g), static_cast<const T*>(g2), static_cast<T*>(dz), nvec)
  if (dtype == VCN_BF16) { if (g2) VCN_RM(__nv_bfloat16, true); else VCN_RM(__nv_bfloat16, false); }
  else { if (g2) VCN_RM(__half, true); else VCN_RM(__half, false); }
#undef VCN_RM
  VCN_LAUNCH_CHECK("relu_mask_bwd_kernel");
  return VCN_OK;
}

extern "C" int vcn_maxpool2x2_relu_bwd(const void* y, const void* gpool, const void* gtap, void* dz, int dtype, int n, int h,
                                       int w, int c, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dtype == VCN_BF16 || dtype == VCN_F16, "maxpool2x2_relu_bwd: 16-bit channels-last tensors only (dtype %d)", dtype);
  VCN_REQUIRE(n >= 0 && h >= 2 && w >= 2 && h % 2 == 0 && w % 2 == 0 && c >= 8 && c % 8 == 0,
              "maxpool2x2_relu_bwd: need even h, w and c %% 8 == 0 (got %dx%dx%d)", h, w, c);
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(y && gpool && dz && aligned16(y) && aligned16(gpool) && aligned16(dz) && (!gtap || aligned16(gtap)),
              "maxpool2x2_relu_bwd: pointers must be 16-byte aligned, y / gpool / dz not NULL");
  const int cg = c / 8, ow = w / 2, oh = h / 2;
  const int64_t total = (int64_t)n * oh * ow * cg;
  const int64_t in_row = (int64_t)w * c;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define VCN_PB(T, H) maxpool2x2_relu_bwd_kernel<T, H><<<grid_for(total), kThreads, 0, st>>>(static_cast<const T*>(y), static_cast<const T*>(gpool), static_cast<const T*>(gtap), static_cast<T*>(dz), total, cg, ow, in_row, c)
  if (dtype == VCN_BF16) { if (gtap) VCN_PB(__nv_bfloat16, true); else VCN_PB(__nv_bfloat16, false); }
  else { if (gtap) VCN_PB(__half, true); else VCN_PB(__half, false); }
#undef VCN_PB
  VCN_LAUNCH_CHECK("maxpool2x2_relu_bwd_kernel");
  return VCN_OK;
}

extern "C" int vcn_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int dtype, int n, int h, int w, int c, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dtype == VCN_BF16 || dtype == VCN_F16, "maxpool3x3s2_fwd: 16-bit channels-last tensors only (dtype %d)", dtype);
  VCN_REQUIRE(n >= 0 && h >= 1 && w >= 1 && c >= 8 && c % 8 == 0, "maxpool3x3s2_fwd: need c %% 8 == 0 (got %dx%dx%d)", h, w, c);
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(x && y && aligned16(x) && aligned16(y) && (!idx || (reinterpret_cast<uintptr_t>(idx) & 7u) == 0),
              "maxpool3x3s2_fwd: x and y must be 16-byte aligned, idx 8-byte aligned");
  const int cg = c / 8, oh = (h - 1) / 2 + 1, ow = (w - 1) / 2 + 1;
  const int64_t total = (int64_t)n * oh * ow * cg;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCN_BF16)
    maxpool3x3s2_fwd_kernel<__nv_bfloat16><<<grid_for(total), kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), static_cast<unsigned char*>(idx), total, cg, h, w, oh, ow, c);
  else
    maxpool3x3s2_fwd_kernel<__half><<<grid_for(total), kThreads, 0, st>>>(static_cast<const __half*>(x), static_cast<__half*>(y), static_cast<unsigned char*>(idx), total, cg, h, w, oh, ow, c);
  VCN_LAUNCH_CHECK("maxpool3x3s2_fwd_kernel");
  return VCN_OK;
}

extern "C" int vcn_maxpool3x3s2_bwd(const void* gy, const void* idx, void* gx, int dtype, int n, int h, int w, int c, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dtype == VCN_BF16 || dtype == VCN_F16, "maxpool3x3s2_bwd: 16-bit channels-last tensors only (dtype %d)", dtype);
  VCN_REQUIRE(n >= 0 && h >= 1 && w >= 1 && c >= 8 && c % 8 == 0, "maxpool3x3s2_bwd: need c %% 8 == 0 (got %dx%dx%d)", h, w, c);
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(gy && idx && gx && aligned16(gy) && aligned16(gx) && (reinterpret_cast<uintptr_t>(idx) & 7u) == 0,
              "maxpool3x3s2_bwd: gy and gx must be 16-byte aligned, idx 8-byte aligned, none NULL");
  const int cg = c / 8, oh = (h - 1) / 2 + 1, ow = (w - 1) / 2 + 1;
  const int64_t total = (int64_t)n * h * w * cg;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == VCN_BF16)
    maxpool3x3s2_bwd_kernel<__nv_bfloat16><<<grid_for(total), kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(gy), static_cast<const unsigned char*>(idx), static_cast<__nv_bfloat16*>(gx), total, cg, h, w, oh, ow, c);
  else
    maxpool3x3s2_bwd_kernel<__half><<<grid_for(total), kThreads, 0, st>>>(static_cast<const __half*>(gy), static_cast<const unsigned char*>(idx), static_cast<__half*>(gx), total, cg, h, w, oh, ow, c);
  VCN_LAUNCH_CHECK("maxpool3x3s2_bwd_kernel");
  return VCN_OK;
}
