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
