// image_ops.cu - the elementwise passes over the [N,3,H,W] fp32 images of the VisCoIN iteration, fused:
//   vcn_image_prep_fwd/bwd   y[n,h,w,3] (bf16, channels-last) = (x[n,c,h,w] - shift[c]) * inv_scale[c]   and its adjoint
//       (the input conversion of the frozen classifier f, shift 0 / scale 1, and LPIPS's ScalingLayer, lpips v0.1.4 through
//       viscoin/training/losses.py:111-123: torch ran sub, div, cast and layout copy as four passes, and four more backward)
//   vcn_l1l2_fwd/bwd         (sum |a-b|, sum (a-b)^2) and d/da of w1*mean|a-b| + w2*mean (a-b)^2
//       (reconstruction_loss, losses.py:125-157: subtract, abs, square, two single-CTA mean reductions; sign / mul / add backward)
// fp32 NCHW in, one thread per pixel (plane reads coalesce), grid-stride.
#include "common.cuh"

namespace vcn {
namespace {

__global__ void __launch_bounds__(256) image_prep_fwd_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                                             const float* __restrict__ shift, const float* __restrict__ inv_scale,
                                                             int64_t total, int hw) {
  const float s0 = shift ? shift[0] : 0.f, s1 = shift ? shift[1] : 0.f, s2 = shift ? shift[2] : 0.f;
  const float k0 = inv_scale ? inv_scale[0] : 1.f, k1 = inv_scale ? inv_scale[1] : 1.f, k2 = inv_scale ? inv_scale[2] : 1.f;
  for (int64_t i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t n = i / hw, p = i - n * hw;
    const float* xp = x + n * 3 * hw + p;
    __nv_bfloat16* yp = y + i * 3;
    yp[0] = __float2bfloat16_rn((xp[0] - s0) * k0);
    yp[1] = __float2bfloat16_rn((xp[hw] - s1) * k1);
    yp[2] = __float2bfloat16_rn((xp[2 * (int64_t)hw] - s2) * k2);
  }
}

__global__ void __launch_bounds__(256) image_prep_bwd_kernel(const __nv_bfloat16* __restrict__ gy, float* __restrict__ gx,
                                                             const float* __restrict__ inv_scale, int64_t total, int hw) {
  const float k0 = inv_scale ? inv_scale[0] : 1.f, k1 = inv_scale ? inv_scale[1] : 1.f, k2 = inv_scale ? inv_scale[2] : 1.f;
  for (int64_t i = blockIdx.x * 256ll + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t n = i / hw, p = i - n * hw;
    const __nv_bfloat16* gp = gy + i * 3;
    float* xp = gx + n * 3 * hw + p;
    xp[0] = __bfloat162float(gp[0]) * k0;
    xp[hw] = __bfloat162float(gp[1]) * k1;
    xp[2 * (int64_t)hw] = __bfloat162float(gp[2]) * k2;
  }
}

__global__ void __launch_bounds__(256) l1l2_fwd_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                       float* __restrict__ out, int64_t nvec) {
  float s1 = 0.f, s2 = 0.f;
  for (int64_t i = blockIdx.x * 256ll + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * 256) {
    const float4 va = reinterpret_cast<const float4*>(a)[i], vb = reinterpret_cast<const float4*>(b)[i];
    const float d0 = va.x - vb.x, d1 = va.y - vb.y, d2 = va.z - vb.z, d3 = va.w - vb.w;
    s1 += fabsf(d0) + fabsf(d1) + fabsf(d2) + fabsf(d3);
    s2 += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
  }
  __shared__ float red[2][8];
  s1 = warp_sum(s1); s2 = warp_sum(s2);
  if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = s1; red[1][threadIdx.x >> 5] = s2; }
  __syncthreads();
  if (threadIdx.x < 2) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += red[threadIdx.x][k];
    atomicAdd(out + threadIdx.x, t);
  }
}

// ga = g[0] * sign(a-b) + g[1] * 2 (a-b)     (g already carries the 1/numel of the means and the loss weights)
__global__ void __launch_bounds__(256) l1l2_bwd_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                       const float* __restrict__ g, float* __restrict__ ga, int64_t nvec) {
  const float g1 = g[0], g2 = 2.f * g[1];
  for (int64_t i = blockIdx.x * 256ll + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * 256) {
    const float4 va = reinterpret_cast<const float4*>(a)[i], vb = reinterpret_cast<const float4*>(b)[i];
    const float d[4] = {va.x - vb.x, va.y - vb.y, va.z - vb.z, va.w - vb.w};
    float o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = (d[k] > 0.f ? g1 : (d[k] < 0.f ? -g1 : 0.f)) + g2 * d[k];
    reinterpret_cast<float4*>(ga)[i] = make_float4(o[0], o[1], o[2], o[3]);
  }
}

inline int grid_of(int64_t work) {
  int64_t blocks = (work + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  return (int)(blocks < 1 ? 1 : blocks);
}

}  // namespace
}  // namespace vcn

extern "C" int vcn_image_prep_fwd(const float* x, void* y, const float* shift, const float* inv_scale, int n, int h, int w,
                                  void* stream) {
  using namespace vcn;
  VCN_REQUIRE(n >= 0 && h >= 1 && w >= 1, "image_prep_fwd: bad shape");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(x && y, "image_prep_fwd: NULL pointer");
  const int64_t total = (int64_t)n * h * w;
  image_prep_fwd_kernel<<<grid_of(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, static_cast<__nv_bfloat16*>(y), shift, inv_scale,
                                                                                         total, h * w);
  VCN_LAUNCH_CHECK("image_prep_fwd_kernel");
  return VCN_OK;
}

extern "C" int vcn_image_prep_bwd(const void* gy, float* gx, const float* inv_scale, int n, int h, int w, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(n >= 0 && h >= 1 && w >= 1, "image_prep_bwd: bad shape");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(gy && gx, "image_prep_bwd: NULL pointer");
  const int64_t total = (int64_t)n * h * w;
  image_prep_bwd_kernel<<<grid_of(total), 256, 0, static_cast<cudaStream_t>(stream)>>>(static_cast<const __nv_bfloat16*>(gy), gx, inv_scale,
                                                                                         total, h * w);
  VCN_LAUNCH_CHECK("image_prep_bwd_kernel");
  return VCN_OK;
}

extern "C" int vcn_l1l2_fwd(const float* a, const float* b, float* out2, int64_t numel, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(numel >= 0 && numel % 4 == 0, "l1l2_fwd: numel (%lld) must be a multiple of 4", (long long)numel);
  if (numel == 0) return VCN_OK;
  VCN_REQUIRE(a && b && out2 && aligned16(a) && aligned16(b), "l1l2_fwd: NULL / misaligned pointer");
  l1l2_fwd_kernel<<<grid_of(numel / 4), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, b, out2, numel / 4);
  VCN_LAUNCH_CHECK("l1l2_fwd_kernel");
  return VCN_OK;
}

extern "C" int vcn_l1l2_bwd(const float* a, const float* b, const float* g2, float* ga, int64_t numel, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(numel >= 0 && numel % 4 == 0, "l1l2_bwd: numel (%lld) must be a multiple of 4", (long long)numel);
  if (numel == 0) return VCN_OK;
  VCN_REQUIRE(a && b && g2 && ga && aligned16(a) && aligned16(b) && aligned16(ga), "l1l2_bwd: NULL / misaligned pointer");
  l1l2_bwd_kernel<<<grid_of(numel / 4), 256, 0, static_cast<cudaStream_t>(stream)>>>(a, b, g2, ga, numel / 4);
  VCN_LAUNCH_CHECK("l1l2_bwd_kernel");
  return VCN_OK;
}
