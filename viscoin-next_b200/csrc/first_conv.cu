// first_conv.cu - the 3-channel first convolutions of the frozen networks (VGG-16 conv1_1: 3x3 / stride 1 / pad 1; the
// ResNet stem: 7x7 / stride 2 / pad 3; both 3 -> 64 channels, bias + ReLU) on 16-bit channels-last images.
//
// These layers are write-bound (K = 27 / 147 multiply-adds per output, 128 bytes written per pixel), but the library path
// pads the image to 8 channels and runs a legacy sm80 implicit-GEMM kernel at 160-180 us per call ([16,3,256,256]); the
// tcgen05 implicit GEMM of modconv.cu wastes 7/8 of a 64-channel k-block on them (measured slower still).  Here:
//   CTA tile = 8 x 16 output pixels x 64 channels, one warp per output row; the (8-1)*S+KS x (16-1)*S+KS x 3 input patch
//   is staged in shared memory as it lies in global memory (channel-interleaved), the im2col row of a pixel is addressed
//   through a K-entry offset table, and the 16 x 64 x K product of a warp runs on mma.sync.m16n8k16 (bf16, fp32
//   accumulation) - warp-level tensor-core MMAs are ample for a memory-bound layer; bias + ReLU in registers, the tile is
//   transposed through shared memory and leaves as 16-byte channels-last stores.  Persistent CTAs (weights staged once).
#include "common.cuh"

namespace vcn {
namespace {

constexpr int kFcTH = 8, kFcTW = 16, kFcCout = 64, kFcWPitch = 72, kFcOutPitch = 72;

struct FirstConvArgs {
  const __nv_bfloat16* x;   // [n, h, w, 3]
  const __nv_bfloat16* wp;  // [KP, 64]: k = (ky*KS + kx)*3 + c, rows >= KS*KS*3 are zero
  const float* bias;        // [64]
  __nv_bfloat16* y;         // [n, oh, ow, 64]
  int n, h, w, oh, ow, tiles_x, tiles_y, relu;
};

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int KS, int S, int PAD>
__global__ void __launch_bounds__(256) first_conv_kernel(const FirstConvArgs a) {
  constexpr int K = KS * KS * 3, KP = (K + 15) / 16 * 16;
  constexpr int PH = (kFcTH - 1) * S + KS, PW = (kFcTW - 1) * S + KS;
  constexpr int PITCH = (PW * 3 + 1) / 2 * 2;  // elements per staged patch row
  __shared__ __align__(16) unsigned short patch[PH * PITCH];
  __shared__ __align__(16) uint32_t wsm[(KP / 2) * kFcWPitch];
  __shared__ __align__(16) __nv_bfloat16 outst[kFcTH * kFcTW * kFcOutPitch];
  __shared__ unsigned short koff[KP];
  __shared__ float sbias[kFcCout];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  // weights -> bf16x2 pairs along k; offset table; bias
  for (int i = tid; i < (KP / 2) * kFcCout; i += 256) {
    const int k2 = i / kFcCout, nn = i - k2 * kFcCout;
    const uint32_t lo = __bfloat16_as_ushort(a.wp[(2 * k2) * kFcCout + nn]), hi = __bfloat16_as_ushort(a.wp[(2 * k2 + 1) * kFcCout + nn]);
    wsm[k2 * kFcWPitch + nn] = lo | (hi << 16);
  }
  for (int k = tid; k < KP; k += 256) {
    int off = 0;
    if (k < K) {
      const int ky = k / (KS * 3), r = k - ky * KS * 3;
      off = ky * PITCH + r;  // r = kx*3 + c: contiguous in the channel-interleaved row
    }
    koff[k] = (unsigned short)off;
  }
  if (tid < kFcCout) sbias[tid] = a.bias ? a.bias[tid] : 0.f;

  const int tiles = a.n * a.tiles_y * a.tiles_x;
  const unsigned short* xg = reinterpret_cast<const unsigned short*>(a.x);
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int bx = tile % a.tiles_x, by = (tile / a.tiles_x) % a.tiles_y, n = tile / (a.tiles_x * a.tiles_y);
    const int oy0 = by * kFcTH, ox0 = bx * kFcTW;
    const int iy0 = oy0 * S - PAD, ix0 = ox0 * S - PAD;
    __syncthreads();  // previous tile's patch / staging reads are done (and the tables above are visible)
    for (int i = tid; i < PH * PW * 3; i += 256) {
      const int r = i / (PW * 3), e = i - r * (PW * 3);
      const int px = e / 3;
      const int iy = iy0 + r, ix = ix0 + px;
      unsigned short v = 0;
      if (iy >= 0 && iy < a.h && ix >= 0 && ix < a.w) v = __ldg(xg + ((int64_t)(n * a.h + iy) * a.w + ix0) * 3 + e);
      patch[r * PITCH + e] = v;
    }
    __syncthreads();

    float acc[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[nt][j] = 0.f;
    const unsigned short* p0 = patch + (warp * S) * PITCH + (g * S) * 3;        // output pixel (row warp, col g)
    const unsigned short* p1 = patch + (warp * S) * PITCH + ((g + 8) * S) * 3;  // col g + 8
#pragma unroll 2
    for (int ks = 0; ks < KP / 16; ++ks) {
      const int k0 = ks * 16 + 2 * t;
      const int o0 = koff[k0], o1 = koff[k0 + 1], o2 = koff[k0 + 8], o3 = koff[k0 + 9];
      uint32_t af[4];
      af[0] = (uint32_t)p0[o0] | ((uint32_t)p0[o1] << 16);
      af[1] = (uint32_t)p1[o0] | ((uint32_t)p1[o1] << 16);
      af[2] = (uint32_t)p0[o2] | ((uint32_t)p0[o3] << 16);
      af[3] = (uint32_t)p1[o2] | ((uint32_t)p1[o3] << 16);
      const uint32_t* wr0 = wsm + (ks * 8 + t) * kFcWPitch + g;
      const uint32_t* wr1 = wr0 + 4 * kFcWPitch;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) mma_bf16_16816(acc[nt], af, wr0[nt * 8], wr1[nt * 8]);
    }
    // bias + ReLU, bf16 pairs -> staging tile [pixel][64 (+8 pad)]
    __nv_bfloat16* s0 = outst + (warp * kFcTW + g) * kFcOutPitch;
    __nv_bfloat16* s1 = s0 + 8 * kFcOutPitch;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const int c = nt * 8 + 2 * t;
      float v0 = acc[nt][0] + sbias[c], v1 = acc[nt][1] + sbias[c + 1], v2 = acc[nt][2] + sbias[c], v3 = acc[nt][3] + sbias[c + 1];
      if (a.relu) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); v2 = fmaxf(v2, 0.f); v3 = fmaxf(v3, 0.f); }
      *reinterpret_cast<__nv_bfloat162*>(s0 + c) = __floats2bfloat162_rn(v0, v1);
      *reinterpret_cast<__nv_bfloat162*>(s1 + c) = __floats2bfloat162_rn(v2, v3);
    }
    __syncthreads();
    for (int v = tid; v < kFcTH * kFcTW * 8; v += 256) {
      const int p = v >> 3, ch = v & 7;
      const int oy = oy0 + p / kFcTW, ox = ox0 + p % kFcTW;
      if (oy < a.oh && ox < a.ow) {
        const uint4 val = *reinterpret_cast<const uint4*>(outst + p * kFcOutPitch + ch * 8);
        st_stream<__nv_bfloat16>(a.y + ((int64_t)(n * a.oh + oy) * a.ow + ox) * kFcCout + ch * 8, *reinterpret_cast<const Vec16<__nv_bfloat16>*>(&val));
      }
    }
  }
}

}  // namespace
}  // namespace vcn

extern "C" int vcn_conv_first3(const void* x, const void* wp, const float* bias, void* y, int dtype, int ks, int stride, int pad,
                               int relu, int n, int h, int w, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dtype == VCN_BF16, "conv_first3: bf16 channels-last tensors only (dtype %d)", dtype);
  VCN_REQUIRE((ks == 3 && stride == 1 && pad == 1) || (ks == 7 && stride == 2 && pad == 3),
              "conv_first3: only 3x3/s1/p1 and 7x7/s2/p3 (got %dx%d/s%d/p%d)", ks, ks, stride, pad);
  VCN_REQUIRE(n >= 0 && h >= 1 && w >= 1, "conv_first3: bad shape");
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(x && wp && y && aligned16(y) && (reinterpret_cast<uintptr_t>(x) & 1u) == 0, "conv_first3: NULL / misaligned pointer");
  FirstConvArgs a;
  a.x = static_cast<const __nv_bfloat16*>(x); a.wp = static_cast<const __nv_bfloat16*>(wp); a.bias = bias;
  a.y = static_cast<__nv_bfloat16*>(y);
  a.n = n; a.h = h; a.w = w; a.relu = relu ? 1 : 0;
  a.oh = (h + 2 * pad - ks) / stride + 1; a.ow = (w + 2 * pad - ks) / stride + 1;
  VCN_REQUIRE(a.oh >= 1 && a.ow >= 1, "conv_first3: empty output");
  a.tiles_x = (a.ow + kFcTW - 1) / kFcTW; a.tiles_y = (a.oh + kFcTH - 1) / kFcTH;
  const int64_t tiles = (int64_t)n * a.tiles_x * a.tiles_y;
  VCN_REQUIRE(tiles <= 0x7fffffffLL, "conv_first3: too many tiles");
  int64_t grid = (int64_t)sm_count() * 4;
  if (grid > tiles) grid = tiles;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (ks == 3) first_conv_kernel<3, 1, 1><<<(unsigned)grid, 256, 0, st>>>(a);
  else first_conv_kernel<7, 2, 3><<<(unsigned)grid, 256, 0, st>>>(a);
  VCN_LAUNCH_CHECK("first_conv_kernel");
  return VCN_OK;
}
