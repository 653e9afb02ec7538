// modconv_wgrad.cu - weight gradient of the (modulated) convolution as a tcgen05 / TMEM GEMM (sm_100a).
//
// Reference: autograd of conv2d_gradfix.conv2d / conv_transpose2d (conv2d_gradfix.py:38,43 -> ATen/cuDNN wgrad) under
// training/networks.py:63-72.  With xhat = x * s (bf16 NHWC) and dy0 = dL/d(conv output) (bf16 NHWC):
//     dW[tap][co][ci] = sum_{n, p}  dy0[n, p, co] * xhat[n, p + shift(tap), ci]
// GEMM view per (tap, co tile, ci tile): D[128 co, N ci] += A^T B over K = pixels, where BOTH operands are "MN-major"
// (the contraction index - the pixel - is the slow one in NHWC, channels are contiguous): one TMA box
// (64 ch, TW, TH, 1) lands in shared memory as [64 pixels][64 ch] = 64 rows of 128 bytes, 128B-swizzled, which is exactly the
// canonical MN-major SWIZZLE_128B layout ((T,8,m),(8,k)):((1,T,LBO),(8T,SBO)) with SBO = 1024 bytes (8 pixels) and
// LBO = the 8 KB between two 64-channel boxes.  No transposed copy of either tensor is ever made; out-of-bounds TMA
// zero fill is the convolution's zero padding.
// K is split over CTAs (split-K): every CTA accumulates its pixel range in TMEM and adds the fp32 tile to dW with
// red.global.add.  kinds: 0 = stride-1 correlation (A = dy0 at p, B = xhat at p + tap - pad);
//                         1 = 3x3 stride-2 transposed conv (A = dz at (2i+ky, 2j+kx), B = xhat at (i, j)).
#include <string.h>

#include "umma.cuh"

namespace vcn {
namespace {

constexpr int kWgThreads = 192;
constexpr int kKPix = 64;                       // pixels per k-block = one TMA box
constexpr uint32_t kBoxBytes = kKPix * 64 * 2;  // [64 pixels][64 ch] bf16 = 8 KB

struct WgradArgs {
  int kind, n, taps, kw, pad;
  int cout, cin, n_tile;          // n_tile: ci columns per CTA tile (multiple of 64, <= 256)
  int tiles_co, tiles_ci;
  int tw, th, pix_w, pix_h;       // pixel box and the pixel grid the sum runs over
  int kblocks_total, kb_per_split, splits;
  int stages, tmem_cols;
  float* dw;                      // [cout][cin][taps] fp32, accumulated
};

// MN-major SWIZZLE_128B descriptor: start, LBO (next 64 channels), SBO = 1024 (next 8 pixels)
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)(1024u >> 4) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;
  return d;
}

__device__ __forceinline__ void red_add_f32(float* p, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

struct WgUnit {
  int tap, co0, ci0, kb0, kb1;
};

__device__ __forceinline__ WgUnit wg_decode(const WgradArgs& a, int u) {
  WgUnit r;
  const int split = u % a.splits; u /= a.splits;
  const int ci_t = u % a.tiles_ci; u /= a.tiles_ci;
  const int co_t = u % a.tiles_co;
  r.tap = u / a.tiles_co;
  r.co0 = co_t * 128; r.ci0 = ci_t * a.n_tile;
  r.kb0 = split * a.kb_per_split;
  r.kb1 = min(a.kblocks_total, r.kb0 + a.kb_per_split);
  return r;
}

__global__ void __launch_bounds__(kWgThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ WgradArgs a) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t a_bytes = 2 * kBoxBytes, b_bytes = (uint32_t)(a.n_tile / 64) * kBoxBytes;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)a.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + a.stages;
  uint64_t* tmem_full_bar = empty_bar + a.stages;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;    // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  // shuffled so that the compiler knows the role branches are warp-uniform (see conv_igemm_kernel): TMA / MMA operands stay
  // in uniform registers and the one-lane issue is an elect.sync predicate instead of per-instruction waterfall loops
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int units = a.taps * a.tiles_co * a.tiles_ci * a.splits;
  const int tiles_x = (a.pix_w + a.tw - 1) / a.tw, tiles_y = (a.pix_h + a.th - 1) / a.th;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < a.stages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(tmem_full_bar + s, 1); mbar_init(tmem_empty_bar + s, 4); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_ptr_smem, (uint32_t)a.tmem_cols);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===== TMA producer (whole warp walks the loop, one elected lane issues) =====
    {
      int s = 0;
      uint32_t ph = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const WgUnit w = wg_decode(a, u);
        const int ky = w.tap / a.kw, kx = w.tap % a.kw;
        for (int kb = w.kb0; kb < w.kb1; ++kb) {
          const int tx = kb % tiles_x, ty = (kb / tiles_x) % tiles_y, n = kb / (tiles_x * tiles_y);
          const int x0 = tx * a.tw, y0 = ty * a.th;
          int ax, ay, bx, by;
          if (a.kind == 0) { ax = x0; ay = y0; bx = x0 + kx - a.pad; by = y0 + ky - a.pad; }
          else { ax = 2 * x0 + kx; ay = 2 * y0 + ky; bx = x0; by = y0; }
          mbar_wait(empty_bar + s, ph ^ 1u);
          unsigned char* sa = smem + (size_t)s * stage_bytes;
          if (elect_one()) {
            mbar_expect_tx(full_bar + s, stage_bytes);
            tma_load_4d(sa, &tmap_a, full_bar + s, w.co0, ax, ay, n);
            tma_load_4d(sa + kBoxBytes, &tmap_a, full_bar + s, w.co0 + 64, ax, ay, n);
            for (int j = 0; j < a.n_tile / 64; ++j)
              tma_load_4d(sa + a_bytes + j * kBoxBytes, &tmap_b, full_bar + s, w.ci0 + 64 * j, bx, by, n);
          }
          __syncwarp();
          if (++s == a.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (whole warp walks the loop, one elected lane issues) =====
    {
      const uint32_t idesc = umma_idesc_bf16_f32(128, a.n_tile) | (1u << 15) | (1u << 16);  // A and B MN-major
      int s = 0, it = 0;
      uint32_t ph = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
        const WgUnit w = wg_decode(a, u);
        const int as = it & 1;
        mbar_wait(tmem_empty_bar + as, ((uint32_t)(it >> 1) & 1u) ^ 1u);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * a.n_tile);
        for (int kb = w.kb0; kb < w.kb1; ++kb) {
          mbar_wait(full_bar + s, ph);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)s * stage_bytes);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < kKPix / 16; ++k) {
              // 16 pixels further along K = 16 rows of 128 bytes
              const uint64_t da = umma_desc_mn_sw128(sa + k * 2048u, kBoxBytes);
              const uint64_t db = umma_desc_mn_sw128(sa + a_bytes + k * 2048u, kBoxBytes);
              umma_bf16_ss(d_tmem, da, db, idesc, (kb > w.kb0 || k > 0) ? 1u : 0u);
            }
            umma_commit(empty_bar + s);
          }
          __syncwarp();
          if (++s == a.stages) { s = 0; ph ^= 1u; }
        }
        if (elect_one()) umma_commit(tmem_full_bar + as);
        __syncwarp();
      }
    }
  } else {
    // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4; lane = output channel =====
    const int q = warp & 3;
    int it = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
      const WgUnit w = wg_decode(a, u);
      const int as = it & 1;
      mbar_wait(tmem_full_bar + as, (uint32_t)(it >> 1) & 1u);
      tcgen05_fence_after();
      const int co = w.co0 + q * 32 + lane;
      for (int c = 0; c < a.n_tile; c += 32) {
        if (w.ci0 + c >= a.cin) break;  // warp-uniform
        uint32_t acc[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * a.n_tile + c), acc);
        tmem_ld_wait();
        if (co < a.cout && w.kb1 > w.kb0) {
          float* dst = a.dw + ((int64_t)co * a.cin + w.ci0 + c) * a.taps + w.tap;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (w.ci0 + c + j < a.cin) red_add_f32(dst + (int64_t)j * a.taps, __uint_as_float(acc[j]));
        }
      }
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty_bar + as);
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, (uint32_t)a.tmem_cols);
  }
}

}  // namespace
}  // namespace vcn

extern "C" int vcn_conv_wgrad(const void* dy0, const void* xhat, float* dw, int kind, int n, int h, int w, int cin_pad,
                              int cin, int cout_pad, int cout, int kh, int kw, int dy_h, int dy_w, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(kind == 0 || kind == 1, "conv_wgrad: kind must be 0 (stride-1 conv) or 1 (stride-2 transposed conv)");
  VCN_REQUIRE(n >= 0 && h >= 1 && w >= 1 && cin >= 1 && cout >= 1, "conv_wgrad: bad shape");
  VCN_REQUIRE(cin_pad >= cin && cin_pad % 8 == 0 && cout_pad >= cout && cout_pad % 8 == 0,
              "conv_wgrad: padded channel counts must be multiples of 8");
  VCN_REQUIRE((kh == 1 && kw == 1) || (kh == 3 && kw == 3), "conv_wgrad: only 1x1 and 3x3 kernels (got %dx%d)", kh, kw);
  if (kind == 0) VCN_REQUIRE(dy_h == h && dy_w == w, "conv_wgrad: stride-1 gradient must be %dx%d", h, w);
  else VCN_REQUIRE(kh == 3 && dy_h == 2 * h + 1 && dy_w == 2 * w + 1, "conv_wgrad: transposed-conv gradient must be 3x3, %dx%d", 2 * h + 1, 2 * w + 1);
  if (n == 0) return VCN_OK;
  VCN_REQUIRE(dy0 && xhat && dw, "conv_wgrad: NULL pointer");
  VCN_REQUIRE(aligned16(dy0) && aligned16(xhat), "conv_wgrad: operands must be 16-byte aligned");
  WgradArgs a;
  memset(&a, 0, sizeof(a));
  a.kind = kind; a.n = n; a.taps = kh * kw; a.kw = kw; a.pad = kh / 2;
  a.cout = cout; a.cin = cin; a.dw = dw;
  a.n_tile = (cin + 63) / 64 * 64;
  if (a.n_tile > 256) a.n_tile = 256;
  a.tiles_co = (cout + 127) / 128;
  a.tiles_ci = (cin + a.n_tile - 1) / a.n_tile;
  a.pix_w = w; a.pix_h = h;
  a.tw = 64;
  while (a.tw > 1 && a.tw / 2 >= w) a.tw /= 2;   // smallest power of two >= w, at most 64
  a.th = kKPix / a.tw;
  const int tiles_x = (w + a.tw - 1) / a.tw, tiles_y = (h + a.th - 1) / a.th;
  a.kblocks_total = n * tiles_x * tiles_y;
  const int base_units = a.taps * a.tiles_co * a.tiles_ci;
  int splits = (2 * sm_count() + base_units - 1) / base_units;   // about two waves of work units
  if (splits > a.kblocks_total) splits = a.kblocks_total;
  if (splits < 1) splits = 1;
  a.kb_per_split = (a.kblocks_total + splits - 1) / splits;
  a.splits = (a.kblocks_total + a.kb_per_split - 1) / a.kb_per_split;
  const size_t stage_bytes = 2 * kBoxBytes + (size_t)(a.n_tile / 64) * kBoxBytes;
  const size_t tail_bytes = (2 * (size_t)8 + 4) * 8 + 64;
  a.stages = (int)((220 * 1024 - 1024 - tail_bytes) / stage_bytes);
  if (a.stages > 8) a.stages = 8;
  a.tmem_cols = 2 * a.n_tile;
  int p2 = 32;
  while (p2 < a.tmem_cols) p2 <<= 1;
  a.tmem_cols = p2;
  size_t smem = 1024 + (size_t)a.stages * stage_bytes + tail_bytes;
  if (smem < 120 * 1024) smem = 120 * 1024;  // one CTA per SM (TMEM allocations above 256 columns rely on it)

  CUtensorMap tmap_a, tmap_b;
  {
    const uint32_t m = kind == 1 ? 2u : 1u;
    const uint64_t dims[4] = {(uint64_t)cout_pad, (uint64_t)dy_w, (uint64_t)dy_h, (uint64_t)n};
    const uint64_t strides[3] = {(uint64_t)cout_pad * 2, (uint64_t)dy_w * cout_pad * 2, (uint64_t)dy_h * dy_w * cout_pad * 2};
    const uint32_t box[4] = {64u, (uint32_t)a.tw * m, (uint32_t)a.th * m, 1u};
    const uint32_t estr[4] = {1u, m, m, 1u};
    const int rc = make_tensor_map(&tmap_a, VCN_BF16, 4, dy0, dims, strides, box, estr, 1);
    if (rc) return rc;
  }
  {
    const uint64_t dims[4] = {(uint64_t)cin_pad, (uint64_t)w, (uint64_t)h, (uint64_t)n};
    const uint64_t strides[3] = {(uint64_t)cin_pad * 2, (uint64_t)w * cin_pad * 2, (uint64_t)h * w * cin_pad * 2};
    const uint32_t box[4] = {64u, (uint32_t)a.tw, (uint32_t)a.th, 1u};
    const int rc = make_tensor_map(&tmap_b, VCN_BF16, 4, xhat, dims, strides, box, nullptr, 1);
    if (rc) return rc;
  }
  VCN_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int units = base_units * a.splits;
  int grid = sm_count();
  if (grid > units) grid = units;
  conv_wgrad_kernel<<<(unsigned)grid, kWgThreads, smem, static_cast<cudaStream_t>(stream)>>>(tmap_a, tmap_b, a);
  VCN_LAUNCH_CHECK("conv_wgrad_kernel");
  return VCN_OK;
}
