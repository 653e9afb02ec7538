// modconv.cu - the convolution of modulated_conv2d as a TMA-fed tcgen05 / TMEM implicit GEMM (sm_100a).
//
// Replaces the ATen/cuDNN convolution under training/networks.py:27-84 (conv2d_resample.py:59-154 ->
// conv2d_gradfix.py:38,43).  GEMM view, per CTA tile:
//   D[128 pixels, N_TILE out-channels] (fp32, TMEM) = sum over taps t, channel blocks cb of
//        A_t,cb[128 pixels, 64 ch] (bf16 smem, K-major, 128B swizzle)  x  B_t,cb[N_TILE, 64 ch]^T (bf16 smem, K-major)
//   A = one TMA box (64ch, TW, TH, TN) of the NHWC activation at the tap's pixel shift: out-of-bounds zero fill IS the
//       convolution's zero padding (negative coordinates included), TW*TH*TN = 128 rows.
//   B = one TMA box (64ch, N_TILE, 1) of the packed weights [tap][cout][cin].
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane, tcgen05.mma
// cta_group::1 kind::f16, M=128), warps 2..9 = epilogue (two per lane quarter) (tcgen05.ld of their TMEM lane quarter, then
// demodulation * d[n,co] + noise, bias + activation + gain + clamp, store).  smem ring of S stages with
// full/empty mbarriers; tcgen05.commit releases a stage and, after the last k-block, hands TMEM to the epilogue.
// A "plan" lists the taps; taps are partitioned into up to 4 accumulator groups (TMEM column ranges): the plain
// convolution has one group of kh*kw taps, the stride-2 transposed convolution has its 4 output phases as groups
// (4+2+2+1 taps) so that it costs exactly the reference's FLOPs (2*9*Cin*Cout per INPUT pixel).
// CTA pairs (`pair`): for wide out-channel tiles the kernel runs as 2-CTA clusters with tcgen05.mma.cta_group::2 - one
// MMA covers 256 pixels (128 accumulator rows in each CTA's TMEM) x N_TILE channels, each CTA stages its own 128-pixel A
// tile and HALF of the weight tile, and the tensor cores of both SMs read the two halves from both shared memories: per
// flop every SM ingests 1/3 less operand bytes through TMA and its shared-memory port.  Both producers complete their TMA
// bytes on the LEADER's full barrier; the leader's MMA thread releases stages / publishes accumulators to both CTAs with
// multicast commits; the peer's epilogue warps hand accumulator stages back with remote mbarrier arrives.
#include <stdlib.h>
#include <string.h>

#include "act_math.cuh"
#include "umma.cuh"

namespace vcn {
namespace {

constexpr int kBlockM = 128;  // pixels per CTA tile = UMMA M
constexpr int kBlockK = 64;   // bf16 channels per k-block = one 128-byte swizzle span
constexpr int kUmmaK = 16;
constexpr int kMaxTaps = 9;
constexpr int kMaxPhases = 4;
constexpr int kEpilogueWarps = 8;  // two warps per TMEM lane quarter, alternating 32-column chunks (the epilogue bounds the small-K tiles)
constexpr int kThreads = 64 + kEpilogueWarps * 32;
constexpr uint32_t kABytes = kBlockM * kBlockK * 2;
#ifndef VCN_CONV_PROF
#define VCN_CONV_PROF 0  // 1: per-role cycle counters (tools/prof_kernels.py, debug builds only); 0 in the product library
#endif
constexpr bool kProf = VCN_CONV_PROF != 0;

struct ConvTap {
  int dy, dx;  // pixel shift of the A tile in the M grid
  int wtap;    // index of the tap in the packed weights
};

struct IgemmArgs {
  ConvTap taps[kMaxTaps];               // sorted by output phase
  int phase_first[kMaxPhases + 1];      // taps of phase p: [phase_first[p], phase_first[p+1])
  int num_phases;
  int g_oy[kMaxPhases], g_ox[kMaxPhases];  // output pixel = (py * o_mul + g_oy, px * o_mul + g_ox)
  int o_mul;
  int a_mul;           // A-tile origin = tile origin * a_mul + tap shift (2 for the stride-2 convolution)
  int grid_h, grid_w;  // extent of the M grid (pixels that own accumulator rows)
  int n, cout, out_h, out_w;
  int tw, th, tn, tiles_w, tiles_h, tiles_m, tiles_total;
  int n_tile, kblocks_c, stages, tmem_cols;
  int swap;            // 1: operands swapped - UMMA M = 128 out-channels (weights), N = 256 pixels (see header)
  int tf32;            // 1: fp32 operands, tcgen05 kind::tf32 (a k-block = 32 channels = one 128-byte swizzle span)
  int kbk;             // channels per k-block: 64 (bf16) or 32 (tf32)
  int pair;            // 1: 2-CTA clusters, cta_group::2 (tw/th/tn describe ONE CTA's 128-pixel half of the pair tile)
  int pair_dim;        // the pair tile = two half tiles stacked along w (0), h (1) or n (2)
  int log_tw, log_th;  // swap: tile extents are powers of two
  uint32_t pix_bytes, w_bytes;  // bytes of the pixel / weight operand in one pipeline stage
  // staged output (16-bit channels-fastest y): the epilogue writes a 128B-swizzled [pixel][64 ch] tile per 64-channel
  // block into shared memory and ONE thread hands it to TMA (full-line writes, bounds clipped by the tensor map)
  int stage_out, out_blocks;
  int prof;
  uint32_t stage_bytes_out, stage_blk_bytes;
  int w_per_sample;
  void* y;
  int y_dtype;
  int64_t ys[4];
  int vec_store;  // y is channels-fastest and 16-byte friendly
  const float* scale;
  const float* noise;
  int64_t noise_stride_n;
  const float* bias;
  int act;
  float alpha, act_gain, clamp;
};

template <typename T> __device__ __forceinline__ T cvt_out(float v);
template <> __device__ __forceinline__ float cvt_out<float>(float v) { return v; }
template <> __device__ __forceinline__ __half cvt_out<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 cvt_out<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// 32 consecutive output channels (co0..) of one pixel: demodulate, add noise, bias + activation, store.
// s_scale / s_bias point at the shared-memory copies for channel co0 (scale may be NULL = 1, bias NULL = 0).
template <typename T, int ACT>
__device__ __forceinline__ void epilogue_store(const IgemmArgs& a, const uint32_t (&acc)[32], T* __restrict__ ybase,
                                               int co0, int co_end, float noise, const float* __restrict__ s_scale,
                                               const float* __restrict__ s_bias, bool has_scale,
                                               unsigned char* __restrict__ stg_row, int stg_xor, int c_tile) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    float r = __uint_as_float(acc[j]);
    if (has_scale) r *= s_scale[j];
    r += noise;
    if (ACT != -1) {
      r += s_bias[j];
      r = act_eval<float, (ACT > 0 ? ACT : 0)>(r, 0.f, 0.f, 1.f, a.act, 0, a.alpha, a.act_gain, a.clamp);
    }
    v[j] = r;
  }
  if (stg_row != nullptr) {
    if constexpr (sizeof(T) == 2) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        Vec16<T> o;
#pragma unroll
        for (int k = 0; k < 8; ++k) o.v[k] = cvt_out<T>(v[j + k]);
        const int ch = c_tile + j;  // channel offset inside the tile: block ch / 64, 16-byte chunk (ch % 64) / 8
        *reinterpret_cast<Vec16<T>*>(stg_row + (size_t)(ch >> 6) * a.stage_blk_bytes +
                                     ((uint32_t)(((ch & 63) >> 3) ^ stg_xor) << 4)) = o;
      }
    }
  } else if (a.vec_store && co0 + 32 <= co_end) {
    constexpr int V = 16 / sizeof(T);
#pragma unroll
    for (int j = 0; j < 32; j += V) {
      Vec16<T> o;
#pragma unroll
      for (int k = 0; k < V; ++k) o.v[k] = cvt_out<T>(v[j + k]);
      *reinterpret_cast<Vec16<T>*>(ybase + (co0 + j)) = o;
    }
  } else {
    const int64_t cs = a.ys[1];
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (co0 + j < co_end) ybase[(int64_t)(co0 + j) * cs] = cvt_out<T>(v[j]);
  }
}

template <typename T>
__device__ __forceinline__ void epilogue_dispatch(const IgemmArgs& a, const uint32_t (&acc)[32], T* ybase, int co0,
                                                  int co_end, float noise, const float* s_scale, const float* s_bias,
                                                  bool has_scale, unsigned char* stg_row, int stg_xor, int c_tile) {
  if (a.act == 0) epilogue_store<T, -1>(a, acc, ybase, co0, co_end, noise, s_scale, s_bias, has_scale, stg_row, stg_xor, c_tile);
  else if (a.act == VCN_ACT_LRELU) epilogue_store<T, VCN_ACT_LRELU>(a, acc, ybase, co0, co_end, noise, s_scale, s_bias, has_scale, stg_row, stg_xor, c_tile);
  else if (a.act == VCN_ACT_LINEAR) epilogue_store<T, VCN_ACT_LINEAR>(a, acc, ybase, co0, co_end, noise, s_scale, s_bias, has_scale, stg_row, stg_xor, c_tile);
  else epilogue_store<T, 0>(a, acc, ybase, co0, co_end, noise, s_scale, s_bias, has_scale, stg_row, stg_xor, c_tile);
}

// Profiling aid (VCN_CONV_PROFILE=1): cycle counters summed over CTAs, read back with vcn_conv_profile_counters().
//  [0] epilogue busy  [1] epilogue waiting for an accumulator  [2] MMA warp waiting for a drained accumulator
//  [3] MMA warp waiting for operands  [4] tiles  [5] TMA producer waiting for a free stage  [6] k-blocks
#if VCN_CONV_PROF
__device__ unsigned long long g_conv_prof[8];
#endif

struct alignas(64) OutMaps {
  CUtensorMap m[kMaxPhases];  // per output phase: the strided view y[n, g_oy::o_mul, g_ox::o_mul, :]
};

struct TileCoord {
  int x0, y0, n0, co0, phase;
};

template <bool PAIR>
__device__ __forceinline__ TileCoord decode_tile(const IgemmArgs& a, int tile, int rank) {
  // m-tiles fastest: consecutive CTAs share the weight tile (L2 reuse of B), then phases, then out-channel tiles.
  // pair mode: `tile` indexes pair tiles (two half tiles along pair_dim); this CTA owns half `rank`.
  int mt, r;
  TileCoord t;
  if (a.num_phases > 1) {
    // transposed conv: the 4 output phases of one pixel tile run on neighbouring CTAs at the same time - they read the same
    // input tile (one DRAM fetch instead of up to four) and interleave their writes into the same output rows (a phase
    // writes every other pixel: apart in time, every 32-byte sector of an fp32 / NCHW output was read-modify-written -
    // ncu: 685 MB written + 807 MB read for 270 MB of output).  The phase is rotated by the group index: with a static
    // walk tile = cta + k * grid and grid % 4 == 0 a CTA would otherwise keep the 4-tap (or the 1-tap) phase for good.
    const int per_co = a.tiles_m * a.num_phases;
    const int rem = tile % per_co;
    mt = rem / a.num_phases;
    t.phase = (rem % a.num_phases + mt) % a.num_phases;
    t.co0 = (tile / per_co) * a.n_tile;
  } else {
    mt = tile % a.tiles_m;
    r = tile / a.tiles_m;
    t.phase = 0;
    t.co0 = r * a.n_tile;
  }
  const int wi = mt % a.tiles_w; mt /= a.tiles_w;
  const int hi = mt % a.tiles_h;
  const int ni = mt / a.tiles_h;
  const int sw = (PAIR && a.pair_dim == 0) ? 1 : 0, sh = (PAIR && a.pair_dim == 1) ? 1 : 0, sn = (PAIR && a.pair_dim == 2) ? 1 : 0;
  t.x0 = ((wi << sw) + (sw ? rank : 0)) * a.tw;
  t.y0 = ((hi << sh) + (sh ? rank : 0)) * a.th;
  t.n0 = ((ni << sn) + (sn ? rank : 0)) * a.tn;
  return t;
}

// Swapped-operand epilogue: this thread owns ONE output channel (TMEM lane) and 32 consecutive tile pixels (columns
// m0..m0+31).  The 32 lanes of the warp hold 32 consecutive channels of the same pixel, so each store instruction of
// the warp writes one contiguous run of channels (y is channels-fastest).  Pixel coordinates are warp-uniform.
template <typename T, int ACT>
__device__ __forceinline__ void swap_store(const IgemmArgs& a, const uint32_t (&acc)[32], T* __restrict__ ybase,
                                           const float* __restrict__ noise_n, int m0, int x0, int y0, int g_oy, int g_ox,
                                           bool co_ok, float sc, float bi) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const int m = m0 + j;
    const int px = x0 + (m & (a.tw - 1)), py = y0 + ((m >> a.log_tw) & (a.th - 1));
    const int oy = py * a.o_mul + g_oy, ox = px * a.o_mul + g_ox;
    if (!(py < a.grid_h && px < a.grid_w && oy < a.out_h && ox < a.out_w)) continue;  // warp-uniform
    float r = __uint_as_float(acc[j]) * sc;
    if (noise_n) r += __ldg(noise_n + (int64_t)oy * a.out_w + ox);
    if (ACT != -1) r = act_eval<float, (ACT > 0 ? ACT : 0)>(r + bi, 0.f, 0.f, 1.f, a.act, 0, a.alpha, a.act_gain, a.clamp);
    if (co_ok) ybase[(int64_t)oy * a.ys[2] + (int64_t)ox * a.ys[3]] = cvt_out<T>(r);
  }
}

// Fast variant for tiles at least 32 pixels wide: the 32 columns are consecutive x of ONE output row, so the row
// pointer, the bounds (nvalid leading pixels exist) and the noise (lane j holds pixel j's value) are hoisted out.
template <typename T, int ACT>
__device__ __forceinline__ void swap_store_row(const IgemmArgs& a, const uint32_t (&acc)[32], T* __restrict__ yrow,
                                               int64_t xstride, int nvalid, float nz_lane, bool has_noise, bool co_ok,
                                               float sc, float bi) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    float r = __uint_as_float(acc[j]) * sc;
    if (has_noise) r += __shfl_sync(0xffffffffu, nz_lane, j);
    if (ACT != -1) r = act_eval<float, (ACT > 0 ? ACT : 0)>(r + bi, 0.f, 0.f, 1.f, a.act, 0, a.alpha, a.act_gain, a.clamp);
    if (co_ok && j < nvalid) yrow[j * xstride] = cvt_out<T>(r);
  }
}

template <typename T>
__device__ __forceinline__ void swap_row_dispatch(const IgemmArgs& a, const uint32_t (&acc)[32], T* yrow, int64_t xstride,
                                                  int nvalid, float nz_lane, bool has_noise, bool co_ok, float sc, float bi) {
  if (a.act == 0) swap_store_row<T, -1>(a, acc, yrow, xstride, nvalid, nz_lane, has_noise, co_ok, sc, bi);
  else if (a.act == VCN_ACT_LRELU) swap_store_row<T, VCN_ACT_LRELU>(a, acc, yrow, xstride, nvalid, nz_lane, has_noise, co_ok, sc, bi);
  else if (a.act == VCN_ACT_LINEAR) swap_store_row<T, VCN_ACT_LINEAR>(a, acc, yrow, xstride, nvalid, nz_lane, has_noise, co_ok, sc, bi);
  else swap_store_row<T, 0>(a, acc, yrow, xstride, nvalid, nz_lane, has_noise, co_ok, sc, bi);
}

// Swapped operands, staged output: this thread's channel goes to byte (cl & 7) * 2 of 16-byte chunk (cl >> 3) ^ (row & 7)
// of staging row `row` (= tile pixel), cl = channel inside its 64-channel block.  row0 is a multiple of 32.
template <typename T, int ACT>
__device__ __forceinline__ void swap_stage(const IgemmArgs& a, const uint32_t (&acc)[32], unsigned char* __restrict__ blk,
                                           int cl, int row0, float nz_lane, bool has_noise, float sc, float bi) {
  if constexpr (sizeof(T) == 2) {
    unsigned char* base = blk + (size_t)row0 * 128 + (cl & 7) * 2;
    const int chunk = cl >> 3;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      float r = __uint_as_float(acc[j]) * sc;
      if (has_noise) r += __shfl_sync(0xffffffffu, nz_lane, j);
      if (ACT != -1) r = act_eval<float, (ACT > 0 ? ACT : 0)>(r + bi, 0.f, 0.f, 1.f, a.act, 0, a.alpha, a.act_gain, a.clamp);
      *reinterpret_cast<T*>(base + j * 128 + ((chunk ^ (j & 7)) << 4)) = cvt_out<T>(r);
    }
  }
}

template <typename T>
__device__ __forceinline__ void swap_stage_dispatch(const IgemmArgs& a, const uint32_t (&acc)[32], unsigned char* blk, int cl,
                                                    int row0, float nz_lane, bool has_noise, float sc, float bi) {
  if (a.act == 0) swap_stage<T, -1>(a, acc, blk, cl, row0, nz_lane, has_noise, sc, bi);
  else if (a.act == VCN_ACT_LRELU) swap_stage<T, VCN_ACT_LRELU>(a, acc, blk, cl, row0, nz_lane, has_noise, sc, bi);
  else if (a.act == VCN_ACT_LINEAR) swap_stage<T, VCN_ACT_LINEAR>(a, acc, blk, cl, row0, nz_lane, has_noise, sc, bi);
  else swap_stage<T, 0>(a, acc, blk, cl, row0, nz_lane, has_noise, sc, bi);
}

template <typename T>
__device__ __forceinline__ void swap_dispatch(const IgemmArgs& a, const uint32_t (&acc)[32], T* ybase, const float* noise_n,
                                              int m0, int x0, int y0, int g_oy, int g_ox, bool co_ok, float sc, float bi) {
  if (a.act == 0) swap_store<T, -1>(a, acc, ybase, noise_n, m0, x0, y0, g_oy, g_ox, co_ok, sc, bi);
  else if (a.act == VCN_ACT_LRELU) swap_store<T, VCN_ACT_LRELU>(a, acc, ybase, noise_n, m0, x0, y0, g_oy, g_ox, co_ok, sc, bi);
  else if (a.act == VCN_ACT_LINEAR) swap_store<T, VCN_ACT_LINEAR>(a, acc, ybase, noise_n, m0, x0, y0, g_oy, g_ox, co_ok, sc, bi);
  else swap_store<T, 0>(a, acc, ybase, noise_n, m0, x0, y0, g_oy, g_ox, co_ok, sc, bi);
}

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// PAIR = false: one CTA per tile, tcgen05 cta_group::1 only.  PAIR = true: 2-CTA clusters, cta_group::2 only.  Two
// instantiations because the driver refuses to launch a kernel that contains cta_group::2 instructions without a cluster
// (cudaErrorInvalidClusterSize), whichever path would run.
template <bool PAIR>
__global__ void __launch_bounds__(kThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ OutMaps om, const __grid_constant__ IgemmArgs a) {
  extern __shared__ unsigned char smem_raw[];
  // 1024-byte alignment: the 128B swizzle pattern repeats every 8 rows = 1024 bytes
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t stage_bytes = a.pix_bytes + a.w_bytes;
  unsigned char* stg = smem + (size_t)a.stages * stage_bytes;  // output staging tile (1024-byte aligned), may be empty
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stg + a.stage_bytes_out);
  uint64_t* empty_bar = full_bar + a.stages;
  uint64_t* tmem_full_bar = empty_bar + a.stages;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;     // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);
  float* s_par = reinterpret_cast<float*>(tmem_ptr_smem + 4);  // [2 acc stages][scale n_tile | bias n_tile]

  // the warp index goes through a shuffle so that the compiler KNOWS it is warp-uniform: the role branches below are then
  // uniform control flow, the TMA / MMA operands (descriptors, barrier addresses, coordinates) live in uniform registers and
  // the one-lane issue is an `elect.sync` predicate - with a divergent `lane == 0` branch every UTMALDG / UTCHMMA sat in
  // its own ELECT / R2UR waterfall loop (~20-40 issue slots per MMA on the single issuing thread)
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int rank = PAIR ? __shfl_sync(0xffffffffu, (int)cluster_ctarank(), 0) : 0;  // 0 = leader of the CTA pair (issues the MMAs)
  const int tile0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;   // persistent walk over (pair) tiles
  const int tile_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < a.stages; ++s) {
      mbar_init(full_bar + s, 1);   // pair mode: only the leader's is used (both producers' TMA bytes land on it)
      mbar_init(empty_bar + s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tmem_full_bar + s, 1);
      mbar_init(tmem_empty_bar + s, PAIR ? 2 * kEpilogueWarps : kEpilogueWarps);  // pair: the peer's warps arrive remotely
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    if constexpr (PAIR) tmem_alloc_2cta(tmem_ptr_smem, (uint32_t)a.tmem_cols);
    else tmem_alloc(tmem_ptr_smem, (uint32_t)a.tmem_cols);
  }
  tcgen05_fence_before();
  if constexpr (PAIR) cluster_sync_all();  // barrier inits + the pair's TMEM allocation visible to both CTAs
  else __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===== TMA producer (whole warp walks the loop, one elected lane issues) =====
    {
      int s = 0;
      uint32_t ph = 0;
      long long prof_wait = 0;
      const int co_half = PAIR ? rank * (a.n_tile >> 1) : 0;  // pair: this CTA stages its half of the weight tile
      for (int tile = tile0; tile < a.tiles_total; tile += tile_step) {
        const TileCoord tc = decode_tile<PAIR>(a, tile, rank);
        const int nw = a.w_per_sample ? tc.n0 : 0;
        for (int t = a.phase_first[tc.phase]; t < a.phase_first[tc.phase + 1]; ++t) {
          const ConvTap tap = a.taps[t];
          const int ax = tc.x0 * a.a_mul + tap.dx, ay = tc.y0 * a.a_mul + tap.dy;
          for (int cb = 0; cb < a.kblocks_c; ++cb) {
            const long long t0 = (kProf && a.prof) ? clock64() : 0;
            mbar_wait(empty_bar + s, ph ^ 1u);
            if (kProf && a.prof) prof_wait += clock64() - t0;
            unsigned char* sa = smem + (size_t)s * stage_bytes;
            if (elect_one()) {
              if constexpr (!PAIR) {
                mbar_expect_tx(full_bar + s, stage_bytes);
                tma_load_4d(sa, &tmap_a, full_bar + s, cb * a.kbk, ax, ay, tc.n0);
                tma_load_4d(sa + a.pix_bytes, &tmap_b, full_bar + s, cb * a.kbk, tc.co0, tap.wtap, nw);
              } else {
                // both CTAs load into their OWN shared memory and complete the bytes on the LEADER's barrier, which the
                // leader arms for the whole pair (a complete_tx that overtakes the expect_tx only drives the count negative)
                if (rank == 0) mbar_expect_tx(full_bar + s, 2u * stage_bytes);
                const uint32_t lead_bar = mapa_u32(full_bar + s, 0);
                tma_load_4d_2cta(sa, &tmap_a, lead_bar, cb * a.kbk, ax, ay, tc.n0);
                tma_load_4d_2cta(sa + a.pix_bytes, &tmap_b, lead_bar, cb * a.kbk, tc.co0 + co_half, tap.wtap, nw);
              }
            }
            __syncwarp();
            if (++s == a.stages) { s = 0; ph ^= 1u; }
          }
        }
      }
#if VCN_CONV_PROF
      if (a.prof && lane == 0) atomicAdd(&g_conv_prof[5], (unsigned long long)prof_wait);
#endif
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (rank == 0) {  // pair mode: the leader issues for both CTAs; the whole warp walks the loop, one elected lane issues
      const int acc_cols = a.swap ? 256 : a.n_tile;
      const uint32_t idesc = a.tf32 ? umma_idesc_tf32_f32(PAIR ? 2 * kBlockM : kBlockM, acc_cols)
                                    : umma_idesc_bf16_f32(PAIR ? 2 * kBlockM : kBlockM, acc_cols);
      int s = 0;
      uint32_t ph = 0;
      int tile_iter = 0;
      long long prof_acc = 0, prof_ops = 0, prof_kb = 0;
      for (int tile = tile0; tile < a.tiles_total; tile += tile_step, ++tile_iter) {
        const TileCoord tc = decode_tile<PAIR>(a, tile, rank);
        const int num_kb = (a.phase_first[tc.phase + 1] - a.phase_first[tc.phase]) * a.kblocks_c;
        const int as = tile_iter & 1;
        const uint32_t aph = (uint32_t)(tile_iter >> 1) & 1u;
        long long t0 = (kProf && a.prof) ? clock64() : 0;
        mbar_wait(tmem_empty_bar + as, aph ^ 1u);  // epilogue (of both CTAs in pair mode) has drained this accumulator stage
        if (kProf && a.prof) prof_acc += clock64() - t0;
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * acc_cols);
        for (int kb = 0; kb < num_kb; ++kb) {
          if (kProf && a.prof) t0 = clock64();
          mbar_wait(full_bar + s, ph);
          if (kProf && a.prof) { prof_ops += clock64() - t0; ++prof_kb; }
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)s * stage_bytes);
          const uint64_t dpix = umma_desc_k_sw128(sa), dwt = umma_desc_k_sw128(sa + a.pix_bytes);
          const uint64_t da = a.swap ? dwt : dpix, db = a.swap ? dpix : dwt;
          if (elect_one()) {
            // advancing K inside the swizzle span = +32 bytes on the start address (>>4 encoded)
            // one MMA = 32 bytes of K per row: 16 bf16 or 8 tf32 channels
            if (a.tf32) {
#pragma unroll
              for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                const uint32_t acc = (kb > 0 || k > 0) ? 1u : 0u;
                if constexpr (PAIR) umma_tf32_ss_2cta(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, acc);
                else umma_tf32_ss(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, acc);
              }
            } else {
#pragma unroll
              for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                const uint32_t acc = (kb > 0 || k > 0) ? 1u : 0u;
                if constexpr (PAIR) umma_bf16_ss_2cta(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, acc);
                else umma_bf16_ss(d_tmem, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, acc);
              }
            }
            if constexpr (PAIR) umma_commit_2cta(empty_bar + s, 3);  // frees stage s in BOTH CTAs
            else umma_commit(empty_bar + s);
            if (kb == num_kb - 1) {  // accumulator ready (pair: both CTAs' epilogues)
              if constexpr (PAIR) umma_commit_2cta(tmem_full_bar + as, 3);
              else umma_commit(tmem_full_bar + as);
            }
          }
          __syncwarp();
          if (++s == a.stages) { s = 0; ph ^= 1u; }
        }
      }
#if VCN_CONV_PROF
      if (a.prof && lane == 0) {
        atomicAdd(&g_conv_prof[2], (unsigned long long)prof_acc);
        atomicAdd(&g_conv_prof[3], (unsigned long long)prof_ops);
        atomicAdd(&g_conv_prof[6], (unsigned long long)prof_kb);
      }
#endif
    }
  } else {
    // ===== epilogue: warps 2..9, two per TMEM lane quarter (warp % 4), alternating 32-column chunks =====
    const int q = warp & 3;
    const int et = threadIdx.x - 64;  // 0..kEpilogueWarps*32-1
    const int cg0 = ((warp - 2) >> 2) * 32, cg_step = (kEpilogueWarps / 4) * 32;  // this warp's 32-column chunks: cg0, cg0 + cg_step, ...
    const int m = q * 32 + lane;
    const int tw_i = m % a.tw;
    const int th_i = (m / a.tw) % a.th;
    const int tn_i = m / (a.tw * a.th);
    const bool stage_scale = a.scale != nullptr && a.tn == 1;
    int tile_iter = 0;
    long long prof_busy = 0, prof_wait = 0;
    if (a.swap) {
      // lane = out channel, columns = the 256 pixels of the tile (one sample per tile: tn == 1)
      for (int tile = tile0; tile < a.tiles_total; tile += tile_step, ++tile_iter) {
        const TileCoord tc = decode_tile<PAIR>(a, tile, rank);
        const int as = tile_iter & 1;
        const uint32_t aph = (uint32_t)(tile_iter >> 1) & 1u;
        const int co = tc.co0 + m;
        const bool co_ok = co < a.cout;
        const float sc = (a.scale && co_ok) ? a.scale[(int64_t)tc.n0 * a.cout + co] : 1.f;
        const float bi = (a.bias && co_ok) ? a.bias[co] : 0.f;
        const float* noise_n = a.noise ? a.noise + (int64_t)tc.n0 * a.noise_stride_n : nullptr;
        const int64_t yoff = (int64_t)tc.n0 * a.ys[0] + (co_ok ? co : 0);
        const int g_oy = a.g_oy[tc.phase], g_ox = a.g_ox[tc.phase];
        if (a.stage_out) {  // the previous tile's TMA stores must have read the staging tile before it is rewritten
          if (et == 0) tma_store_wait_read();
          named_bar_sync(1, kEpilogueWarps * 32);
        }
        long long t0 = (kProf && a.prof) ? clock64() : 0;
        mbar_wait(tmem_full_bar + as, aph);
        if (kProf && a.prof) { const long long t1 = clock64(); prof_wait += t1 - t0; t0 = t1; }
        tcgen05_fence_after();
        for (int c = cg0; c < 256; c += cg_step) {
          uint32_t acc[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * 256 + c);
          if (a.stage_out) {
            tmem_ld_32x32(taddr, acc);
            float nz_lane = 0.f;
            if (noise_n) {  // lane j fetches the noise of tile pixel c + j
              const int ml = c + lane;
              const int px = tc.x0 + (ml & (a.tw - 1)), py = tc.y0 + ((ml >> a.log_tw) & (a.th - 1));
              const int oy = py * a.o_mul + g_oy, ox = px * a.o_mul + g_ox;
              if (py < a.grid_h && px < a.grid_w && oy < a.out_h && ox < a.out_w) nz_lane = __ldg(noise_n + (int64_t)oy * a.out_w + ox);
            }
            unsigned char* blk = stg + (size_t)(m >> 6) * a.stage_blk_bytes;
            tmem_ld_wait();
            if (a.y_dtype == VCN_BF16) swap_stage_dispatch<__nv_bfloat16>(a, acc, blk, m & 63, c, nz_lane, noise_n != nullptr, sc, bi);
            else if (a.y_dtype == VCN_F16) swap_stage_dispatch<__half>(a, acc, blk, m & 63, c, nz_lane, noise_n != nullptr, sc, bi);
            continue;
          }
          if (a.tw >= 32) {
            const int px0 = tc.x0 + (c & (a.tw - 1)), py = tc.y0 + ((c >> a.log_tw) & (a.th - 1));
            const int oy = py * a.o_mul + g_oy, ox0 = px0 * a.o_mul + g_ox;
            int nvalid = min(32, min(a.grid_w - px0, (a.out_w - ox0 + a.o_mul - 1) / a.o_mul));
            if (!(py < a.grid_h && oy < a.out_h)) nvalid = 0;
            if (nvalid <= 0) continue;  // warp-uniform
            tmem_ld_32x32(taddr, acc);
            const float nz_lane = (noise_n && lane < nvalid) ? __ldg(noise_n + (int64_t)oy * a.out_w + ox0 + lane * a.o_mul) : 0.f;
            const int64_t roff = yoff + (int64_t)oy * a.ys[2] + (int64_t)ox0 * a.ys[3], xs = (int64_t)a.o_mul * a.ys[3];
            tmem_ld_wait();
            if (a.y_dtype == VCN_F32) swap_row_dispatch<float>(a, acc, static_cast<float*>(a.y) + roff, xs, nvalid, nz_lane, noise_n != nullptr, co_ok, sc, bi);
            else if (a.y_dtype == VCN_BF16) swap_row_dispatch<__nv_bfloat16>(a, acc, static_cast<__nv_bfloat16*>(a.y) + roff, xs, nvalid, nz_lane, noise_n != nullptr, co_ok, sc, bi);
            else if (a.y_dtype == VCN_F16) swap_row_dispatch<__half>(a, acc, static_cast<__half*>(a.y) + roff, xs, nvalid, nz_lane, noise_n != nullptr, co_ok, sc, bi);
            continue;
          }
          tmem_ld_32x32(taddr, acc);
          tmem_ld_wait();
          if (a.y_dtype == VCN_F32) swap_dispatch<float>(a, acc, static_cast<float*>(a.y) + yoff, noise_n, c, tc.x0, tc.y0, g_oy, g_ox, co_ok, sc, bi);
          else if (a.y_dtype == VCN_BF16) swap_dispatch<__nv_bfloat16>(a, acc, static_cast<__nv_bfloat16*>(a.y) + yoff, noise_n, c, tc.x0, tc.y0, g_oy, g_ox, co_ok, sc, bi);
          else swap_dispatch<__half>(a, acc, static_cast<__half*>(a.y) + yoff, noise_n, c, tc.x0, tc.y0, g_oy, g_ox, co_ok, sc, bi);
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tmem_empty_bar + as);
        if (kProf && a.prof) prof_busy += clock64() - t0;
        if (a.stage_out) {
          fence_proxy_async();  // generic-proxy writes of the staging tile -> visible to TMA
          named_bar_sync(2, kEpilogueWarps * 32);
          if (et == 0) {
            for (int b = 0; b < a.out_blocks; ++b)
              if (tc.co0 + 64 * b < a.cout) tma_store_4d(&om.m[tc.phase], stg + (size_t)b * a.stage_blk_bytes, tc.co0 + 64 * b, tc.x0, tc.y0, tc.n0);
            tma_store_commit();
          }
        }
      }
    } else
    for (int tile = tile0; tile < a.tiles_total; tile += tile_step, ++tile_iter) {
      const TileCoord tc = decode_tile<PAIR>(a, tile, rank);
      const int as = tile_iter & 1;
      const uint32_t aph = (uint32_t)(tile_iter >> 1) & 1u;
      float* sp = s_par + as * 2 * a.n_tile;
      // per-tile channel parameters -> shared memory (read while the MMAs of this tile are still running)
      for (int j = et; j < a.n_tile; j += kEpilogueWarps * 32) {
        const int co = tc.co0 + j;
        float sc = 1.f, bi = 0.f;
        if (co < a.cout) {
          if (stage_scale) sc = a.scale[(int64_t)tc.n0 * a.cout + co];
          if (a.bias) bi = a.bias[co];
        }
        sp[j] = sc;
        sp[a.n_tile + j] = bi;
      }
      const int px = tc.x0 + tw_i, py = tc.y0 + th_i, n = tc.n0 + tn_i;
      const int oy = py * a.o_mul + a.g_oy[tc.phase], ox = px * a.o_mul + a.g_ox[tc.phase];
      const bool ok = n < a.n && py < a.grid_h && px < a.grid_w && oy < a.out_h && ox < a.out_w;
      float noise = 0.f;
      if (ok && a.noise) noise = a.noise[(int64_t)n * a.noise_stride_n + (int64_t)oy * a.out_w + ox];
      const int64_t yoff = ok ? (int64_t)n * a.ys[0] + (int64_t)oy * a.ys[2] + (int64_t)ox * a.ys[3] : 0;
      const float* gscale = (a.scale && !stage_scale && ok) ? a.scale + (int64_t)n * a.cout : nullptr;
      const int co_end = min(a.cout, tc.co0 + a.n_tile);  // channels past the tile belong to another CTA
      if (a.stage_out && et == 0) tma_store_wait_read();  // previous tile's TMA stores have read the staging tile
      named_bar_sync(1, kEpilogueWarps * 32);
      long long t0 = (kProf && a.prof) ? clock64() : 0;
      mbar_wait(tmem_full_bar + as, aph);
      if (kProf && a.prof) { const long long t1 = clock64(); prof_wait += t1 - t0; t0 = t1; }
      tcgen05_fence_after();
      for (int c = cg0; c < a.n_tile; c += cg_step) {
        if (tc.co0 + c >= a.cout) break;  // warp-uniform
        uint32_t acc[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * a.n_tile + c);
        if (a.n_tile - c >= 32) {
          tmem_ld_32x32(taddr, acc);
        } else {  // n_tile is a multiple of 16: last half chunk
          uint32_t h16[16];
          tmem_ld_32x16(taddr, h16);
#pragma unroll
          for (int j = 0; j < 16; ++j) { acc[j] = h16[j]; acc[16 + j] = 0u; }
        }
        tmem_ld_wait();
        if (ok || a.stage_out) {  // staged: every row is written, TMA clips the pixels that do not exist
          unsigned char* stg_row = a.stage_out ? stg + (size_t)m * 128 : nullptr;
          float lsc[32];
          const float* scp = sp + c;
          bool has_scale = stage_scale;
          if (gscale) {  // several samples in one tile (tiny resolutions): per-row scale from global memory
#pragma unroll
            for (int j = 0; j < 32; ++j) lsc[j] = (tc.co0 + c + j < a.cout) ? gscale[tc.co0 + c + j] : 1.f;
            scp = lsc;
            has_scale = true;
          }
          const float* bip = sp + a.n_tile + c;
          if (a.y_dtype == VCN_F32) epilogue_dispatch<float>(a, acc, static_cast<float*>(a.y) + yoff, tc.co0 + c, co_end, noise, scp, bip, has_scale, stg_row, m & 7, c);
          else if (a.y_dtype == VCN_BF16) epilogue_dispatch<__nv_bfloat16>(a, acc, static_cast<__nv_bfloat16*>(a.y) + yoff, tc.co0 + c, co_end, noise, scp, bip, has_scale, stg_row, m & 7, c);
          else if (a.y_dtype == VCN_F16) epilogue_dispatch<__half>(a, acc, static_cast<__half*>(a.y) + yoff, tc.co0 + c, co_end, noise, scp, bip, has_scale, stg_row, m & 7, c);
        }
      }
      // hand the accumulator stage back to the MMA warp (pair mode: the leader's barrier, remotely from the peer CTA)
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (PAIR) {
          if (rank != 0) mbar_arrive_cluster(mapa_u32(tmem_empty_bar + as, 0));
          else mbar_arrive(tmem_empty_bar + as);
        } else {
          mbar_arrive(tmem_empty_bar + as);
        }
      }
      if (kProf && a.prof) prof_busy += clock64() - t0;
      if (a.stage_out) {
        fence_proxy_async();
        named_bar_sync(2, kEpilogueWarps * 32);
        if (et == 0) {
          for (int b = 0; b < a.out_blocks; ++b)
            if (tc.co0 + 64 * b < a.cout) tma_store_4d(&om.m[tc.phase], stg + (size_t)b * a.stage_blk_bytes, tc.co0 + 64 * b, tc.x0, tc.y0, tc.n0);
          tma_store_commit();
        }
      }
    }
    if (a.stage_out && et == 0) tma_store_wait_all();
#if VCN_CONV_PROF
    if (a.prof && et == 0) {
      atomicAdd(&g_conv_prof[0], (unsigned long long)prof_busy);
      atomicAdd(&g_conv_prof[1], (unsigned long long)prof_wait);
      atomicAdd(&g_conv_prof[4], (unsigned long long)tile_iter);
    }
#endif
  }
  tcgen05_fence_before();
  if constexpr (PAIR) cluster_sync_all();  // the peer may still be arriving on this CTA's barriers / the leader's MMAs read the peer's smem
  else __syncthreads();
  if (warp == 2) {
    tcgen05_fence_after();
    if constexpr (PAIR) tmem_dealloc_2cta(tmem_base, (uint32_t)a.tmem_cols);
    else tmem_dealloc(tmem_base, (uint32_t)a.tmem_cols);
  }
}

inline int pow2_ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

// (tw, th, tn) with tw*th*tn = pixels (128, or 256 with swapped operands) minimising the number of CTA tiles; ties -> wider rows
void choose_tile(int gw, int gh, int n, int one_sample, int pixels, int max_w, int min_w, int* tw, int* th, int* tn) {
  int64_t best = -1;
  for (int w = pixels < max_w ? pixels : max_w; w >= min_w; w >>= 1) {
    for (int h = 1; w * h <= pixels; h <<= 1) {
      const int nn = pixels / (w * h);
      if (one_sample && nn != 1) continue;  // per-sample weights: a tile must not straddle samples
      const int64_t tiles = (int64_t)((gw + w - 1) / w) * ((gh + h - 1) / h) * ((n + nn - 1) / nn);
      if (best < 0 || tiles < best) { best = tiles; *tw = w; *th = h; *tn = nn; }
    }
  }
}

// CTA pairs: (tw, th, tn) = ONE CTA's 128-pixel half tile and the dimension along which the two halves sit side by side,
// minimising the number of pair tiles (ties -> wider rows, then stacking along h).  Returns the pair-tile count.
int64_t choose_pair_tile(int gw, int gh, int n, int one_sample, int max_w, int* tw, int* th, int* tn, int* pair_dim) {
  int64_t best = -1;
  for (int w = 128 < max_w ? 128 : max_w; w >= 1; w >>= 1) {
    for (int h = 1; w * h <= 128; h <<= 1) {
      const int nn = 128 / (w * h);
      if (one_sample && nn != 1) continue;
      static const int kOrder[3] = {1, 0, 2};  // prefer stacking along h, then w, then n
      for (int oi = 0; oi < 3; ++oi) {
        const int pd = kOrder[oi];
        if (pd == 2 && one_sample) continue;
        if (pd == 0 && 2 * w > max_w) continue;
        const int pw = pd == 0 ? 2 * w : w, phh = pd == 1 ? 2 * h : h, pn = pd == 2 ? 2 * nn : nn;
        const int64_t tiles = (int64_t)((gw + pw - 1) / pw) * ((gh + phh - 1) / phh) * ((n + pn - 1) / pn);
        if (best < 0 || tiles < best) { best = tiles; *tw = w; *th = h; *tn = nn; *pair_dim = pd; }
      }
    }
  }
  return best;
}

// environment switches (experiments), read ONCE per process - never on the launch path
struct ConvEnv {
  int swap = -1;   // VCN_CONV_SWAP: 0 never, 1 not for the transposed conv, unset = policy
  int stage = -1;  // VCN_CONV_STAGE: 0 = no TMA-store epilogue
  int pair = -1;   // VCN_CONV_PAIR: 0 never, 2 = every tile of >= 64 out channels, unset = policy (n_tile == 256)
  int prof = 0;    // VCN_CONV_PROFILE (only meaningful in -DVCN_CONV_PROF=1 builds)
  ConvEnv() {
    if (const char* e = getenv("VCN_CONV_SWAP")) swap = atoi(e);
    if (const char* e = getenv("VCN_CONV_STAGE")) stage = atoi(e);
    if (const char* e = getenv("VCN_CONV_PAIR")) pair = atoi(e);
    if (const char* e = getenv("VCN_CONV_PROFILE")) prof = atoi(e);
  }
};
const ConvEnv& conv_env() {
  static const ConvEnv env;
  return env;
}

// SMs this kernel leaves free (vcn_set_conv_sm_reserve): the persistent grid takes one CTA per SM with a STATIC tile walk, so
// a CTA that cannot become resident - a collective's CTAs own its SM - holds back its whole share of the tiles until the
// collective ends.  Data-parallel runs reserve as many SMs as the communication kernels occupy.
int g_sm_reserve = 0;

}  // namespace
}  // namespace vcn

extern "C" int vcn_set_conv_sm_reserve(int sms) {
  using namespace vcn;
  VCN_REQUIRE(sms >= 0 && sms <= 64, "set_conv_sm_reserve: 0..64 SMs (got %d)", sms);
  g_sm_reserve = sms;
  return VCN_OK;
}

extern "C" int vcn_conv_igemm(const void* xhat, const void* wt, void* y, const vcn_conv_desc* d,
                              const vcn_upfirdn2d_epilogue* epi, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(d != nullptr, "conv_igemm: desc must not be NULL");
  VCN_REQUIRE(d->n >= 0 && d->h >= 1 && d->w >= 1 && d->cout >= 1, "conv_igemm: bad shape");
  VCN_REQUIRE(d->cin_pad >= 8 && d->cin_pad % 8 == 0, "conv_igemm: cin_pad (%d) must be a positive multiple of 8", d->cin_pad);
  VCN_REQUIRE((d->kh == 1 && d->kw == 1) || (d->kh == 3 && d->kw == 3), "conv_igemm: only 1x1 and 3x3 kernels (got %dx%d)",
              d->kh, d->kw);
  VCN_REQUIRE(d->y_dtype == VCN_F32 || d->y_dtype == VCN_BF16 || d->y_dtype == VCN_F16, "conv_igemm: bad output dtype %d",
              d->y_dtype);
  IgemmArgs a;
  memset(&a, 0, sizeof(a));
  const int kh = d->kh, kw = d->kw;
  int num_taps = 0;
  if (d->kind == VCN_CONV_FPROP || d->kind == VCN_CONV_DGRAD) {
    VCN_REQUIRE(d->out_h == d->h && d->out_w == d->w, "conv_igemm: stride-1 output must be %dx%d", d->h, d->w);
    for (int ky = 0; ky < kh; ++ky)
      for (int kx = 0; kx < kw; ++kx) {
        ConvTap& t = a.taps[num_taps++];
        if (d->kind == VCN_CONV_FPROP) { t.dy = ky - kh / 2; t.dx = kx - kw / 2; }
        else { t.dy = kh / 2 - ky; t.dx = kw / 2 - kx; }
        t.wtap = ky * kw + kx;
      }
    a.num_phases = 1; a.phase_first[1] = num_taps; a.o_mul = 1; a.grid_h = d->h; a.grid_w = d->w;
  } else if (d->kind == VCN_CONV_SCONV2) {
    VCN_REQUIRE(kh == 3 && kw == 3, "conv_igemm: stride-2 conv needs a 3x3 kernel");
    VCN_REQUIRE(d->h >= 3 && d->w >= 3 && d->h % 2 == 1 && d->w % 2 == 1 && d->out_h == (d->h - 1) / 2 && d->out_w == (d->w - 1) / 2,
                "conv_igemm: stride-2 conv maps (2h+1, 2w+1) to (h, w); got %dx%d -> %dx%d", d->h, d->w, d->out_h, d->out_w);
    for (int ky = 0; ky < 3; ++ky)
      for (int kx = 0; kx < 3; ++kx) {
        ConvTap& t = a.taps[num_taps++];
        t.dy = ky; t.dx = kx; t.wtap = ky * 3 + kx;
      }
    a.num_phases = 1; a.phase_first[1] = num_taps; a.o_mul = 1; a.grid_h = d->out_h; a.grid_w = d->out_w;
  } else if (d->kind == VCN_CONV_TCONV2) {
    VCN_REQUIRE(kh == 3 && kw == 3, "conv_igemm: transposed conv needs a 3x3 kernel");
    VCN_REQUIRE(d->out_h == 2 * d->h + 1 && d->out_w == 2 * d->w + 1, "conv_igemm: transposed output must be %dx%d",
                2 * d->h + 1, 2 * d->w + 1);
    // output phase (a,b) = (oy&1, ox&1) collects the taps with ky%2==a, kx%2==b; input shift -(k>>1)
    for (int g = 0; g < 4; ++g) {
      const int pa = g >> 1, pb = g & 1;
      a.g_oy[g] = pa; a.g_ox[g] = pb;
      a.phase_first[g] = num_taps;
      for (int ky = pa; ky < 3; ky += 2)
        for (int kx = pb; kx < 3; kx += 2) {
          ConvTap& t = a.taps[num_taps++];
          t.dy = -(ky >> 1); t.dx = -(kx >> 1); t.wtap = ky * 3 + kx;
        }
    }
    a.phase_first[4] = num_taps;
    a.num_phases = 4; a.o_mul = 2; a.grid_h = d->h + 1; a.grid_w = d->w + 1;
  } else {
    set_error("conv_igemm: unknown kind %d", d->kind);
    return VCN_ERR_INVALID;
  }
  if (d->n == 0) return VCN_OK;
  VCN_REQUIRE(xhat && wt && y, "conv_igemm: xhat, wt and y must not be NULL");
  VCN_REQUIRE(aligned16(xhat) && aligned16(wt), "conv_igemm: xhat and wt must be 16-byte aligned");

  a.a_mul = d->kind == VCN_CONV_SCONV2 ? 2 : 1;
  a.tf32 = d->tf32 ? 1 : 0;
  a.kbk = a.tf32 ? 32 : kBlockK;  // either way a k-block is one 128-byte swizzle span per row
  const int op_dtype = a.tf32 ? VCN_F32 : VCN_BF16;
  const uint64_t op_size = a.tf32 ? 4 : 2;
  a.w_per_sample = d->w_per_sample ? 1 : 0;
  a.n = d->n; a.cout = d->cout; a.out_h = d->out_h; a.out_w = d->out_w;
  // Operand swap (UMMA M = out channels, N = 256 pixels) for 65..128 out channels and a channels-fastest output: a
  // 128 x 128 tile pulls 1/64 byte of operands per flop through TMA, a 128 x 256 tile 1/85 - and the measured ceiling
  // of this kernel is the ~50 B/clk an SM ingests from L2, not the tensor pipe.
  // Also for the transposed conv (256 -> 128 at 128^2: 0.104 ms swapped vs 0.110 ms; profiles/r2i_rooflines_conv_policy.txt) -
  // with the waterfall issue loops of round 1 the swapped form lost there (0.168 vs 0.131 ms).
  const ConvEnv& env = conv_env();
  a.swap = d->cout > 64 && d->cout <= 128 && d->ys[1] == 1 && (int64_t)a.grid_w * a.grid_h >= 256;
  if (env.swap >= 0) {  // 0: never, 1: not for the transposed conv (experiments)
    a.swap = a.swap && env.swap != 0 && !(env.swap == 1 && d->kind == VCN_CONV_TCONV2);
  }
  // two accumulator stages of n_tile fp32 columns each must fit the 512 TMEM columns
  a.n_tile = (d->cout + 15) / 16 * 16;
  if (a.n_tile > 256) a.n_tile = 256;
  if (a.swap) a.n_tile = 128;  // out channels per tile = the 128 UMMA rows (TMA zero-fills rows past cout)
  // CTA pairs (cta_group::2): worth it when the weight tile is wide - with N = 256 each SM stages 16 + 16 KB per k-block
  // instead of 16 + 32 - and when there are enough pair tiles to fill the machine
  {
    int ptw = 0, pth = 0, ptn = 0, pd = 0;
    const int64_t ptiles = choose_pair_tile(a.grid_w, a.grid_h, d->n, a.w_per_sample, 256 / a.a_mul, &ptw, &pth, &ptn, &pd) *
                           a.num_phases * ((d->cout + a.n_tile - 1) / a.n_tile);
    // measured (B200, L2 flushed; profiles/README.md): FPROP / DGRAD 256->256 and 512->512 gain 5-8 % as pairs; the transposed
    // conv's short-K phase tiles are epilogue-bound and lose (512->256: 0.52 vs 0.56 of peak), so it stays on single CTAs
    bool want = !a.swap && a.n_tile == 256 && d->kind != VCN_CONV_TCONV2;
    if (env.pair == 0) want = false;
    if (env.pair == 2) want = !a.swap && a.n_tile >= 64;
    if (want && a.n_tile % 32 == 0 && ptiles >= sm_count() / 2) {
      a.pair = 1; a.pair_dim = pd; a.tw = ptw; a.th = pth; a.tn = ptn;
    }
  }
  if (!a.pair)
    choose_tile(a.grid_w, a.grid_h, d->n, a.w_per_sample || a.swap, a.swap ? 256 : 128, 256 / a.a_mul,
                (a.swap && a.grid_w > 16) ? 32 : 1 /* swapped epilogue: 32 consecutive x per TMEM chunk */, &a.tw, &a.th, &a.tn);
  for (a.log_tw = 0; (1 << a.log_tw) < a.tw; ++a.log_tw) {}
  for (a.log_th = 0; (1 << a.log_th) < a.th; ++a.log_th) {}
  {
    const int pw = a.tw << (a.pair && a.pair_dim == 0), phh = a.th << (a.pair && a.pair_dim == 1), pn = a.tn << (a.pair && a.pair_dim == 2);
    a.tiles_w = (a.grid_w + pw - 1) / pw;
    a.tiles_h = (a.grid_h + phh - 1) / phh;
    a.tiles_m = a.tiles_w * a.tiles_h * ((d->n + pn - 1) / pn);
  }
  // few M tiles (tiny resolutions): narrower out-channel tiles spread the weight stream over more SMs - one CTA
  // ingests only ~75 GB/s through TMA, so a 4.7 MB weight tensor on 2 CTAs is latency/bandwidth bound
  while (!a.swap && !a.pair && a.n_tile > 32 && a.n_tile % 32 == 0 &&
         (int64_t)a.tiles_m * a.num_phases * ((d->cout + a.n_tile - 1) / a.n_tile) < sm_count() / 2)
    a.n_tile /= 2;
  const int tiles_co = (d->cout + a.n_tile - 1) / a.n_tile;
  const int64_t total = (int64_t)a.tiles_m * a.num_phases * tiles_co;
  VCN_REQUIRE(total <= 0x7fffffffLL, "conv_igemm: too many tiles");
  a.tiles_total = (int)total;
  a.kblocks_c = (d->cin_pad + a.kbk - 1) / a.kbk;
  a.tmem_cols = a.swap ? 512 : pow2_ceil(2 * a.n_tile);
  if (a.tmem_cols < 32) a.tmem_cols = 32;
  a.pix_bytes = (a.swap ? 2u : 1u) * kABytes;
  a.w_bytes = (uint32_t)(a.pair ? a.n_tile / 2 : a.n_tile) * kBlockK * 2;  // pair: each CTA stages half of the weight tile
  const size_t stage_bytes = (size_t)a.pix_bytes + a.w_bytes;
  const size_t tail_bytes = (2 * (size_t)8 + 4) * 8 + 16 + 4 * (size_t)a.n_tile * sizeof(float) + 64;

  a.y = y; a.y_dtype = d->y_dtype;
  for (int i = 0; i < 4; ++i) a.ys[i] = d->ys[i];
  const size_t oe = dtype_size(d->y_dtype);
  const int64_t vq = 16 / (int64_t)oe;
  a.vec_store = d->ys[1] == 1 && aligned16(y) && d->ys[0] % vq == 0 && d->ys[2] % vq == 0 && d->ys[3] % vq == 0;
  // staged TMA-store epilogue: 16-bit channels-fastest output whose 64-channel blocks fill whole staging rows
  a.stage_out = a.vec_store && oe == 2 && (a.swap || a.n_tile % 64 == 0);
  if (env.stage >= 0) a.stage_out = a.stage_out && env.stage != 0;
  if (a.stage_out) {
    a.out_blocks = a.n_tile / 64;
    a.stage_blk_bytes = (a.swap ? 256u : 128u) * 128u;
    a.stage_bytes_out = (uint32_t)a.out_blocks * a.stage_blk_bytes;
  }
  a.stages = (int)((232448 - 1024 - (int64_t)tail_bytes - (int64_t)a.stage_bytes_out) / (int64_t)stage_bytes);
  if (a.stages > 8) a.stages = 8;
  if (a.stages < 1) a.stages = 1;
  // one CTA per SM (launch bounds); TMEM allocations above 256 columns additionally rely on it
  size_t smem = 1024 + (size_t)a.stages * stage_bytes + a.stage_bytes_out + tail_bytes;
  if (smem < 120 * 1024) smem = 120 * 1024;
  if (epi) {
    a.scale = epi->scale; a.noise = epi->noise; a.noise_stride_n = epi->noise_stride_n; a.bias = epi->bias;
    a.act = epi->act; a.alpha = epi->alpha; a.act_gain = epi->act_gain; a.clamp = epi->clamp;
    VCN_REQUIRE(a.act == 0 || (a.act >= VCN_ACT_LINEAR && a.act <= VCN_ACT_SWISH), "conv_igemm: bad epilogue act %d", a.act);
  }

  a.prof = kProf && env.prof != 0;
  CUtensorMap tmap_a, tmap_b;
  {
    const uint64_t dims[4] = {(uint64_t)d->cin_pad, (uint64_t)d->w, (uint64_t)d->h, (uint64_t)d->n};
    const uint64_t strides[3] = {(uint64_t)d->cin_pad * op_size, (uint64_t)d->w * d->cin_pad * op_size,
                                 (uint64_t)d->h * d->w * d->cin_pad * op_size};
    // stride-2 gather: traversal strides (1,2,2,1); the box spans 2*TW x 2*TH input pixels and delivers TW x TH
    const uint32_t m = (uint32_t)a.a_mul;
    const uint32_t box[4] = {(uint32_t)a.kbk, (uint32_t)a.tw * m, (uint32_t)a.th * m, (uint32_t)a.tn};
    const uint32_t estr[4] = {1u, m, m, 1u};
    const int rc = make_tensor_map(&tmap_a, op_dtype, 4, xhat, dims, strides, box, estr, 1);
    if (rc) return rc;
  }
  {
    const uint64_t taps = (uint64_t)(kh * kw);
    const uint64_t dims[4] = {(uint64_t)d->cin_pad, (uint64_t)d->cout, taps, (uint64_t)(a.w_per_sample ? d->n : 1)};
    const uint64_t strides[3] = {(uint64_t)d->cin_pad * op_size, (uint64_t)d->cout * d->cin_pad * op_size,
                                 taps * d->cout * d->cin_pad * op_size};
    const uint32_t box[4] = {(uint32_t)a.kbk, (uint32_t)(a.pair ? a.n_tile / 2 : a.n_tile), 1u, 1u};
    const int rc = make_tensor_map(&tmap_b, op_dtype, 4, wt, dims, strides, box, nullptr, 1);
    if (rc) return rc;
  }
  OutMaps om;
  memset(&om, 0, sizeof(om));
  if (a.stage_out) {
    for (int g = 0; g < a.num_phases; ++g) {
      // phase view: pixel (py, px) of the M grid is output pixel (py * o_mul + g_oy, px * o_mul + g_ox)
      const int64_t vw = (d->out_w - a.g_ox[g] + a.o_mul - 1) / a.o_mul, vh = (d->out_h - a.g_oy[g] + a.o_mul - 1) / a.o_mul;
      const char* base = static_cast<const char*>(y) + ((int64_t)a.g_oy[g] * d->ys[2] + (int64_t)a.g_ox[g] * d->ys[3]) * 2;
      const uint64_t dims[4] = {(uint64_t)d->cout, (uint64_t)vw, (uint64_t)vh, (uint64_t)d->n};
      const uint64_t strides[3] = {(uint64_t)(d->ys[3] * a.o_mul * 2), (uint64_t)(d->ys[2] * a.o_mul * 2), (uint64_t)(d->ys[0] * 2)};
      const uint32_t box[4] = {64u, (uint32_t)a.tw, (uint32_t)a.th, (uint32_t)a.tn};
      const int rc = make_tensor_map(&om.m[g], d->y_dtype, 4, base, dims, strides, box, nullptr, 1);
      if (rc) return rc;
    }
  }
  {
    // opt in to the full dynamic shared-memory carve-out ONCE per device and kernel instead of on every launch
    static bool smem_set[2][64] = {};
    int dev = 0;
    VCN_CUDA(cudaGetDevice(&dev));
    const int p = a.pair ? 1 : 0;
    if (dev < 0 || dev >= 64 || !smem_set[p][dev]) {
      if (a.pair) VCN_CUDA(cudaFuncSetAttribute(conv_igemm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
      else VCN_CUDA(cudaFuncSetAttribute(conv_igemm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
      if (dev >= 0 && dev < 64) smem_set[p][dev] = true;
    }
  }
  if (!a.pair) {
    int grid = sm_count() - g_sm_reserve;
    if (grid < 1) grid = 1;
    if (grid > a.tiles_total) grid = a.tiles_total;
    conv_igemm_kernel<false><<<(unsigned)grid, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(tmap_a, tmap_b, om, a);
  } else {
    int pairs = (sm_count() - g_sm_reserve) / 2;
    if (pairs < 1) pairs = 1;
    if (pairs > a.tiles_total) pairs = a.tiles_total;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(2 * pairs), 1, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = static_cast<cudaStream_t>(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    VCN_CUDA(cudaLaunchKernelEx(&cfg, conv_igemm_kernel<true>, tmap_a, tmap_b, om, a));
  }
  VCN_LAUNCH_CHECK("conv_igemm_kernel");
  return VCN_OK;
}

extern "C" int vcn_conv_profile_counters(uint64_t out[8], int reset) {
  using namespace vcn;
  VCN_REQUIRE(out != nullptr, "conv_profile_counters: out must not be NULL");
#if VCN_CONV_PROF
  unsigned long long h[8];
  VCN_CUDA(cudaMemcpyFromSymbol(h, g_conv_prof, sizeof(h)));
  for (int i = 0; i < 8; ++i) out[i] = h[i];
  if (reset) {
    memset(h, 0, sizeof(h));
    VCN_CUDA(cudaMemcpyToSymbol(g_conv_prof, h, sizeof(h)));
  }
  return VCN_OK;
#else
  (void)reset;
  for (int i = 0; i < 8; ++i) out[i] = 0;
  set_error("conv_profile_counters: this build has no profiling counters (compile csrc/modconv.cu with -DVCN_CONV_PROF=1)");
  return VCN_ERR_UNSUPPORTED;
#endif
}
