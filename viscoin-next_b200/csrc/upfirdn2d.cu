// upfirdn2d.cu - pad -> zero-insert upsample -> 2-D FIR -> decimate, for sm_100a.
//
// Replaces upfirdn2d_kernel_small / upfirdn2d_kernel_large (stylegan2_ada/torch_utils/ops/upfirdn2d.cu:29-200)
// and the spec table of upfirdn2d.cu:205-341.  Index contract (upfirdn2d.cu:141-144,176-183):
//   out[oy,ox] = gain * sum_{fy,fx} f'[fy,fx] * U[oy*down + fy - pad0, ...],  U = zero-inserted input,
//   f' = f flipped unless `flip` (flip=0 is true convolution).
//
// Two kernels:
//   tile    : dense NCHW, filter <= 4x4, (up,down) in {(1,1),(2,1),(1,2)}.  One CTA stages the input halo tile of
//             TP planes x IH x IWp in shared memory - by ONE TMA box load (cp.async.bulk.tensor.3d; out-of-bounds
//             zero fill IS the op's zero padding, negative coordinates included) when the row pitch is 16-byte
//             aligned, else by coalesced loads - then every thread produces a 4 x MC register micro-tile from
//             16-byte shared-memory reads and writes it with 16-byte stores.
//   generic : any filter / factors / strides (channels_last included): one output per thread, gather.
// HBM-bound: algorithmic bytes = (N*C*Hin*Win + N*C*Hout*Wout) * e.
#include <string.h>

#include "act_math.cuh"

namespace vcn {
namespace {

struct UpfirdnArgs {
  const void* x;
  const float* f;
  void* y;
  int n, c, in_h, in_w, out_h, out_w;
  int64_t xs[4], ys[4];
  int fh, fw, upx, upy, downx, downy, padx0, pady0, flip;
  float gain;
  // fused epilogue (all optional)
  const float* scale;
  const float* noise;
  int64_t noise_stride_n;
  const float* bias;
  int act;
  float alpha, act_gain, clamp;
};

template <typename S>
__device__ __forceinline__ S epilogue(const UpfirdnArgs& a, S v, int plane, int oy, int ox) {
  v *= S(a.gain);
  if (a.scale) v *= S(a.scale[plane]);
  if (a.noise) v += S(a.noise[(int64_t)(plane / a.c) * a.noise_stride_n + (int64_t)oy * a.out_w + ox]);
  if (a.act) {
    if (a.bias) v += S(a.bias[plane % a.c]);
    v = act_eval<S, 0>(v, S(0), S(0), S(1), a.act, 0, S(a.alpha), S(a.act_gain), S(a.clamp));
  }
  return v;
}

__device__ __forceinline__ int floor_div(int a, int b) {
  const int q = a / b;
  return (a % b != 0 && ((a < 0) != (b < 0))) ? q - 1 : q;
}

// ---------------------------------------------------------------------------------------------
// generic gather kernel
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) upfirdn2d_generic_kernel(const UpfirdnArgs a, const int channels_fastest) {
  using S = typename Acc<T>::type;
  const int64_t total = (int64_t)a.n * a.c * a.out_h * a.out_w;
  const T* __restrict__ px = static_cast<const T*>(a.x);
  T* __restrict__ py = static_cast<T*>(a.y);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    int ox, oy, ch, nb;
    int64_t r = i;
    if (channels_fastest) {
      ch = (int)(r % a.c); r /= a.c;
      ox = (int)(r % a.out_w); r /= a.out_w;
      oy = (int)(r % a.out_h); nb = (int)(r / a.out_h);
    } else {
      ox = (int)(r % a.out_w); r /= a.out_w;
      oy = (int)(r % a.out_h); r /= a.out_h;
      ch = (int)(r % a.c); nb = (int)(r / a.c);
    }
    const T* xb = px + nb * a.xs[0] + ch * a.xs[1];
    S acc = S(0);
    const int my = oy * a.downy - a.pady0, mx = ox * a.downx - a.padx0;
    for (int fy = 0; fy < a.fh; ++fy) {
      const int uy = my + fy;
      if (uy < 0 || uy % a.upy != 0) continue;
      const int iy = uy / a.upy;
      if (iy >= a.in_h) continue;
      for (int fx = 0; fx < a.fw; ++fx) {
        const int ux = mx + fx;
        if (ux < 0 || ux % a.upx != 0) continue;
        const int ix = ux / a.upx;
        if (ix >= a.in_w) continue;
        const float w = a.flip ? a.f[fy * a.fw + fx] : a.f[(a.fh - 1 - fy) * a.fw + (a.fw - 1 - fx)];
        acc += S(w) * to_acc<T>(xb[iy * a.xs[2] + ix * a.xs[3]]);
      }
    }
    acc = epilogue<S>(a, acc, nb * a.c + ch, oy, ox);
    py[nb * a.ys[0] + ch * a.ys[1] + oy * a.ys[2] + ox * a.ys[3]] = from_acc<T>(acc);
  }
}

// ---------------------------------------------------------------------------------------------
// tile kernel
// ---------------------------------------------------------------------------------------------
constexpr int kF = 4;   // filter taps per axis (smaller filters are zero-extended)
constexpr int kMR = 4;  // output rows per thread

struct TileGeom {
  int tw, th, tp;        // output tile: cols, rows, planes
  int tx, ty;            // threads along cols / rows (tp along planes)
  int iw, ih, iwp;       // staged input tile: valid cols, rows, pitch (elements)
  int tiles_x, tiles_y, tiles_p;
  int use_tma, vec_store;
};

// EPI: 0 = plain FIR (gain folded into the taps; nothing after the accumulation), 1 = fused scale / noise / bias with a
// linear or lrelu activation (the synthesis layers), 2 = any activation (run-time dispatch).
// STG: 0 = halo tile staged by TMA / cp.async / plain loads into two buffers; 1 = fp32, one plane per tile, row pitch not
// 16-byte aligned (the 2H+1 intermediate): the NEXT tile is prefetched into registers with coalesced 4-byte loads while
// this one is filtered, then written to the single buffer - 2 instructions per element instead of ~14 for 4-byte
// cp.async (address math + the LDGSTS issue sequence), which made the kernel issue-bound.
template <typename T, int UP, int DOWN, int PHX, int PHY, int EPI, int STG>
__global__ void __launch_bounds__(256, STG != 0 ? 3 : 1)
upfirdn2d_tile_kernel(const UpfirdnArgs a, const TileGeom g, const __grid_constant__ CUtensorMap tmap) {
  using S = typename Acc<T>::type;
  constexpr int V = Vec16<T>::N;   // elements per 16 bytes
  constexpr int MC = V;            // output cols per thread
  constexpr int WR = ((kMR - 1) * DOWN + kF - 1 - PHY) / UP + 1;  // window rows a thread reads
  constexpr int WC = ((MC - 1) * DOWN + kF - 1 - PHX) / UP + 1;   // window cols
  constexpr int NV = (UP == 1) ? (WC + V - 1) / V : 0;             // 16-byte loads per window row (UP==1)
  constexpr int WCL = (UP == 1) ? NV * V : WC;

  // Persistent CTA: tiles blockIdx.x, blockIdx.x + gridDim.x, ... with two shared-memory buffers - the halo tile of
  // the next tile is in flight (TMA or cp.async) while this one is filtered and stored.
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int tile_elems = g.tp * g.ih * g.iwp;
  const size_t buf_bytes = ((size_t)tile_elems * sizeof(T) + 127) & ~(size_t)127;
  float* sflt = reinterpret_cast<float*>(smem_raw + 2 * buf_bytes);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sflt + kF * kF);  // [2]

  const int tid = threadIdx.x;
  const int planes = a.n * a.c;
  const int total_tiles = g.tiles_x * g.tiles_y * g.tiles_p;

  // tile index -> (bx, by, bp), advanced incrementally by gridDim.x (no integer divisions in the loop)
  struct TileIdx { int bx, by, bp; };
  auto decompose = [&](int t) {
    TileIdx r;
    r.bx = t % g.tiles_x; r.by = (t / g.tiles_x) % g.tiles_y; r.bp = t / (g.tiles_x * g.tiles_y);
    return r;
  };
  const TileIdx stride = decompose((int)gridDim.x);
  auto advance = [&](TileIdx v) {
    v.bx += stride.bx;
    if (v.bx >= g.tiles_x) { v.bx -= g.tiles_x; ++v.by; }
    v.by += stride.by;
    if (v.by >= g.tiles_y) { v.by -= g.tiles_y; ++v.bp; }
    v.bp += stride.bp;
    return v;
  };
  constexpr int KMAX = 5;  // staged row pitch <= 160 elements on the cp.async path (host guarantees it)

  // register path shapes: STG 1 = up to 40 rows x 160 columns (wide tiles), STG 2 = up to 72 rows x 96 columns (64-wide)
  constexpr int RMAX = STG == 2 ? 9 : 5, RK = STG == 2 ? 3 : 5;
  T pre[STG != 0 ? RMAX : 1][STG != 0 ? RK : 1];
  auto fetch = [&](const TileIdx ti) {  // STG == 1: tile ti -> registers (tp == 1, blockDim == 256)
    if constexpr (STG != 0) {
      const int warp = tid >> 5, lane = tid & 31;
      const int ix0 = ti.bx * g.tw * DOWN - a.padx0, iy0 = ti.by * g.th * DOWN - a.pady0;  // UP == 1
      const T* __restrict__ base = static_cast<const T*>(a.x) + (int64_t)ti.bp * a.in_h * a.in_w;
      const int gx0 = ix0 + lane;
      uint32_t in_image = 0;
#pragma unroll
      for (int k = 0; k < RK; ++k) {
        const int gx = gx0 + 32 * k;
        if (lane + 32 * k < g.iwp && gx >= 0 && gx < a.in_w) in_image |= 1u << k;
      }
#pragma unroll
      for (int j = 0; j < RMAX; ++j) {
        const int gy = iy0 + warp + 8 * j;
        const bool row_ok = warp + 8 * j < g.ih && gy >= 0 && gy < a.in_h;
        const T* srow = base + (int64_t)gy * a.in_w + gx0;
#pragma unroll
        for (int k = 0; k < RK; ++k)
          pre[j][k] = (row_ok && ((in_image >> k) & 1u)) ? __ldg(srow + 32 * k) : T(0);
      }
    }
  };
  auto stash = [&]() {  // registers -> the (single) shared buffer
    if constexpr (STG != 0) {
      const int warp = tid >> 5, lane = tid & 31;
      T* tile = reinterpret_cast<T*>(smem_raw);
#pragma unroll
      for (int j = 0; j < RMAX; ++j) {
        if (warp + 8 * j < g.ih) {
#pragma unroll
          for (int k = 0; k < RK; ++k)
            if (lane + 32 * k < g.iwp) tile[(warp + 8 * j) * g.iwp + lane + 32 * k] = pre[j][k];
        }
      }
    }
  };

  auto issue = [&](const TileIdx ti, int b) {
    T* tile = reinterpret_cast<T*>(smem_raw + b * buf_bytes);
    const int plane0 = ti.bp * g.tp;
    // first staged input column/row: ceil((tile_o*DOWN - pad0) / UP)
    const int ix0 = -floor_div(-(ti.bx * g.tw * DOWN - a.padx0), UP);
    const int iy0 = -floor_div(-(ti.by * g.th * DOWN - a.pady0), UP);
    if (g.use_tma) {
      if (tid == 0) {
        mbar_expect_tx(bar + b, (uint32_t)(tile_elems * sizeof(T)));
        tma_load_3d(tile, &tmap, bar + b, ix0, iy0, plane0);
      }
      return;
    }
    const T* __restrict__ px = static_cast<const T*>(a.x);
    const int warp = tid >> 5, lane = tid & 31, nwarps = blockDim.x >> 5;
    const int nrows = g.tp * g.ih;
    if constexpr (sizeof(T) == 4) {
      // odd row pitch (the 2H+1 transposed-conv intermediate): rows are only 4-byte aligned, so neither TMA nor
      // 16-byte loads apply.  4-byte cp.async (LDGSTS) with zero fill: the whole halo tile is in flight at once,
      // no registers staged, a warp covers 128 contiguous bytes per instruction.  Column validity is a per-thread
      // bit mask computed once per tile; the row loop only moves two pointers.
      const int gx0 = ix0 + lane;
      uint32_t in_tile = 0, in_image = 0;
#pragma unroll
      for (int k = 0; k < KMAX; ++k) {
        const int cc = lane + 32 * k, gx = gx0 + 32 * k;
        if (cc < g.iwp) in_tile |= 1u << k;
        if (cc < g.iwp && gx >= 0 && gx < a.in_w) in_image |= 1u << k;
      }
      for (int rr = warp; rr < nrows; rr += nwarps) {
        const int p = g.tp == 1 ? 0 : rr / g.ih, r = rr - p * g.ih;
        const int gy = iy0 + r, pl = plane0 + p;
        const uint32_t row_mask = (gy >= 0 && gy < a.in_h && pl < planes) ? in_image : 0u;
        const T* srow = px + ((int64_t)pl * a.in_h + gy) * a.in_w + gx0;
        const uint32_t drow = smem_u32(tile + (size_t)rr * g.iwp + lane);
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
          if (in_tile & (1u << k)) {
            const bool ok = (row_mask >> k) & 1u;
            cp_async4_zfill(drow + 128u * k, ok ? srow + 32 * k : px, ok);
          }
        }
      }
    } else {
      for (int rr = warp; rr < nrows; rr += nwarps) {
        const int p = g.tp == 1 ? 0 : rr / g.ih, r = rr - p * g.ih;
        const int gy = iy0 + r, pl = plane0 + p;
        const bool row_ok = gy >= 0 && gy < a.in_h && pl < planes;
        const T* src = px + ((int64_t)pl * a.in_h + gy) * a.in_w;
        T* dst = tile + (size_t)rr * g.iwp;
        for (int cc = lane; cc < g.iwp; cc += 32) {
          const int gx = ix0 + cc;
          dst[cc] = (row_ok && gx >= 0 && gx < a.in_w) ? src[gx] : from_acc<T>(S(0));
        }
      }
    }
  };

  if (g.use_tma && tid == 0) {
    mbar_init(bar, 1);
    mbar_init(bar + 1, 1);
    fence_mbar_init();
  }
  if (tid < kF * kF) {
    const int fy = tid / kF, fx = tid % kF;
    float w = 0.f;
    if (fy < a.fh && fx < a.fw) w = a.flip ? a.f[fy * a.fw + fx] : a.f[(a.fh - 1 - fy) * a.fw + (a.fw - 1 - fx)];
    sflt[tid] = w * a.gain;  // the reference scales the filter too (upfirdn2d.py:193)
  }
  __syncthreads();
  TileIdx cur = decompose((int)blockIdx.x);
  if constexpr (STG != 0) {
    fetch(cur);
    stash();
    __syncthreads();
  } else {
    issue(cur, 0);
    cp_async_commit();
  }

  S flt[kF][kF];
#pragma unroll
  for (int i = 0; i < kF; ++i)
#pragma unroll
    for (int j = 0; j < kF; ++j) flt[i][j] = S(sflt[i * kF + j]);

  const int t_x = tid % g.tx;
  const int t_y = (tid / g.tx) % g.ty;
  const int t_p = tid / (g.tx * g.ty);
  const int rx = t_x * MC, ry = t_y * kMR;
  int it = 0;
  for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
  const int b = STG != 0 ? 0 : (it & 1);
  const TileIdx nxt = advance(cur);
  const bool has_next = t + (int)gridDim.x < total_tiles;
  if constexpr (STG != 0) {
    if (has_next) fetch(nxt);   // loads stay in flight while this tile is filtered
  } else {
    if (has_next) issue(nxt, b ^ 1);
    cp_async_commit();
    cp_async_wait_1();       // this thread's copies of tile t have landed ...
    __syncthreads();         // ... and everybody else's
    if (g.use_tma) mbar_wait(bar + b, (uint32_t)(it >> 1) & 1u);
  }
  const T* tile = reinterpret_cast<const T*>(smem_raw + b * buf_bytes);
  const int tile_ox = cur.bx * g.tw, tile_oy = cur.by * g.th;
  const int plane0 = cur.bp * g.tp;
  cur = nxt;
  // register-staged tiles may use fewer than blockDim.x compute threads (balanced tiles of a 2^k+1 output): the rest only
  // help with the staging
  if (tid < g.tx * g.ty * g.tp) {
  const T* wbase = tile + ((size_t)t_p * g.ih + (ry * DOWN) / UP) * g.iwp + (rx * DOWN) / UP;

  S acc[kMR][MC];
#pragma unroll
  for (int r = 0; r < kMR; ++r)
#pragma unroll
    for (int c = 0; c < MC; ++c) acc[r][c] = S(0);

#pragma unroll
  for (int wr = 0; wr < WR; ++wr) {
    S w[WCL];
    const T* srow = wbase + (size_t)wr * g.iwp;
    if constexpr (UP == 1) {
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const Vec16<T> v = *reinterpret_cast<const Vec16<T>*>(srow + k * V);
#pragma unroll
        for (int j = 0; j < V; ++j) w[k * V + j] = to_acc<T>(v.v[j]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < WC; ++j) w[j] = to_acc<T>(srow[j]);
    }
#pragma unroll
    for (int r = 0; r < kMR; ++r) {
      const int fy = UP * wr + PHY - r * DOWN;
      if (fy < 0 || fy >= kF) continue;
#pragma unroll
      for (int c = 0; c < MC; ++c) {
#pragma unroll
        for (int wc = 0; wc < WC; ++wc) {
          const int fx = UP * wc + PHX - c * DOWN;
          if (fx < 0 || fx >= kF) continue;
          acc[r][c] = fma(w[wc], flt[fy][fx], acc[r][c]);
        }
      }
    }
  }

  const int plane = plane0 + t_p;
  const int ox = tile_ox + rx, oy_base = tile_oy + ry;
  if (plane < planes && ox < a.out_w) {
  T* __restrict__ py = static_cast<T*>(a.y) + (int64_t)plane * a.out_h * a.out_w;
  // per-thread constants of the fused epilogue (the plane is fixed per thread)
  S sc = S(1), bi = S(0);
  const float* nrow = nullptr;
  if constexpr (EPI != 0) {
    if (a.scale) sc = S(a.scale[plane]);
    if (a.act && a.bias) bi = S(a.bias[plane % a.c]);
    if (a.noise) nrow = a.noise + (int64_t)(plane / a.c) * a.noise_stride_n;
  }
  const bool full = ox + MC <= a.out_w;
#pragma unroll
  for (int r = 0; r < kMR; ++r) {
    const int oy = oy_base + r;
    if (oy >= a.out_h) break;
    Vec16<T> out;
#pragma unroll
    for (int c = 0; c < MC; ++c) {
      S v = acc[r][c];
      if constexpr (EPI != 0) {
        v *= sc;
        if (nrow) v += S(nrow[(int64_t)oy * a.out_w + (ox + c < a.out_w ? ox + c : a.out_w - 1)]);
        if constexpr (EPI == 1) {
          if (a.act) {  // linear (alpha = 1) or lrelu
            v += bi;
            v = (v > S(0) ? v : v * S(a.act == VCN_ACT_LRELU ? a.alpha : 1.f)) * S(a.act_gain);
            if (a.clamp >= 0.f) v = (v > S(-a.clamp) && v < S(a.clamp)) ? v : (v >= S(0) ? S(a.clamp) : S(-a.clamp));
          }
        } else {
          if (a.act) v = act_eval<S, 0>(v + bi, S(0), S(0), S(1), a.act, 0, S(a.alpha), S(a.act_gain), S(a.clamp));
        }
      }
      out.v[c] = from_acc<T>(v);
    }
    T* dst = py + (int64_t)oy * a.out_w + ox;
    if (g.vec_store && full) {
      st_stream(dst, out);
    } else {
#pragma unroll
      for (int c = 0; c < MC; ++c)
        if (ox + c < a.out_w) dst[c] = out.v[c];
    }
  }
  }
  }
  __syncthreads();  // buffer b is free for the tile after next
  if constexpr (STG != 0) {
    if (has_next) stash();
    __syncthreads();
  }
  }
}


// ---------------------------------------------------------------------------------------------
// channels-last kernel: 16-bit storage, up = down = 1, filter <= 4x4 (the FIR after the transposed convolution and its
// adjoint).  x is NHWC-strided (xs[1] == 1).  One CTA = 16 x 32 output pixels x 64 channels; the (16+3) x (32+3) x 64ch
// halo tile is ONE TMA 4-D box load (inner coordinate = channel offset, always 16-byte aligned; out-of-bounds pixels
// zero-fill = the op's padding).  Thread = 8 channels (one 16-byte vector) x four 2x2 pixel blocks: each staged vector
// is read once from shared memory per 2x2 block (25 LDS.128 for 4 outputs) - the 8 threads of a quarter warp read the
// 128 contiguous bytes of one pixel, so the reads are bank-conflict free.
// ---------------------------------------------------------------------------------------------
constexpr int kClTx = 32, kClTy = 16, kClCb = 64;

template <typename T, typename TO>
__global__ void __launch_bounds__(256, 2)
upfirdn2d_cl_kernel(const UpfirdnArgs a, const __grid_constant__ CUtensorMap tmap, const int tiles_x, const int tiles_y,
                    const int ncb) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int HX = kClTx + kF - 1, HY = kClTy + kF - 1;
  T* tile = reinterpret_cast<T*>(smem_raw);                                   // [HY][HX][64]
  float* s_par = reinterpret_cast<float*>(smem_raw + (size_t)HY * HX * kClCb * sizeof(T));  // scale[64] | bias[64] | flt[16]
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_par + 2 * kClCb + kF * kF);

  const int tid = threadIdx.x;
  // channel block fastest: the CTAs that read the two 128-byte halves of the same 256-byte L2 line (TMA promotes every
  // request to 256 bytes) run at the same time - with the channel block as the SLOWEST index the promoted half was fetched
  // from DRAM a second time one whole tensor pass later (ncu: 1.49x read amplification on [8,128,257,257])
  int bid = blockIdx.x;
  const int c0 = (bid % ncb) * kClCb; bid /= ncb;
  const int bx = bid % tiles_x; bid /= tiles_x;
  const int by = bid % tiles_y;
  const int n = bid / tiles_y;
  const int ox0 = bx * kClTx, oy0 = by * kClTy;

  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    mbar_expect_tx(bar, (uint32_t)(HY * HX * kClCb * sizeof(T)));
    tma_load_4d(tile, &tmap, bar, c0, ox0 - a.padx0, oy0 - a.pady0, n);
  }
  if (tid < kF * kF) {
    const int fy = tid / kF, fx = tid % kF;
    float w = 0.f;
    if (fy < a.fh && fx < a.fw) w = a.flip ? a.f[fy * a.fw + fx] : a.f[(a.fh - 1 - fy) * a.fw + (a.fw - 1 - fx)];
    s_par[2 * kClCb + tid] = w * a.gain;
  } else if (tid >= 64 && tid < 64 + kClCb) {
    const int c = c0 + tid - 64;
    s_par[tid - 64] = (a.scale && c < a.c) ? a.scale[(int64_t)n * a.c + c] : 1.f;
  } else if (tid >= 128 && tid < 128 + kClCb) {
    const int c = c0 + tid - 128;
    s_par[kClCb + tid - 128] = (a.bias && a.act && c < a.c) ? a.bias[c] : 0.f;
  }
  __syncthreads();
  mbar_wait(bar, 0);

  float flt[kF][kF];
#pragma unroll
  for (int i = 0; i < kF; ++i)
#pragma unroll
    for (int j = 0; j < kF; ++j) flt[i][j] = s_par[2 * kClCb + i * kF + j];

  const int cg = tid & 7;          // channel group: channels c0 + cg*8 .. +7
  const int pb = tid >> 3;         // 0..31
  const int c = c0 + cg * 8;
  if (c >= a.c) return;
  float sc[8], bi[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) { sc[k] = s_par[cg * 8 + k]; bi[k] = s_par[kClCb + cg * 8 + k]; }

#pragma unroll 1
  for (int it = 0; it < 4; ++it) {
    const int blk = pb + it * 32;               // 128 blocks of 2x2 pixels: 16 wide x 8 high
    const int bxx = blk % (kClTx / 2), byy = blk / (kClTx / 2);
    const int oy = oy0 + 2 * byy, ox = ox0 + 2 * bxx;
    if (oy >= a.out_h || ox >= a.out_w) continue;
    float acc[2][2][8];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[i][j][k] = 0.f;
#pragma unroll
    for (int wy = 0; wy < kF + 1; ++wy) {
#pragma unroll
      for (int wx = 0; wx < kF + 1; ++wx) {
        const Vec16<T> v = *reinterpret_cast<const Vec16<T>*>(tile + ((size_t)(2 * byy + wy) * HX + (2 * bxx + wx)) * kClCb + cg * 8);
        float f[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) f[k] = to_acc<T>(v.v[k]);
#pragma unroll
        for (int dy = 0; dy < 2; ++dy) {
          const int fy = wy - dy;
          if (fy < 0 || fy >= kF) continue;
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) {
            const int fx = wx - dx;
            if (fx < 0 || fx >= kF) continue;
#pragma unroll
            for (int k = 0; k < 8; ++k) acc[dy][dx][k] = fmaf(f[k], flt[fy][fx], acc[dy][dx][k]);
          }
        }
      }
    }
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int yy = oy + dy, xx = ox + dx;
        if (yy >= a.out_h || xx >= a.out_w) continue;
        float nz = 0.f;
        if (a.noise) nz = a.noise[(int64_t)n * a.noise_stride_n + (int64_t)yy * a.out_w + xx];
        TO* dst = static_cast<TO*>(a.y) + n * a.ys[0] + yy * a.ys[2] + xx * a.ys[3] + c;
        constexpr int VO = 16 / sizeof(TO);
#pragma unroll
        for (int h = 0; h < 8 / VO; ++h) {
          Vec16<TO> o;
#pragma unroll
          for (int k = 0; k < VO; ++k) {
            float r = acc[dy][dx][h * VO + k] * sc[h * VO + k] + nz;
            if (a.act) r = act_eval<float, 0>(r + bi[h * VO + k], 0.f, 0.f, 1.f, a.act, 0, a.alpha, a.act_gain, a.clamp);
            o.v[k] = from_acc<TO>(r);
          }
          *reinterpret_cast<Vec16<TO>*>(dst + h * VO) = o;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// channels-last kernel, separable form.  Same tile, TMA box and thread/channel mapping as upfirdn2d_cl_kernel, but the
// 4x4 filter is factored inside the kernel into gy (x) gx (every filter setup_filter() builds from a 1-D kernel is an
// outer product; the factorisation is checked against all 16 taps and a non-separable filter takes the direct loop at the
// end).  Thread = 8 channels x 2 columns x 8 rows, walked as a sliding window over 11 staged rows: per row 5 LDS.128, a
// 4-tap horizontal pass into two column sums and a 4-tap vertical pass into a ring of four output-row accumulators -
// 8 FMAs per output instead of 16, issued as packed fma.rn.f32x2 (FFMA2: the two channels of one bf16x2 word), i.e.
// ~10 issue slots per output instead of ~23.  The direct kernel was issue-bound (ncu: issue-active 70+ %).
// EPI: 0 = plain FIR, 1 = scale / noise / bias with linear or lrelu, 2 = any activation (run-time dispatch).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t pack_f2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack_f2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t mul_f2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
template <typename T> __device__ __forceinline__ uint64_t widen_f2(uint32_t w);
template <> __device__ __forceinline__ uint64_t widen_f2<__nv_bfloat16>(uint32_t w) {
  return pack_f2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
template <> __device__ __forceinline__ uint64_t widen_f2<__half>(uint32_t w) {
  const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w));
  return pack_f2(f.x, f.y);
}

// one output pixel x 8 channels: epilogue + 16-byte stores
template <typename TO, int EPI>
__device__ __forceinline__ void cl_emit(const UpfirdnArgs& a, const float (&v)[8], const float* s_par, int cg, int n, int c,
                                        int yy, int xx) {
  if (yy >= a.out_h || xx >= a.out_w) return;
  float r[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) r[k] = v[k];
  if constexpr (EPI != 0) {
    const float nz = a.noise ? __ldg(a.noise + (int64_t)n * a.noise_stride_n + (int64_t)yy * a.out_w + xx) : 0.f;
    const float4 s0 = *reinterpret_cast<const float4*>(s_par + cg * 8), s1 = *reinterpret_cast<const float4*>(s_par + cg * 8 + 4);
    const float4 b0 = *reinterpret_cast<const float4*>(s_par + kClCb + cg * 8), b1 = *reinterpret_cast<const float4*>(s_par + kClCb + cg * 8 + 4);
    const float sc[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    const float bi[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float t = fmaf(r[k], sc[k], nz);
      if (a.act) t = act_eval<float, 0>(t + bi[k], 0.f, 0.f, 1.f, a.act, 0, a.alpha, a.act_gain, a.clamp);
      r[k] = t;
    }
  }
  TO* dst = static_cast<TO*>(a.y) + n * a.ys[0] + yy * a.ys[2] + xx * a.ys[3] + c;
  constexpr int VO = 16 / sizeof(TO);
#pragma unroll
  for (int h = 0; h < 8 / VO; ++h) {
    Vec16<TO> o;
#pragma unroll
    for (int k = 0; k < VO; ++k) o.v[k] = from_acc<TO>(r[h * VO + k]);
    *reinterpret_cast<Vec16<TO>*>(dst + h * VO) = o;
  }
}

// EPI = 1, packed: lrelu / linear are positively homogeneous, so with g = act_gain > 0 and 0 <= alpha <= 1
//   clamp(lrelu(acc*sc + nz + b) * g) = clamp(max(t, alpha*t)),  t = acc*(sc*g) + (nz*g + b*g)
// (sc*g and b*g are staged per channel, alpha = 1 for linear / no activation): add2 + fma2 + mul2 per channel pair.
template <typename TO>
__device__ __forceinline__ void cl_emit_packed(const UpfirdnArgs& a, const uint64_t (&v)[4], const float* s_par, int cg, int n,
                                               int c, int yy, int xx, float g, uint64_t alpha2, bool has_clamp) {
  if (yy >= a.out_h || xx >= a.out_w) return;
  float nz = 0.f;
  if (a.noise) nz = __ldg(a.noise + (int64_t)n * a.noise_stride_n + (int64_t)yy * a.out_w + xx) * g;
  const uint64_t nz2 = pack_f2(nz, nz);
  const ulonglong2 s0 = *reinterpret_cast<const ulonglong2*>(s_par + cg * 8), s1 = *reinterpret_cast<const ulonglong2*>(s_par + cg * 8 + 4);
  const ulonglong2 b0 = *reinterpret_cast<const ulonglong2*>(s_par + kClCb + cg * 8), b1 = *reinterpret_cast<const ulonglong2*>(s_par + kClCb + cg * 8 + 4);
  const uint64_t sc[4] = {s0.x, s0.y, s1.x, s1.y}, bi[4] = {b0.x, b0.y, b1.x, b1.y};
  float r[8];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    uint64_t off;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(off) : "l"(nz2), "l"(bi[p]));
    const uint64_t t = fma_f2(v[p], sc[p], off);
    const uint64_t u = mul_f2(t, alpha2);
    float t0, t1, u0, u1;
    unpack_f2(t, t0, t1);
    unpack_f2(u, u0, u1);
    r[2 * p] = fmaxf(t0, u0);
    r[2 * p + 1] = fmaxf(t1, u1);
  }
  if (has_clamp) {
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = fminf(fmaxf(r[k], -a.clamp), a.clamp);
  }
  TO* dst = static_cast<TO*>(a.y) + n * a.ys[0] + yy * a.ys[2] + xx * a.ys[3] + c;
  constexpr int VO = 16 / sizeof(TO);
#pragma unroll
  for (int h = 0; h < 8 / VO; ++h) {
    Vec16<TO> o;
#pragma unroll
    for (int k = 0; k < VO; ++k) o.v[k] = from_acc<TO>(r[h * VO + k]);
    *reinterpret_cast<Vec16<TO>*>(dst + h * VO) = o;
  }
}

template <typename T, typename TO, int EPI>
__global__ void __launch_bounds__(256, 2)
upfirdn2d_cl_sep_kernel(const UpfirdnArgs a, const __grid_constant__ CUtensorMap tmap, const int tiles_x, const int tiles_y,
                        const int ncb) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int HX = kClTx + kF - 1, HY = kClTy + kF - 1;
  constexpr int RT = kClTy / 2;  // output rows per thread
  T* tile = reinterpret_cast<T*>(smem_raw);                                   // [HY][HX][64]
  float* s_par = reinterpret_cast<float*>(smem_raw + (size_t)HY * HX * kClCb * sizeof(T));  // scale[64] | bias[64] | flt[16]
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_par + 2 * kClCb + kF * kF);

  const int tid = threadIdx.x;
  // channel block fastest: the CTAs that read the two 128-byte halves of the same 256-byte L2 line (TMA promotes every
  // request to 256 bytes) run at the same time - with the channel block as the SLOWEST index the promoted half was fetched
  // from DRAM a second time one whole tensor pass later (ncu: 1.49x read amplification on [8,128,257,257])
  int bid = blockIdx.x;
  const int c0 = (bid % ncb) * kClCb; bid /= ncb;
  const int bx = bid % tiles_x; bid /= tiles_x;
  const int by = bid % tiles_y;
  const int n = bid / tiles_y;
  const int ox0 = bx * kClTx, oy0 = by * kClTy;

  const float epi_g = (EPI == 1 && a.act) ? a.act_gain : 1.f;  // EPI 1 folds the activation gain into scale / bias / noise
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0) {
    mbar_expect_tx(bar, (uint32_t)(HY * HX * kClCb * sizeof(T)));
    tma_load_4d(tile, &tmap, bar, c0, ox0 - a.padx0, oy0 - a.pady0, n);
  }
  if (tid < kF * kF) {
    const int fy = tid / kF, fx = tid % kF;
    float w = 0.f;
    if (fy < a.fh && fx < a.fw) w = a.flip ? a.f[fy * a.fw + fx] : a.f[(a.fh - 1 - fy) * a.fw + (a.fw - 1 - fx)];
    s_par[2 * kClCb + tid] = w * a.gain;
  } else if (tid >= 64 && tid < 64 + kClCb) {
    const int c = c0 + tid - 64;
    s_par[tid - 64] = ((a.scale && c < a.c) ? a.scale[(int64_t)n * a.c + c] : 1.f) * epi_g;
  } else if (tid >= 128 && tid < 128 + kClCb) {
    const int c = c0 + tid - 128;
    s_par[kClCb + tid - 128] = ((a.bias && a.act && c < a.c) ? a.bias[c] : 0.f) * epi_g;
  }
  __syncthreads();

  // rank-1 factorisation flt[i][j] = gy[i] * gx[j] around the largest tap (while the TMA load is in flight)
  float gx[kF], gy[kF];
  bool sep = true;
  {
    float flt[kF * kF];
    int piv = 0;
    float pv = 0.f;
#pragma unroll
    for (int i = 0; i < kF * kF; ++i) {
      flt[i] = s_par[2 * kClCb + i];
      if (fabsf(flt[i]) > fabsf(pv)) { pv = flt[i]; piv = i; }
    }
    const int pr = piv / kF, pc = piv % kF;
    const float inv = pv != 0.f ? 1.f / pv : 0.f;
#pragma unroll
    for (int j = 0; j < kF; ++j) {
      gx[j] = 0.f; gy[j] = 0.f;
#pragma unroll
      for (int i = 0; i < kF; ++i) {
        if (i == pr) gx[j] = flt[i * kF + j];
        if (i == pc) gy[j] = flt[j * kF + i] * inv;
      }
    }
#pragma unroll
    for (int i = 0; i < kF; ++i)
#pragma unroll
      for (int j = 0; j < kF; ++j) sep = sep && fabsf(flt[i * kF + j] - gy[i] * gx[j]) <= 1e-6f * fabsf(pv);
  }

  const int cg = tid & 7;           // channel group: channels c0 + cg*8 .. +7
  const int cp = (tid >> 3) & 15;   // column pair: tile columns 2cp, 2cp+1
  const int rh = tid >> 7;          // row half: tile rows 8rh .. 8rh+7
  const int c = c0 + cg * 8;
  mbar_wait(bar, 0);
  if (c >= a.c) return;
  const T* base = tile + ((size_t)(RT * rh) * HX + 2 * cp) * kClCb + cg * 8;
  const int oyb = oy0 + RT * rh, oxb = ox0 + 2 * cp;

  if (sep) {
    uint64_t gx2[kF], gy2[kF];
#pragma unroll
    for (int j = 0; j < kF; ++j) { gx2[j] = pack_f2(gx[j], gx[j]); gy2[j] = pack_f2(gy[j], gy[j]); }
    uint64_t acc[4][2][4];  // [output row & 3][column][channel pair]
    const float alpha = a.act == VCN_ACT_LRELU ? a.alpha : 1.f;
    const uint64_t alpha2 = pack_f2(alpha, alpha);
    const bool has_clamp = a.act != 0 && a.clamp >= 0.f;
#pragma unroll
    for (int r = 0; r < RT + kF - 1; ++r) {
      uint64_t h[2][4];
#pragma unroll
      for (int j = 0; j < kF + 1; ++j) {
        const uint4 v = *reinterpret_cast<const uint4*>(base + ((size_t)r * HX + j) * kClCb);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          const uint64_t f = widen_f2<T>(w[p]);
          if (j < kF) h[0][p] = (j == 0) ? mul_f2(f, gx2[0]) : fma_f2(f, gx2[j], h[0][p]);
          if (j >= 1) h[1][p] = (j == 1) ? mul_f2(f, gx2[0]) : fma_f2(f, gx2[j - 1], h[1][p]);
        }
      }
#pragma unroll
      for (int i = 0; i < kF; ++i) {
        const int o = r - i;  // staged row r is tap i of output row o
        if (o < 0 || o >= RT) continue;
#pragma unroll
        for (int col = 0; col < 2; ++col)
#pragma unroll
          for (int p = 0; p < 4; ++p)
            acc[o & 3][col][p] = (i == 0) ? mul_f2(h[col][p], gy2[0]) : fma_f2(h[col][p], gy2[i], acc[o & 3][col][p]);
      }
      if (r >= kF - 1) {
        const int o = r - (kF - 1);
#pragma unroll
        for (int col = 0; col < 2; ++col) {
          if constexpr (EPI == 1) {
            cl_emit_packed<TO>(a, acc[o & 3][col], s_par, cg, n, c, oyb + o, oxb + col, epi_g, alpha2, has_clamp);
          } else {
            float v[8];
#pragma unroll
            for (int p = 0; p < 4; ++p) unpack_f2(acc[o & 3][col][p], v[2 * p], v[2 * p + 1]);
            cl_emit<TO, EPI>(a, v, s_par, cg, n, c, oyb + o, oxb + col);
          }
        }
      }
    }
  } else {
    // non-separable filter: direct 16-tap accumulation, one pixel at a time (correctness path)
#pragma unroll 1
    for (int o = 0; o < RT; ++o) {
#pragma unroll 1
      for (int col = 0; col < 2; ++col) {
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = 0.f;
#pragma unroll 1
        for (int t = 0; t < kF * kF; ++t) {
          const float w = s_par[2 * kClCb + t];
          const Vec16<T> in = *reinterpret_cast<const Vec16<T>*>(base + ((size_t)(o + t / kF) * HX + col + t % kF) * kClCb);
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = fmaf(to_acc<T>(in.v[k]), w, v[k]);
        }
        if constexpr (EPI == 1) {
          uint64_t vp[4];
#pragma unroll
          for (int p = 0; p < 4; ++p) vp[p] = pack_f2(v[2 * p], v[2 * p + 1]);
          const float alpha = a.act == VCN_ACT_LRELU ? a.alpha : 1.f;
          cl_emit_packed<TO>(a, vp, s_par, cg, n, c, oyb + o, oxb + col, epi_g, pack_f2(alpha, alpha), a.act != 0 && a.clamp >= 0.f);
        } else {
          cl_emit<TO, EPI>(a, v, s_par, cg, n, c, oyb + o, oxb + col);
        }
      }
    }
  }
}

template <typename T, typename TO, int EPI>
int launch_cl_sep(const UpfirdnArgs& a, const CUtensorMap& tmap, dim3 grid, size_t smem, int tiles_x, int tiles_y, int ncb,
                  cudaStream_t st) {
  auto kern = upfirdn2d_cl_sep_kernel<T, TO, EPI>;
  VCN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<grid, 256, smem, st>>>(a, tmap, tiles_x, tiles_y, ncb);
  VCN_LAUNCH_CHECK("upfirdn2d_cl_sep_kernel");
  return VCN_OK;
}

// Eligibility + launch of the channels-last kernel.  Returns -1 when the configuration is not covered.
template <typename T>
int try_launch_cl(const UpfirdnArgs& a, int dtype, int y_dtype, cudaStream_t st) {
  if (sizeof(T) != 2) return -1;
  if (!(a.upx == 1 && a.upy == 1 && a.downx == 1 && a.downy == 1 && a.fw <= kF && a.fh <= kF)) return -1;
  if (!(a.xs[1] == 1 && a.ys[1] == 1 && a.c % 8 == 0)) return -1;
  if (!(a.xs[0] % 8 == 0 && a.xs[2] % 8 == 0 && a.xs[3] % 8 == 0 && aligned16(a.x) && aligned16(a.y))) return -1;
  const int64_t vo = y_dtype == VCN_F32 ? 4 : 8;
  if (!(a.ys[0] % vo == 0 && a.ys[2] % vo == 0 && a.ys[3] % vo == 0)) return -1;
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  const uint64_t dims[4] = {(uint64_t)a.c, (uint64_t)a.in_w, (uint64_t)a.in_h, (uint64_t)a.n};
  const uint64_t strides[3] = {(uint64_t)a.xs[3] * 2, (uint64_t)a.xs[2] * 2, (uint64_t)a.xs[0] * 2};
  const uint32_t box[4] = {(uint32_t)kClCb, (uint32_t)(kClTx + kF - 1), (uint32_t)(kClTy + kF - 1), 1u};
  const int rc = make_tensor_map(&tmap, dtype, 4, a.x, dims, strides, box, nullptr, 0);
  if (rc) return rc;
  const int tiles_x = (a.out_w + kClTx - 1) / kClTx, tiles_y = (a.out_h + kClTy - 1) / kClTy;
  const size_t smem = (size_t)(kClTy + kF - 1) * (kClTx + kF - 1) * kClCb * 2 + (2 * kClCb + kF * kF) * sizeof(float) + 16;
  const int ncb = (a.c + kClCb - 1) / kClCb;
  if ((int64_t)tiles_x * tiles_y * a.n * ncb > 0x7fffffffLL) return -1;
  dim3 grid((unsigned)((int64_t)tiles_x * tiles_y * a.n * ncb), 1u);
  static const bool use_sep = [] { const char* e = getenv("VCN_FIR_CL_SEP"); return !e || atoi(e) != 0; }();
  if constexpr (sizeof(T) == 2) {
    if (use_sep && y_dtype != VCN_F32) {
      if (!a.scale && !a.noise && !a.act) return launch_cl_sep<T, T, 0>(a, tmap, grid, smem, tiles_x, tiles_y, ncb, st);
      const bool homog = a.act == 0 || (a.act_gain > 0.f && (a.act == VCN_ACT_LINEAR || (a.act == VCN_ACT_LRELU && a.alpha >= 0.f && a.alpha <= 1.f)));
      if (homog) return launch_cl_sep<T, T, 1>(a, tmap, grid, smem, tiles_x, tiles_y, ncb, st);
      return launch_cl_sep<T, T, 2>(a, tmap, grid, smem, tiles_x, tiles_y, ncb, st);
    }
  }
  if (y_dtype == VCN_F32) {
    auto kern = upfirdn2d_cl_kernel<T, float>;
    VCN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, 256, smem, st>>>(a, tmap, tiles_x, tiles_y, ncb);
  } else {
    auto kern = upfirdn2d_cl_kernel<T, T>;
    VCN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, 256, smem, st>>>(a, tmap, tiles_x, tiles_y, ncb);
  }
  VCN_LAUNCH_CHECK("upfirdn2d_cl_kernel");
  return VCN_OK;
}


// ---------------------------------------------------------------------------------------------
// separable long filters: one 1-D pass along x (fh == 1) or along y (fw == 1), 5..32 taps, up / down in {1, 2} along the
// axis (the reference's 1-D specialisations upfirdn2d.cu:219-246,264-338 - AugmentPipe's 12-tap sym6 resampling,
// augment.py:167,290,301 - and the two passes of every `separable` setup_filter, upfirdn2d.py:239-240).  Dense NCHW,
// fp32 / fp16 / bf16.  The CTA stages the input segment of its output tile in shared memory (zero outside the image = the
// op's padding), every thread keeps the taps ITS output phase uses in registers (gain folded in, zero-extended to the
// template size) and reads consecutive shared-memory words: K = FWMAX / UP shared loads + FMAs per output, one coalesced
// global read and write per element.
// ---------------------------------------------------------------------------------------------
constexpr int k1dTW = 128, k1dTL = 16;   // x pass: 128 outputs x 16 lines per CTA
constexpr int k1dTX = 64, k1dTH = 64;    // y pass: 64 columns x 64 outputs per CTA

template <int FWMAX, int UP>
__device__ __forceinline__ void load_taps_1d(const UpfirdnArgs& a, int ntaps, int fx0, float (&w)[(FWMAX + UP - 1) / UP]) {
  // tap fx of the pass: the filter as given when `flip`, reversed otherwise (upfirdn2d.cu:125-127); gain folded in
#pragma unroll
  for (int k = 0; k < (FWMAX + UP - 1) / UP; ++k) {
    const int fx = fx0 + UP * k;
    w[k] = fx < ntaps ? __ldg(a.f + (a.flip ? fx : ntaps - 1 - fx)) * a.gain : 0.f;
  }
}

template <typename T, int FWMAX, int UP>
__global__ void __launch_bounds__(256) upfirdn1d_x_kernel(const UpfirdnArgs a, int down, int lines, int tiles_x) {
  constexpr int K = (FWMAX + UP - 1) / UP;
  constexpr int SEG = (k1dTW - 1) * 2 / UP + K + 4;
  constexpr int PITCH = SEG + 1;
  __shared__ float s[k1dTL * PITCH];
  __shared__ int64_t s_in[k1dTL], s_out[k1dTL];  // element offsets of the CTA's lines (-1: past the end)
  const int bx = blockIdx.x % tiles_x, bl = blockIdx.x / tiles_x;
  const int ox0 = bx * k1dTW, line0 = bl * k1dTL;
  const int seg_lo = floor_div(ox0 * down - a.padx0, UP);
  const T* __restrict__ px = static_cast<const T*>(a.x);
  T* __restrict__ py = static_cast<T*>(a.y);
  if (threadIdx.x < k1dTL) {
    const int line = line0 + threadIdx.x;
    int64_t oi = -1, oo = -1;
    if (line < lines) {
      const int y = line % a.in_h, pc = line / a.in_h;
      oi = (int64_t)(pc / a.c) * a.xs[0] + (int64_t)(pc % a.c) * a.xs[1] + (int64_t)y * a.xs[2];
      oo = (int64_t)(pc / a.c) * a.ys[0] + (int64_t)(pc % a.c) * a.ys[1] + (int64_t)y * a.ys[2];
    }
    s_in[threadIdx.x] = oi; s_out[threadIdx.x] = oo;
  }
  __syncthreads();
  // stage the input segments: one flat index space so that every thread has ~17 independent loads in flight
  constexpr int NLD = (k1dTL * SEG + 255) / 256;
  T v[NLD];
#pragma unroll
  for (int i = 0; i < NLD; ++i) {  // all loads first (independent), then the shared-memory writes
    const int e = threadIdx.x + i * 256;
    const int l = e / SEG, j = e - l * SEG;
    const int ix = seg_lo + j;
    const int64_t off = e < k1dTL * SEG ? s_in[l] : -1;
    v[i] = (off >= 0 && ix >= 0 && ix < a.in_w) ? px[off + ix] : T(0);
  }
#pragma unroll
  for (int i = 0; i < NLD; ++i) {
    const int e = threadIdx.x + i * 256;
    const int l = e / SEG, j = e - l * SEG;
    if (e < k1dTL * SEG) s[l * PITCH + j] = to_acc<T>(v[i]);
  }
  __syncthreads();
  const int tx = threadIdx.x & (k1dTW - 1), lg = threadIdx.x >> 7;
  const int ox = ox0 + tx;
  if (ox >= a.out_w) return;
  const int mx = ox * down - a.padx0;
  const int fx0 = ((-mx) % UP + UP) % UP;
  float w[K];
  load_taps_1d<FWMAX, UP>(a, a.fw, fx0, w);
  const int base = (mx + fx0) / UP - seg_lo;  // exact division: mx + fx0 is a multiple of UP
#pragma unroll 2
  for (int q = 0; q < k1dTL / 2; ++q) {
    const int l = lg * (k1dTL / 2) + q;
    const int64_t oo = s_out[l];
    if (oo < 0) break;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) acc = fmaf(w[k], s[l * PITCH + base + k], acc);
    py[oo + ox] = from_acc<T>(acc);
  }
}

template <typename T, int FWMAX, int UP>
__global__ void __launch_bounds__(256) upfirdn1d_y_kernel(const UpfirdnArgs a, int down, int tiles_x, int tiles_y) {
  constexpr int K = (FWMAX + UP - 1) / UP;
  constexpr int SEG = (k1dTH - 1) * 2 / UP + K + 4;
  __shared__ float s[SEG][k1dTX + 1];
  int b = blockIdx.x;
  const int bx = b % tiles_x; b /= tiles_x;
  const int by = b % tiles_y;
  const int pc = b / tiles_y;
  const int x0 = bx * k1dTX, oy0 = by * k1dTH;
  const int seg_lo = floor_div(oy0 * down - a.pady0, UP);
  // rows this tile reads, zero-extended taps included (they multiply by 0 but must not meet uninitialised NaN bit patterns)
  const int seg_n = min(SEG, ((min(k1dTH, a.out_h - oy0) - 1) * down) / UP + K + 2);
  const T* __restrict__ src = static_cast<const T*>(a.x) + (int64_t)(pc / a.c) * a.xs[0] + (int64_t)(pc % a.c) * a.xs[1];
  T* __restrict__ dst = static_cast<T*>(a.y) + (int64_t)(pc / a.c) * a.ys[0] + (int64_t)(pc % a.c) * a.ys[1];
  const int tx = threadIdx.x & (k1dTX - 1), tg = threadIdx.x >> 6;  // 4 row groups
  const int x = x0 + tx;
#pragma unroll 1
  for (int j0 = tg; j0 < seg_n; j0 += 32) {  // 8 independent loads per thread per round
    T v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int j = j0 + 4 * i, iy = seg_lo + j;
      v[i] = (j < seg_n && x < a.in_w && iy >= 0 && iy < a.in_h) ? src[(int64_t)iy * a.xs[2] + x] : T(0);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int j = j0 + 4 * i;
      if (j < seg_n) s[j][tx] = to_acc<T>(v[i]);
    }
  }
  __syncthreads();
  if (x >= a.out_w) return;
  float w[2][K];  // the two output phases of an up = 2 pass use different taps (UP = 1: both rows are the same)
  load_taps_1d<FWMAX, UP>(a, a.fh, 0, w[0]);
  load_taps_1d<FWMAX, UP>(a, a.fh, UP - 1, w[1]);
#pragma unroll 2
  for (int q = 0; q < k1dTH / 4; ++q) {
    const int oy = oy0 + tg + 4 * q;
    if (oy >= a.out_h) break;
    const int my = oy * down - a.pady0;
    const int fy0 = ((-my) % UP + UP) % UP;
    const int base = (my + fy0) / UP - seg_lo;
    float acc = 0.f;
    if (UP == 1 || fy0 == 0) {
#pragma unroll
      for (int k = 0; k < K; ++k) acc = fmaf(w[0][k], s[base + k][tx], acc);
    } else {
#pragma unroll
      for (int k = 0; k < K; ++k) acc = fmaf(w[1][k], s[base + k][tx], acc);
    }
    dst[(int64_t)oy * a.ys[2] + x] = from_acc<T>(acc);
  }
}

// Eligibility + launch of the 1-D kernels.  Returns -1 when the configuration is not covered.
template <typename T>
int try_launch_1d(const UpfirdnArgs& a, cudaStream_t st) {
  if (a.scale || a.noise || a.act) return -1;
  if (!(a.xs[3] == 1 && a.ys[3] == 1)) return -1;
  const bool xpass = a.fh == 1 && a.fw > kF && a.upy == 1 && a.downy == 1 && a.pady0 == 0 && a.out_h == a.in_h;
  const bool ypass = a.fw == 1 && a.fh > kF && a.upx == 1 && a.downx == 1 && a.padx0 == 0 && a.out_w == a.in_w;
  if (!xpass && !ypass) return -1;
  const int taps = xpass ? a.fw : a.fh, up = xpass ? a.upx : a.upy, down = xpass ? a.downx : a.downy;
  if (taps > 32 || up > 2 || down > 2) return -1;
  const int64_t planes = (int64_t)a.n * a.c;
#define VCN_1D(FW, UPV)                                                                                              \
  do {                                                                                                               \
    if (xpass) {                                                                                                     \
      const int64_t lines = planes * a.in_h;                                                                         \
      const int tiles_x = (a.out_w + k1dTW - 1) / k1dTW;                                                             \
      const int64_t blocks = (int64_t)tiles_x * ((lines + k1dTL - 1) / k1dTL);                                       \
      if (blocks > 0x7fffffffLL) return -1;                                                                          \
      upfirdn1d_x_kernel<T, FW, UPV><<<(unsigned)blocks, 256, 0, st>>>(a, down, (int)lines, tiles_x);               \
    } else {                                                                                                         \
      const int tiles_x = (a.out_w + k1dTX - 1) / k1dTX, tiles_y = (a.out_h + k1dTH - 1) / k1dTH;                    \
      const int64_t blocks = (int64_t)tiles_x * tiles_y * planes;                                                    \
      if (blocks > 0x7fffffffLL) return -1;                                                                          \
      upfirdn1d_y_kernel<T, FW, UPV><<<(unsigned)blocks, 256, 0, st>>>(a, down, tiles_x, tiles_y);                  \
    }                                                                                                                \
  } while (0)
  if (up == 1) { if (taps <= 12) VCN_1D(12, 1); else if (taps <= 24) VCN_1D(24, 1); else VCN_1D(32, 1); }
  else { if (taps <= 12) VCN_1D(12, 2); else if (taps <= 24) VCN_1D(24, 2); else VCN_1D(32, 2); }
#undef VCN_1D
  VCN_LAUNCH_CHECK("upfirdn1d_kernel");
  return VCN_OK;
}

inline int pow2_ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

template <typename T>
TileGeom make_geom(const UpfirdnArgs& a, int up, int down) {
  constexpr int V = Vec16<T>::N;
  constexpr int MC = V;
  TileGeom g{};
  const int tw_cap = down == 2 ? 64 : 128;
  g.tw = pow2_ceil(a.out_w); if (g.tw > tw_cap) g.tw = tw_cap; if (g.tw < MC) g.tw = MC;
  g.tx = g.tw / MC;
  const int max_threads = down == 2 ? 128 : 256;
  int ty_cap = max_threads / g.tx;
  g.th = pow2_ceil(a.out_h); if (g.th < kMR) g.th = kMR;
  if (g.th > kMR * ty_cap) g.th = kMR * ty_cap;
  g.ty = g.th / kMR;
  // Balanced tiles for fp32 outputs just above a multiple of the power-of-two tile (the 2^k+1 transposed-conv intermediate
  // as an OUTPUT: pattern A', the backward of the FIR after the up-sampling convolution): 257 = 3 x 88 columns and 6 x 44
  // rows = 18 nearly full tiles per plane instead of 3 x 9 = 27 tiles of 128 x 32 of which 11 carry one row or column.
  // Sized for the register-staged path STG 2 (<= 72 staged rows x 96 columns).
  if (sizeof(T) == 4 && up == 1 && down == 1 && a.out_w > tw_cap && a.out_h > 64) {
    const int cur_tiles = ((a.out_w + g.tw - 1) / g.tw) * ((a.out_h + g.th - 1) / g.th);
    const int nx = (a.out_w + 87) / 88;
    const int bw = ((a.out_w + nx - 1) / nx + MC - 1) / MC * MC;        // <= 88
    const int btx = bw / MC;
    int bty = max_threads / btx; if (bty > 17) bty = 17;                 // 17 * 4 + 3 = 71 staged rows
    const int ny = (a.out_h + kMR * bty - 1) / (kMR * bty);
    const int bh = ((a.out_h + ny - 1) / ny + kMR - 1) / kMR * kMR;
    if (nx * ny * 5 <= cur_tiles * 4 && btx * (bh / kMR) >= 192) {  // at least 20 % fewer tiles, most threads computing
      g.tw = bw; g.tx = btx; g.th = bh; g.ty = bh / kMR;
    }
  }
  // down = 2, fp32 (pattern B': FIR + decimation, the backward of the RGB up-sampling and the discriminator-side down-sampling):
  // a 32 x 32 output tile stages 66 x 66 inputs - the register-staged path STG 2 with 64 of the 256 threads filtering (the
  // 4-byte cp.async path it replaces was issue-bound at 128 threads: 0.33 of the HBM peak)
  bool reg_down2 = false;
  if (sizeof(T) == 4 && up == 1 && down == 2 && a.out_w >= 32 && a.out_h >= 32) {
    g.tw = 32; g.tx = 32 / MC; g.th = 32; g.ty = 32 / kMR;
    reg_down2 = true;
  }
  g.tp = reg_down2 ? 1 : max_threads / (g.tx * g.ty);
  const int planes = a.n * a.c;
  while (g.tp > 1 && g.tp / 2 >= planes && g.tx * g.ty * (g.tp / 2) >= 32) g.tp /= 2;
  const int span_x = (g.tw - 1) * down + kF, span_y = (g.th - 1) * down + kF;
  g.iw = (span_x + up - 1) / up + (up > 1 ? 1 : 0);
  g.ih = (span_y + up - 1) / up + (up > 1 ? 1 : 0);
  // pitch: 16-byte multiple, and wide enough for the last thread's vector window reads
  const int wc = ((MC - 1) * down + kF - 1) / up + 1;
  const int nv = (wc + V - 1) / V;
  int need = ((g.tw - MC) * down) / up + (up == 1 ? nv * V : wc);
  if (need < g.iw) need = g.iw;
  g.iwp = (need + V - 1) / V * V;
  g.tiles_x = (a.out_w + g.tw - 1) / g.tw;
  g.tiles_y = (a.out_h + g.th - 1) / g.th;
  g.tiles_p = (planes + g.tp - 1) / g.tp;
  return g;
}

template <typename T, int UP, int DOWN, int PHX, int PHY, int EPI>
int launch_tile_epi(const UpfirdnArgs& a, TileGeom g, int dtype, cudaStream_t st) {
  constexpr bool kCanReg = sizeof(T) == 4 && UP == 1;
  const size_t e = sizeof(T);
  const int planes = a.n * a.c;
  CUtensorMap tmap;
  memset(&tmap, 0, sizeof(tmap));
  // TMA needs the innermost box coordinate to be a multiple of 16 bytes (measured on B200: an unaligned inner
  // coordinate raises cudaErrorIllegalInstruction; negative / out-of-range coordinates are fine and zero-fill).
  // ix0 of tile bx is ceil((bx*tw*DOWN - padx0) / UP); it is aligned for every bx iff it is for bx = 0 and 1.
  constexpr int V = Vec16<T>::N;
  auto ix0_of = [&](int bx) {
    const int m = bx * g.tw * DOWN - a.padx0;
    return m >= 0 ? (m + UP - 1) / UP : -((-m) / UP);
  };
  const bool coord_ok = (((ix0_of(0) % V) + V) % V == 0) && (((ix0_of(1) % V) + V) % V == 0);
  g.use_tma = coord_ok && aligned16(a.x) && ((size_t)a.in_w * e) % 16 == 0 && g.iwp <= 256 && g.ih <= 256 && g.tp <= 256;
  if (g.use_tma) {
    const uint64_t dims[3] = {(uint64_t)a.in_w, (uint64_t)a.in_h, (uint64_t)planes};
    const uint64_t strides[2] = {(uint64_t)a.in_w * e, (uint64_t)a.in_w * a.in_h * e};
    const uint32_t box[3] = {(uint32_t)g.iwp, (uint32_t)g.ih, (uint32_t)g.tp};
    const int rc = make_tensor_map(&tmap, dtype, 3, a.x, dims, strides, box, nullptr, 0);
    if (rc) return rc;
  }
  g.vec_store = aligned16(a.y) && ((size_t)a.out_w * e) % 16 == 0;
  if (!g.use_tma && e == 4 && g.iwp > 160) { set_error("upfirdn2d: staged pitch %d exceeds the cp.async path", g.iwp); return VCN_ERR_UNSUPPORTED; }
  const size_t buf_bytes = ((size_t)g.tp * g.ih * g.iwp * e + 127) & ~(size_t)127;
  const size_t smem = 2 * buf_bytes + kF * kF * sizeof(float) + 16;
  const bool reg_ok = kCanReg && !g.use_tma && g.tp == 1 && g.tx * g.ty <= 256 && (g.tx * g.ty >= 192 || DOWN == 2);
  const int reg_path = !reg_ok ? 0 : (g.ih <= 40 && g.iwp <= 160) ? 1 : (g.ih <= 72 && g.iwp <= 96) ? 2 : 0;
  const int64_t tiles = (int64_t)g.tiles_x * g.tiles_y * g.tiles_p;
  if (tiles > 0x7fffffffLL) { set_error("upfirdn2d: too many tiles"); return VCN_ERR_INVALID; }
  if (reg_path) {
    if constexpr (kCanReg) {
      auto kern = reg_path == 1 ? upfirdn2d_tile_kernel<T, UP, DOWN, PHX, PHY, EPI, 1> : upfirdn2d_tile_kernel<T, UP, DOWN, PHX, PHY, EPI, 2>;
      static int per_sm_cache[3] = {0, 0, 0};  // resident CTAs per SM of each instantiation (registers bound it)
      int& per_sm = per_sm_cache[reg_path];
      if (per_sm == 0) {
        int nb = 0;
        VCN_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, 256, smem));
        per_sm = nb < 1 ? 1 : nb;
      }
      int64_t grid = (int64_t)sm_count() * per_sm;
      if (grid > tiles) grid = tiles;
      kern<<<(unsigned)grid, 256, smem, st>>>(a, g, tmap);
    }
  } else {
    auto kern = upfirdn2d_tile_kernel<T, UP, DOWN, PHX, PHY, EPI, 0>;
    if (smem > 48 * 1024) VCN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // persistent: as many CTAs as stay resident (at most 4 per SM; shared memory may allow fewer)
    int per_sm = (int)((200 * 1024) / (smem + 1024));
    if (per_sm > 4) per_sm = 4;
    if (per_sm < 1) per_sm = 1;
    int64_t grid = (int64_t)sm_count() * per_sm;
    if (grid > tiles) grid = tiles;
    kern<<<(unsigned)grid, g.tx * g.ty * g.tp, smem, st>>>(a, g, tmap);
  }
  VCN_LAUNCH_CHECK("upfirdn2d_tile_kernel");
  return VCN_OK;
}

template <typename T, int UP, int DOWN, int PHX, int PHY>
int launch_tile(const UpfirdnArgs& a, const TileGeom& g, int dtype, cudaStream_t st) {
  if (!a.scale && !a.noise && !a.act) return launch_tile_epi<T, UP, DOWN, PHX, PHY, 0>(a, g, dtype, st);
  if (a.act == 0 || a.act == VCN_ACT_LINEAR || a.act == VCN_ACT_LRELU) return launch_tile_epi<T, UP, DOWN, PHX, PHY, 1>(a, g, dtype, st);
  return launch_tile_epi<T, UP, DOWN, PHX, PHY, 2>(a, g, dtype, st);
}

template <typename T>
int run(const UpfirdnArgs& a, int dtype, cudaStream_t st) {
  const bool dense_in = a.xs[3] == 1 && a.xs[2] == a.in_w && a.xs[1] == (int64_t)a.in_h * a.in_w &&
                        (a.n == 1 || a.xs[0] == (int64_t)a.c * a.in_h * a.in_w);
  const bool dense_out = a.ys[3] == 1 && a.ys[2] == a.out_w && a.ys[1] == (int64_t)a.out_h * a.out_w &&
                         (a.n == 1 || a.ys[0] == (int64_t)a.c * a.out_h * a.out_w);
  const bool small_f = a.fw <= kF && a.fh <= kF;
  const bool sq = a.upx == a.upy && a.downx == a.downy;
  const int up = a.upx, down = a.downx;
  const bool combo = (up == 1 && down == 1) || (up == 2 && down == 1) || (up == 1 && down == 2);
  if constexpr (sizeof(T) == 2) {
    const int rc = try_launch_cl<T>(a, dtype, dtype, st);
    if (rc >= 0) return rc;
  }
  if constexpr (sizeof(T) != 8)
  if (dense_in && dense_out && small_f && sq && combo && a.out_w >= 4 && a.out_h >= 4) {
    const TileGeom g = make_geom<T>(a, up, down);
    if (up == 1 && down == 1) return launch_tile<T, 1, 1, 0, 0>(a, g, dtype, st);
    if (up == 1 && down == 2) return launch_tile<T, 1, 2, 0, 0>(a, g, dtype, st);
    const int phx = ((a.padx0 % 2) + 2) % 2, phy = ((a.pady0 % 2) + 2) % 2;
    if (phx == 0 && phy == 0) return launch_tile<T, 2, 1, 0, 0>(a, g, dtype, st);
    if (phx == 1 && phy == 0) return launch_tile<T, 2, 1, 1, 0>(a, g, dtype, st);
    if (phx == 0 && phy == 1) return launch_tile<T, 2, 1, 0, 1>(a, g, dtype, st);
    return launch_tile<T, 2, 1, 1, 1>(a, g, dtype, st);
  }
  if constexpr (sizeof(T) != 8) {  // separable long filters (one 1-D pass per call)
    const int rc = try_launch_1d<T>(a, st);
    if (rc >= 0) return rc;
  }
  const int64_t total = (int64_t)a.n * a.c * a.out_h * a.out_w;
  int64_t grid = (total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (grid > cap) grid = cap;
  const int channels_fastest = (a.ys[1] == 1 && a.c > 1) ? 1 : 0;
  upfirdn2d_generic_kernel<T><<<(unsigned)grid, 256, 0, st>>>(a, channels_fastest);
  VCN_LAUNCH_CHECK("upfirdn2d_generic_kernel");
  return VCN_OK;
}

}  // namespace
}  // namespace vcn

extern "C" int vcn_upfirdn2d(const void* x, const float* f, void* y, int dtype, int n, int c, int in_h, int in_w,
                             const int64_t xs[4], int out_h, int out_w, const int64_t ys[4], int fh, int fw, int upx,
                             int upy, int downx, int downy, int padx0, int padx1, int pady0, int pady1, int flip,
                             float gain, const vcn_upfirdn2d_epilogue* epi, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(dtype_size(dtype) != 0, "upfirdn2d: unknown dtype %d", dtype);
  VCN_REQUIRE(n >= 0 && c >= 0 && in_h >= 1 && in_w >= 1, "upfirdn2d: bad input shape [%d,%d,%d,%d]", n, c, in_h, in_w);
  VCN_REQUIRE(fh >= 1 && fw >= 1, "upfirdn2d: filter must be at least 1x1");
  VCN_REQUIRE(upx >= 1 && upy >= 1 && downx >= 1 && downy >= 1, "upfirdn2d: up/down factors must be >= 1");
  const int want_w = (in_w * upx + padx0 + padx1 - fw + downx) / downx;
  const int want_h = (in_h * upy + pady0 + pady1 - fh + downy) / downy;
  VCN_REQUIRE(want_w >= 1 && want_h >= 1, "upfirdn2d: output would be empty (%d x %d)", want_h, want_w);
  VCN_REQUIRE(out_w == want_w && out_h == want_h, "upfirdn2d: output shape %dx%d does not match expected %dx%d", out_h,
              out_w, want_h, want_w);
  if (n == 0 || c == 0) return VCN_OK;
  VCN_REQUIRE(x && f && y && xs && ys, "upfirdn2d: x, f, y and strides must not be NULL");
  VCN_REQUIRE((int64_t)n * c * out_h * out_w <= 0x7fffffffLL && (int64_t)n * c * in_h * in_w <= 0x7fffffffLL,
              "upfirdn2d: tensor too large");
  UpfirdnArgs a{};
  a.x = x; a.f = f; a.y = y;
  a.n = n; a.c = c; a.in_h = in_h; a.in_w = in_w; a.out_h = out_h; a.out_w = out_w;
  for (int i = 0; i < 4; ++i) { a.xs[i] = xs[i]; a.ys[i] = ys[i]; }
  a.fh = fh; a.fw = fw; a.upx = upx; a.upy = upy; a.downx = downx; a.downy = downy;
  a.padx0 = padx0; a.pady0 = pady0; a.flip = flip ? 1 : 0; a.gain = gain;
  if (epi) {
    a.scale = epi->scale; a.noise = epi->noise; a.noise_stride_n = epi->noise_stride_n; a.bias = epi->bias;
    a.act = epi->act; a.alpha = epi->alpha; a.act_gain = epi->act_gain; a.clamp = epi->clamp;
    VCN_REQUIRE(a.act == 0 || (a.act >= VCN_ACT_LINEAR && a.act <= VCN_ACT_SWISH), "upfirdn2d: bad epilogue act %d", a.act);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case VCN_F32: return run<float>(a, dtype, st);
    case VCN_F16: return run<__half>(a, dtype, st);
    case VCN_BF16: return run<__nv_bfloat16>(a, dtype, st);
    case VCN_F64: return run<double>(a, dtype, st);
  }
  return VCN_ERR_INVALID;
}
