// bias_act.cu - fused bias + activation + gain + clamp (forward, grad 1, grad 2) for sm_100a.
//
// Replaces bias_act_kernel<T,A> (stylegan2_ada/torch_utils/ops/bias_act.cu:23-147) and the separate
// `dx.sum(...)` bias-gradient reduction of bias_act.py:173,204.
//
// Layout: x is dense; viewed as [outer, size_b, step_b] in memory order, bias index = (i / step_b) % size_b.
// Kernels (all HBM-bound, 16-byte vector accesses, no shared-memory staging because nothing is reused):
//   rows  : step_b is a multiple of the vector width and >= kRowMin.  One CTA = one chunk of one
//           (outer, channel) row, so the bias is a CTA-uniform scalar and the bias gradient is a CTA
//           reduction (warp shuffles + one smem hop) written to a partial buffer, summed by `finalize`
//           in a fixed order (deterministic, no atomics).
//   flat  : any other aligned tensor; bias looked up per vector / per element.
//   scalar: unaligned base pointers.
//   chansum: bias gradient for the flat/scalar cases (second pass over the result, L2 resident at those sizes).
// Algorithmic bytes: forward 2*S*e; grad 1 and 2: 3*S*e (+S*e when both xref and yref are read).
#include "act_math.cuh"

#include <mutex>

namespace vcn {
namespace {

constexpr int kThreads = 256;
constexpr int64_t kRowMin = 1024;  // elements; below this rows are too short to give each CTA useful work

struct BiasActArgs {
  const void* x;
  const void* b;
  const void* xref;
  const void* yref;
  const void* dy;
  void* y;
  void* partial;  // acc-typed partial sums or NULL
  int64_t size_x, size_b, step_b;
  int act, grad;
  float alpha, gain, clamp;
};

template <typename T, int ACT, int UNROLL>
__global__ void __launch_bounds__(kThreads) bias_act_rows_kernel(const BiasActArgs a, const int chunks_per_row) {
  using S = typename Acc<T>::type;
  constexpr int V = Vec16<T>::N;
  const int64_t row = blockIdx.x / chunks_per_row;
  const int chunk = blockIdx.x - row * chunks_per_row;
  const int64_t row_vecs = a.step_b / V;
  const int64_t base = row * a.step_b;
  const T* __restrict__ px = static_cast<const T*>(a.x) + base;
  const T* __restrict__ pxr = a.xref ? static_cast<const T*>(a.xref) + base : nullptr;
  const T* __restrict__ pyr = a.yref ? static_cast<const T*>(a.yref) + base : nullptr;
  const T* __restrict__ pdy = a.dy ? static_cast<const T*>(a.dy) + base : nullptr;
  T* __restrict__ py = static_cast<T*>(a.y) + base;
  const S bias = a.b ? to_acc<T>(static_cast<const T*>(a.b)[row % a.size_b]) : S(0);
  const S alpha = S(a.alpha), gain = S(a.gain), clamp = S(a.clamp);
  const int grad = a.grad;

  int64_t vi[UNROLL];
  Vec16<T> vx[UNROLL], vxr[UNROLL], vyr[UNROLL], vdy[UNROLL];
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    vi[u] = (int64_t)chunk * (kThreads * UNROLL) + u * kThreads + threadIdx.x;
    if (vi[u] < row_vecs) vx[u] = ld_stream(px + vi[u] * V);
  }
  if (pxr) {
#pragma unroll
    for (int u = 0; u < UNROLL; ++u)
      if (vi[u] < row_vecs) vxr[u] = ld_stream(pxr + vi[u] * V);
  }
  if (pyr) {
#pragma unroll
    for (int u = 0; u < UNROLL; ++u)
      if (vi[u] < row_vecs) vyr[u] = ld_stream(pyr + vi[u] * V);
  }
  if (pdy) {
#pragma unroll
    for (int u = 0; u < UNROLL; ++u)
      if (vi[u] < row_vecs) vdy[u] = ld_stream(pdy + vi[u] * V);
  }

  S local = S(0);
#pragma unroll
  for (int u = 0; u < UNROLL; ++u) {
    if (vi[u] < row_vecs) {
      Vec16<T> out;
#pragma unroll
      for (int k = 0; k < V; ++k) {
        S x = to_acc<T>(vx[u].v[k]);
        S xr = pxr ? to_acc<T>(vxr[u].v[k]) : S(0);
        const S yr = pyr ? to_acc<T>(vyr[u].v[k]) : S(0);
        const S dy = pdy ? to_acc<T>(vdy[u].v[k]) : S(1);
        if (grad == 0) x += bias; else xr += bias;
        out.v[k] = from_acc<T>(act_eval<S, ACT>(x, xr, yr, dy, a.act, grad, alpha, gain, clamp));
        local += to_acc<T>(out.v[k]);
      }
      st_stream(py + vi[u] * V, out);
    }
  }

  if (a.partial) {
    __shared__ S red[kThreads / 32];
    local = warp_sum(local);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x < 32) {
      S v = threadIdx.x < kThreads / 32 ? red[threadIdx.x] : S(0);
      v = warp_sum(v);
      if (threadIdx.x == 0) static_cast<S*>(a.partial)[blockIdx.x] = v;
    }
  }
}

// partial is [outer, size_b, chunks]; db[b] = sum over outer and chunks, fixed order.
template <typename T>
__global__ void __launch_bounds__(128) bias_act_finalize_kernel(const void* partial, void* db, int64_t outer,
                                                                int64_t size_b, int chunks) {
  using S = typename Acc<T>::type;
  const int64_t b = blockIdx.x;
  const S* p = static_cast<const S*>(partial);
  S acc = S(0);
  const int64_t per = (int64_t)chunks;
  for (int64_t j = threadIdx.x; j < outer * per; j += blockDim.x) {
    const int64_t o = j / per, c = j - o * per;
    acc += p[(o * size_b + b) * per + c];
  }
  __shared__ S red[4];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) static_cast<T*>(db)[b] = from_acc<T>(red[0] + red[1] + red[2] + red[3]);
}

// bias lookup modes of the flat kernel
enum { kBiasNone = 0, kBiasPerVector = 1, kBiasContiguous = 2, kBiasPerElement = 3 };

template <typename T, int ACT>
__global__ void __launch_bounds__(kThreads) bias_act_flat_kernel(const BiasActArgs a, const int bias_mode) {
  using S = typename Acc<T>::type;
  constexpr int V = Vec16<T>::N;
  constexpr int U = 2;
  const T* __restrict__ px = static_cast<const T*>(a.x);
  const T* __restrict__ pb = static_cast<const T*>(a.b);
  const T* __restrict__ pxr = static_cast<const T*>(a.xref);
  const T* __restrict__ pyr = static_cast<const T*>(a.yref);
  const T* __restrict__ pdy = static_cast<const T*>(a.dy);
  T* __restrict__ py = static_cast<T*>(a.y);
  const S alpha = S(a.alpha), gain = S(a.gain), clamp = S(a.clamp);
  const int grad = a.grad;
  const int64_t nvec = a.size_x / V;
  const int64_t stride = (int64_t)gridDim.x * kThreads;

  for (int64_t v0 = (int64_t)blockIdx.x * kThreads + threadIdx.x; v0 < nvec; v0 += stride * U) {
    Vec16<T> vx[U], vxr[U], vyr[U], vdy[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t v = v0 + u * stride;
      if (v < nvec) {
        vx[u] = ld_stream(px + v * V);
        if (pxr) vxr[u] = ld_stream(pxr + v * V);
        if (pyr) vyr[u] = ld_stream(pyr + v * V);
        if (pdy) vdy[u] = ld_stream(pdy + v * V);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t v = v0 + u * stride;
      if (v >= nvec) continue;
      const int64_t e0 = v * V;
      S bias_u = S(0);
      int64_t cb = 0;
      if (bias_mode == kBiasPerVector) bias_u = to_acc<T>(pb[(e0 / a.step_b) % a.size_b]);
      else if (bias_mode == kBiasContiguous) cb = e0 % a.size_b;
      Vec16<T> out;
#pragma unroll
      for (int k = 0; k < V; ++k) {
        S bias = bias_u;
        if (bias_mode == kBiasContiguous) bias = to_acc<T>(pb[cb + k]);
        else if (bias_mode == kBiasPerElement) bias = to_acc<T>(pb[((e0 + k) / a.step_b) % a.size_b]);
        S x = to_acc<T>(vx[u].v[k]);
        S xr = pxr ? to_acc<T>(vxr[u].v[k]) : S(0);
        const S yr = pyr ? to_acc<T>(vyr[u].v[k]) : S(0);
        const S dy = pdy ? to_acc<T>(vdy[u].v[k]) : S(1);
        if (grad == 0) x += bias; else xr += bias;
        out.v[k] = from_acc<T>(act_eval<S, ACT>(x, xr, yr, dy, a.act, grad, alpha, gain, clamp));
      }
      st_stream(py + v * V, out);
    }
  }
  // ragged tail (< V elements), first CTA only
  if (blockIdx.x == 0) {
    const int64_t e = nvec * V + threadIdx.x;
    if (e < a.size_x) {
      const S bias = pb ? to_acc<T>(pb[(e / a.step_b) % a.size_b]) : S(0);
      S x = to_acc<T>(px[e]);
      S xr = pxr ? to_acc<T>(pxr[e]) : S(0);
      const S yr = pyr ? to_acc<T>(pyr[e]) : S(0);
      const S dy = pdy ? to_acc<T>(pdy[e]) : S(1);
      if (grad == 0) x += bias; else xr += bias;
      py[e] = from_acc<T>(act_eval<S, 0>(x, xr, yr, dy, a.act, grad, alpha, gain, clamp));
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) bias_act_scalar_kernel(const BiasActArgs a) {
  using S = typename Acc<T>::type;
  const T* __restrict__ px = static_cast<const T*>(a.x);
  const T* __restrict__ pb = static_cast<const T*>(a.b);
  const T* __restrict__ pxr = static_cast<const T*>(a.xref);
  const T* __restrict__ pyr = static_cast<const T*>(a.yref);
  const T* __restrict__ pdy = static_cast<const T*>(a.dy);
  T* __restrict__ py = static_cast<T*>(a.y);
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (int64_t e = (int64_t)blockIdx.x * kThreads + threadIdx.x; e < a.size_x; e += stride) {
    const S bias = pb ? to_acc<T>(pb[(e / a.step_b) % a.size_b]) : S(0);
    S x = to_acc<T>(px[e]);
    S xr = pxr ? to_acc<T>(pxr[e]) : S(0);
    const S yr = pyr ? to_acc<T>(pyr[e]) : S(0);
    const S dy = pdy ? to_acc<T>(pdy[e]) : S(1);
    if (a.grad == 0) x += bias; else xr += bias;
    py[e] = from_acc<T>(act_eval<S, 0>(x, xr, yr, dy, a.act, a.grad, S(a.alpha), S(a.gain), S(a.clamp)));
  }
}

// db[b] = sum_{o,s} y[(o*size_b + b)*step_b + s].  step_b == 1: CTA = 32 channels x 8 row lanes (coalesced across
// channels); otherwise CTA = one channel, threads over the flattened (o, s) index.
template <typename T>
__global__ void __launch_bounds__(kThreads) bias_act_chansum_kernel(const void* y, void* db, int64_t outer,
                                                                    int64_t size_b, int64_t step_b) {
  using S = typename Acc<T>::type;
  const T* __restrict__ p = static_cast<const T*>(y);
  __shared__ S red[kThreads];
  if (step_b == 1) {
    const int lane = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int64_t b = (int64_t)blockIdx.x * 32 + lane;
    S acc = S(0);
    if (b < size_b)
      for (int64_t o = ry; o < outer; o += kThreads / 32) acc += to_acc<T>(p[o * size_b + b]);
    red[threadIdx.x] = acc;
    __syncthreads();
    if (ry == 0 && b < size_b) {
      S t = S(0);
#pragma unroll
      for (int j = 0; j < kThreads / 32; ++j) t += red[j * 32 + lane];
      static_cast<T*>(db)[b] = from_acc<T>(t);
    }
  } else {
    const int64_t b = blockIdx.x;
    S acc = S(0);
    const int64_t total = outer * step_b;
    for (int64_t j = threadIdx.x; j < total; j += kThreads) {
      const int64_t o = j / step_b, s = j - o * step_b;
      acc += to_acc<T>(p[(o * size_b + b) * step_b + s]);
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      S t = S(0);
#pragma unroll
      for (int j = 0; j < kThreads / 32; ++j) t += red[j];
      static_cast<T*>(db)[b] = from_acc<T>(t);
    }
  }
}

struct Plan {
  bool rows;
  int unroll;
  int chunks_per_row;
  int64_t outer;
};

template <typename T> Plan make_plan(const BiasActArgs& a, bool aligned, bool want_db) {
  constexpr int V = Vec16<T>::N;
  Plan p{};
  const int64_t sb = a.size_b > 0 ? a.size_b : 1, st = a.step_b > 0 ? a.step_b : 1;
  p.outer = a.size_x / (sb * st);
  p.rows = aligned && (a.b != nullptr || want_db) && (st % V == 0) && st >= kRowMin && (a.size_x % (sb * st) == 0);
  if (p.rows) {
    const int64_t row_vecs = st / V;
    p.unroll = row_vecs >= 4 * kThreads ? 4 : (row_vecs >= 2 * kThreads ? 2 : 1);
    p.chunks_per_row = (int)((row_vecs + (int64_t)kThreads * p.unroll - 1) / ((int64_t)kThreads * p.unroll));
  }
  return p;
}

template <typename T, int ACT>
int launch_rows(const BiasActArgs& a, const Plan& p, cudaStream_t st) {
  const int64_t grid = p.outer * a.size_b * p.chunks_per_row;
  if (grid > 0x7fffffffLL) { set_error("bias_act: grid too large"); return VCN_ERR_INVALID; }
  if (p.unroll == 4) bias_act_rows_kernel<T, ACT, 4><<<(unsigned)grid, kThreads, 0, st>>>(a, p.chunks_per_row);
  else if (p.unroll == 2) bias_act_rows_kernel<T, ACT, 2><<<(unsigned)grid, kThreads, 0, st>>>(a, p.chunks_per_row);
  else bias_act_rows_kernel<T, ACT, 1><<<(unsigned)grid, kThreads, 0, st>>>(a, p.chunks_per_row);
  VCN_LAUNCH_CHECK("bias_act_rows_kernel");
  return VCN_OK;
}

template <typename T, int ACT>
int launch_flat(const BiasActArgs& a, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  int mode = kBiasNone;
  if (a.b) {
    if (a.step_b % V == 0) mode = kBiasPerVector;
    else if (a.step_b == 1 && a.size_b % V == 0) mode = kBiasContiguous;
    else mode = kBiasPerElement;
  }
  const int64_t nvec = a.size_x / V;
  int64_t grid = (nvec + kThreads * 2 - 1) / (kThreads * 2);
  const int64_t cap = (int64_t)sm_count() * 32;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  bias_act_flat_kernel<T, ACT><<<(unsigned)grid, kThreads, 0, st>>>(a, mode);
  VCN_LAUNCH_CHECK("bias_act_flat_kernel");
  return VCN_OK;
}

template <typename T>
int run(BiasActArgs a, void* db, void* ws, size_t ws_bytes, cudaStream_t st) {
  using S = typename Acc<T>::type;
  const bool aligned = aligned16(a.x) && aligned16(a.y) && (!a.xref || aligned16(a.xref)) &&
                       (!a.yref || aligned16(a.yref)) && (!a.dy || aligned16(a.dy));
  const Plan p = make_plan<T>(a, aligned, db != nullptr);
  int rc = VCN_OK;
  if (p.rows) {
    const size_t need = (size_t)(p.outer * a.size_b * p.chunks_per_row) * sizeof(S);
    if (db) {
      VCN_REQUIRE(ws != nullptr && ws_bytes >= need, "bias_act: workspace of %zu bytes required, got %zu", need, ws_bytes);
      a.partial = ws;
    }
    if (a.act == VCN_ACT_LRELU) rc = launch_rows<T, VCN_ACT_LRELU>(a, p, st);
    else if (a.act == VCN_ACT_LINEAR) rc = launch_rows<T, VCN_ACT_LINEAR>(a, p, st);
    else rc = launch_rows<T, 0>(a, p, st);
    if (rc) return rc;
    if (db) {
      bias_act_finalize_kernel<T><<<(unsigned)a.size_b, 128, 0, st>>>(ws, db, p.outer, a.size_b, p.chunks_per_row);
      VCN_LAUNCH_CHECK("bias_act_finalize_kernel");
    }
    return VCN_OK;
  }
  if (aligned) {
    if (a.act == VCN_ACT_LRELU) rc = launch_flat<T, VCN_ACT_LRELU>(a, st);
    else if (a.act == VCN_ACT_LINEAR) rc = launch_flat<T, VCN_ACT_LINEAR>(a, st);
    else if (a.act == VCN_ACT_RELU) rc = launch_flat<T, VCN_ACT_RELU>(a, st);  // conv + bias + ReLU of the frozen f / VGG
    else rc = launch_flat<T, 0>(a, st);
    if (rc) return rc;
  } else {
    int64_t grid = (a.size_x + kThreads - 1) / kThreads;
    const int64_t cap = (int64_t)sm_count() * 32;
    if (grid > cap) grid = cap;
    bias_act_scalar_kernel<T><<<(unsigned)grid, kThreads, 0, st>>>(a);
    VCN_LAUNCH_CHECK("bias_act_scalar_kernel");
  }
  if (db) {
    const unsigned grid = a.step_b == 1 ? (unsigned)((a.size_b + 31) / 32) : (unsigned)a.size_b;
    bias_act_chansum_kernel<T><<<grid, kThreads, 0, st>>>(a.y, db, p.outer, a.size_b, a.step_b);
    VCN_LAUNCH_CHECK("bias_act_chansum_kernel");
  }
  return VCN_OK;
}

}  // namespace
}  // namespace vcn

extern "C" {

size_t vcn_bias_act_workspace(int dtype, int64_t size_x, int64_t size_b, int64_t step_b) {
  using namespace vcn;
  if (size_b <= 0 || step_b <= 0 || size_x <= 0) return 0;
  const size_t e = dtype_size(dtype);
  if (e == 0) return 0;
  const int64_t V = 16 / (int64_t)e;
  if (step_b % V != 0 || step_b < kRowMin || size_x % (size_b * step_b) != 0) return 0;
  const int64_t row_vecs = step_b / V;
  const int unroll = row_vecs >= 4 * kThreads ? 4 : (row_vecs >= 2 * kThreads ? 2 : 1);
  const int64_t chunks = (row_vecs + (int64_t)kThreads * unroll - 1) / ((int64_t)kThreads * unroll);
  return (size_t)((size_x / step_b) * chunks) * (dtype == VCN_F64 ? 8 : 4);
}

int vcn_bias_act(const void* x, const void* b, const void* xref, const void* yref, const void* dy, void* y, void* db,
                 void* workspace, size_t workspace_bytes, int dtype, int grad, int act, float alpha, float gain,
                 float clamp, int64_t size_x, int64_t size_b, int64_t step_b, void* stream) {
  using namespace vcn;
  VCN_REQUIRE(size_x >= 0 && size_x <= 0x7fffffffLL, "bias_act: x is too large (%lld elements)", (long long)size_x);
  if (size_x == 0 && !db) return VCN_OK;
  VCN_REQUIRE(x != nullptr && y != nullptr, "bias_act: x and y must not be NULL");
  VCN_REQUIRE(grad >= 0 && grad <= 2, "bias_act: grad must be 0, 1 or 2 (got %d)", grad);
  VCN_REQUIRE(act >= VCN_ACT_LINEAR && act <= VCN_ACT_SWISH, "bias_act: unknown activation id %d", act);
  VCN_REQUIRE(dtype_size(dtype) != 0, "bias_act: unknown dtype %d", dtype);
  if (b) VCN_REQUIRE(size_b >= 1 && step_b >= 1, "bias_act: size_b and step_b must be >= 1 when b is given");
  if (db) VCN_REQUIRE(size_b >= 1 && step_b >= 1, "bias_act: db needs size_b and step_b");
  if (!b && !db) { size_b = 1; step_b = 1; }
  VCN_REQUIRE(grad != 2 || dy != nullptr, "bias_act: grad 2 needs dy");
  BiasActArgs a{x, b, xref, yref, dy, y, nullptr, size_x, size_b, step_b, act, grad, alpha, gain, clamp};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case VCN_F32: return run<float>(a, db, workspace, workspace_bytes, st);
    case VCN_F16: return run<__half>(a, db, workspace, workspace_bytes, st);
    case VCN_BF16: return run<__nv_bfloat16>(a, db, workspace, workspace_bytes, st);
    case VCN_F64: return run<double>(a, db, workspace, workspace_bytes, st);
  }
  return VCN_ERR_INVALID;
}

}  // extern "C"

// ---- allocation-free variant -----------------------------------------------------------------------------------
// The partial-sum scratch of the fused db reduction lives in a file-scope device array instead of a caller-provided
// workspace.  Every device that loads the library holds VCN_BIAS_ACT_STATIC_SCRATCH_BYTES of it until the process exits.
// In host code `g_bias_act_static_scratch` is only a stand-in for the symbol: its device address is looked up once per
// device with cudaGetSymbolAddress (with that device current) and that pointer is what the kernels receive.
namespace vcn {
namespace {
constexpr size_t kStaticScratchBytes = 1u << 20;  // 1 MiB == VCN_BIAS_ACT_STATIC_SCRATCH_BYTES in include/vcn_b200.h
constexpr int kMaxDevices = 64;
}  // namespace
}  // namespace vcn

static __device__ __align__(256) unsigned char g_bias_act_static_scratch[vcn::kStaticScratchBytes];

namespace vcn {
namespace {
std::mutex g_static_scratch_mu;
void* g_static_scratch_ptr[kMaxDevices] = {};

// Device address of the scratch array on the CURRENT device (0 + error text on failure).
int static_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_error("bias_act_static: cudaGetDevice: %s", cudaGetErrorString(e)); return VCN_ERR_CUDA; }
  VCN_REQUIRE(dev >= 0 && dev < kMaxDevices, "bias_act_static: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (!g_static_scratch_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_bias_act_static_scratch);
    if (e != cudaSuccess || !p) {
      set_error("bias_act_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return VCN_ERR_CUDA;
    }
    g_static_scratch_ptr[dev] = p;
  }
  *out = g_static_scratch_ptr[dev];
  return VCN_OK;
}
}  // namespace
}  // namespace vcn

extern "C" {

size_t vcn_bias_act_static_scratch_bytes(void) { return vcn::kStaticScratchBytes; }

int vcn_bias_act_static(const void* x, const void* b, const void* xref, const void* yref, const void* dy, void* y,
                        void* db, int dtype, int grad, int act, float alpha, float gain, float clamp, int64_t size_x,
                        int64_t size_b, int64_t step_b, void* stream) {
  using namespace vcn;
  size_t need = 0;
  void* scratch = nullptr;
  if (db) {
    need = vcn_bias_act_workspace(dtype, size_x, size_b, step_b);
    VCN_REQUIRE(need <= kStaticScratchBytes,
                "bias_act_static: this shape needs %zu bytes of reduction scratch, the built-in array holds %zu "
                "(use vcn_bias_act with a caller workspace)", need, (size_t)kStaticScratchBytes);
    // every slot the finalize pass reads is written by the rows kernel of the same call: nothing has to start at zero
    if (int rc = static_scratch_for_current_device(&scratch)) return rc;
  }
  return vcn_bias_act(x, b, xref, yref, dy, y, db, scratch, db ? kStaticScratchBytes : 0, dtype, grad, act, alpha,
                      gain, clamp, size_x, size_b, step_b, stream);
}

}  // extern "C"
