// api.cu - library-level entry points: status/error text, device probe, TMA descriptor factory.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace vcn {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error in %s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  return VCN_ERR_CUDA;
}

static unsigned long long g_launches = 0;
void count_launch() { __atomic_fetch_add(&g_launches, 1ULL, __ATOMIC_RELAXED); }

int sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached = n;
    cached_dev = dev;
  }
  return cached;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int make_tensor_map(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* elem_strides,
                    int swizzle_128b) {
  // Autograd worker threads may reach this point before the runtime has bound the primary context to the thread;
  // the driver entry point needs one (CUDA_ERROR_INVALID_CONTEXT otherwise).
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(nullptr);
    ctx_bound = true;
  }
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return VCN_ERR_CUDA;
  }
  CUtensorMapDataType dt;
  switch (dtype) {
    case VCN_F32: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32; break;
    case VCN_F16: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT16; break;
    case VCN_BF16: dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; break;
    case VCN_F64: dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT64; break;
    default: set_error("make_tensor_map: bad dtype %d", dtype); return VCN_ERR_INVALID;
  }
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = elem_strides ? elem_strides[i] : 1u;
  }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  // Immutable descriptor cache (per thread, direct-mapped): an encoded map is a pure function of the arguments below, and a
  // training loop presents the same (pointer, shape, box) tuples every iteration (caching allocator), so the steady state
  // of the eager path does no driver call here.  The full key is compared, a hash collision only costs a re-encode.
  struct Key {
    const void* base; uint64_t d[5], s[4]; uint32_t b[5], e[5]; int dtype, rank, swz;
  };
  struct Slot { Key k; CUtensorMap m; bool used; };
  constexpr int kSlots = 256;
  static thread_local Slot* cache = nullptr;
  if (!cache) cache = static_cast<Slot*>(calloc(kSlots, sizeof(Slot)));
  Key key;
  memset(&key, 0, sizeof(key));
  key.base = base; key.dtype = dtype; key.rank = rank; key.swz = swizzle_128b;
  for (int i = 0; i < rank; ++i) { key.d[i] = gd[i]; key.b[i] = bx[i]; key.e[i] = es[i]; }
  for (int i = 0; i + 1 < rank; ++i) key.s[i] = gs[i];
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < sizeof(Key) / 8; ++i) { h ^= reinterpret_cast<const uint64_t*>(&key)[i]; h *= 1099511628211ull; }
  Slot* slot = cache ? cache + (h >> 20) % kSlots : nullptr;
  if (slot && slot->used && memcmp(&slot->k, &key, sizeof(Key)) == 0) {
    memcpy(out, &slot->m, sizeof(CUtensorMap));
    return VCN_OK;
  }
  CUresult r = fn(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle_128b ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (CUresult %d): rank %d dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u]", (int)r,
              rank, (unsigned long long)gd[0], (unsigned long long)(rank > 1 ? gd[1] : 0),
              (unsigned long long)(rank > 2 ? gd[2] : 0), (unsigned long long)(rank > 3 ? gd[3] : 0), bx[0],
              rank > 1 ? bx[1] : 0, rank > 2 ? bx[2] : 0, rank > 3 ? bx[3] : 0);
    return VCN_ERR_CUDA;
  }
  if (slot) {
    slot->k = key;
    memcpy(&slot->m, out, sizeof(CUtensorMap));
    slot->used = true;
  }
  return VCN_OK;
}

}  // namespace vcn

extern "C" {

int vcn_abi_version(void) { return VCN_ABI_VERSION; }

const char* vcn_last_error(void) { return vcn::g_err; }

uint64_t vcn_launch_count(void) { return __atomic_load_n(&vcn::g_launches, __ATOMIC_RELAXED); }

int vcn_device_info(int device, int* sm_count, int64_t* l2_bytes, int* cc_major, int* cc_minor) {
  int n = 0;
  VCN_CUDA(cudaGetDeviceCount(&n));
  VCN_REQUIRE(device >= 0 && device < n, "device %d out of range (count %d)", device, n);
  int major = 0, minor = 0, sms = 0, l2 = 0;
  VCN_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  VCN_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  VCN_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  VCN_CUDA(cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, device));
  if (sm_count) *sm_count = sms;
  if (l2_bytes) *l2_bytes = l2;
  if (cc_major) *cc_major = major;
  if (cc_minor) *cc_minor = minor;
  if (major != 10) {
    vcn::set_error("device %d is sm_%d%d; libvcn_b200 holds sm_100a code only", device, major, minor);
    return VCN_ERR_DEVICE;
  }
  return VCN_OK;
}

}  // extern "C"
