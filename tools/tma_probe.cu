// Probe: which TMA 3-D box configurations run on this driver (debug aid, not product code).
#include <cstdio>
#include <cstring>
#include <vector>
#include "../viscoin-next_b200/csrc/common.cuh"
using namespace vcn;

__global__ void probe(const __grid_constant__ CUtensorMap tmap, float* out, int elems, int c0, int c1, int c2) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* tile = reinterpret_cast<float*>(smem);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + ((elems * 4 + 15) & ~15));
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  __syncthreads();
  if (threadIdx.x == 0) { mbar_expect_tx(bar, elems * 4); tma_load_3d(tile, &tmap, bar, c0, c1, c2); }
  mbar_wait(bar, 0);
  for (int i = threadIdx.x; i < elems; i += blockDim.x) out[i] = tile[i];
}

int run(int w, int h, int p, int bw, int bh, int bp, int c0, int c1, int c2) {
  std::vector<float> host((size_t)w * h * p);
  for (size_t i = 0; i < host.size(); ++i) host[i] = (float)i + 1;
  float *dx, *dout;
  cudaMalloc(&dx, host.size() * 4);
  cudaMemcpy(dx, host.data(), host.size() * 4, cudaMemcpyHostToDevice);
  int elems = bw * bh * bp;
  cudaMalloc(&dout, elems * 4);
  CUtensorMap tm; memset(&tm, 0, sizeof(tm));
  uint64_t dims[3] = {(uint64_t)w, (uint64_t)h, (uint64_t)p};
  uint64_t strides[2] = {(uint64_t)w * 4, (uint64_t)w * h * 4};
  uint32_t box[3] = {(uint32_t)bw, (uint32_t)bh, (uint32_t)bp};
  int rc = make_tensor_map(&tm, VCN_F32, 3, dx, dims, strides, box, nullptr, 0);
  if (rc) { printf("dims[%d,%d,%d] box[%d,%d,%d]: encode failed: %s\n", w, h, p, bw, bh, bp, vcn_last_error()); return 1; }
  probe<<<1, 128, elems * 4 + 64>>>(tm, dout, elems, c0, c1, c2);
  cudaError_t e = cudaDeviceSynchronize();
  printf("dims[%d,%d,%d] box[%d,%d,%d] at(%d,%d,%d): %s", w, h, p, bw, bh, bp, c0, c1, c2, cudaGetErrorName(e));
  if (e == cudaSuccess) {
    std::vector<float> o(elems);
    cudaMemcpy(o.data(), dout, elems * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int k = 0; k < bp; ++k) for (int j = 0; j < bh; ++j) for (int i = 0; i < bw; ++i) {
      int gi = c0 + i, gj = c1 + j, gk = c2 + k;
      float want = (gi >= 0 && gi < w && gj >= 0 && gj < h && gk >= 0 && gk < p) ? host[((size_t)gk * h + gj) * w + gi] : 0.f;
      if (o[((size_t)k * bh + j) * bw + i] != want) ++bad;
    }
    printf("  mismatches=%d", bad);
  }
  printf("\n");
  return e != cudaSuccess;
}

int main(int argc, char** argv) {
  int which = argc > 1 ? atoi(argv[1]) : 0;
  switch (which) {
    case 0: return run(64, 64, 8, 32, 16, 4, 0, 0, 0);
    case 1: return run(64, 64, 8, 32, 16, 4, -2, -3, 6);
    case 2: return run(8, 8, 6, 20, 19, 8, -2, -2, 0);
    case 3: return run(8, 8, 6, 8, 8, 4, 0, 0, 0);
    case 4: return run(8, 8, 6, 20, 8, 4, 0, 0, 0);
    case 5: return run(8, 8, 6, 8, 19, 4, 0, 0, 0);
    case 6: return run(8, 8, 6, 8, 8, 8, 0, 0, 0);
    case 7: return run(16, 16, 6, 20, 19, 8, -2, -2, 0);
    case 8: return run(32, 32, 8, 36, 36, 2, -2, -2, 0);
    case 9: return run(64, 64, 8, 32, 16, 4, 2, 0, 0);
    case 10: return run(64, 64, 8, 32, 16, 4, -4, 0, 0);
    case 11: return run(64, 64, 8, 32, 16, 4, 0, -3, 0);
    case 12: return run(64, 64, 8, 32, 16, 4, 0, 0, 6);
    case 13: return run(64, 64, 8, 32, 16, 4, 48, 0, 0);
    case 14: return run(64, 64, 8, 32, 16, 4, 0, 60, 0);
    case 15: return run(64, 64, 8, 32, 16, 4, -2, 0, 0);
    case 16: return run(64, 64, 8, 32, 16, 4, 0, 0, -1);
  }
  return 0;
}
