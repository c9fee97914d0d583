// Probe: TMA store clipping. usage: tma_probe_store tensor_rows boxrows x y
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include "../csrc/kernels/device_utils.cuh"
using namespace b2i::dev;
__global__ void probe(const __grid_constant__ CUtensorMap m, int x, int y, int bytes) {
  extern __shared__ __align__(1024) uint8_t smem[];
  for (int i = threadIdx.x; i < bytes; i += blockDim.x) smem[i] = 7;
  fence_proxy_async_smem();
  __syncthreads();
  if (threadIdx.x == 0) { tma_store_2d(&m, x, y, smem); tma_store_commit(); tma_store_wait_all(); }
}
int main(int argc, char** argv) {
  int trows = atoi(argv[1]), boxr = atoi(argv[2]), x = atoi(argv[3]), y = atoi(argv[4]);
  int W = 3072, boxw = 768, H = 64;
  uint8_t* d; cudaMalloc(&d, (size_t)W * H); cudaMemset(d, 0, (size_t)W * H);
  void* p; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  auto enc = (PFN_cuTensorMapEncodeTiled_v12000)p;
  CUtensorMap m;
  cuuint64_t gd[2] = {(cuuint64_t)W / 4, (cuuint64_t)trows}; cuuint64_t gs[1] = {(cuuint64_t)W};
  cuuint32_t box[2] = {(cuuint32_t)boxw / 4, (cuuint32_t)boxr}; cuuint32_t es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d + 8 * W, gd, gs, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode=%d ", (int)r);
  int smem = boxw * boxr;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe<<<1, 128, smem>>>(m, x, y, smem);
  cudaError_t e = cudaDeviceSynchronize();
  std::vector<uint8_t> h((size_t)W * H);
  cudaMemcpy(h.data(), d, h.size(), cudaMemcpyDeviceToHost);
  printf("tensor rows %d box rows %d at (%d,%d): %s; rows written:", trows, boxr, x, y, cudaGetErrorString(e));
  for (int r2 = 0; r2 < H; ++r2) { int c = 0; for (int i = 0; i < W; ++i) c += h[(size_t)r2 * W + i] == 7; if (c) printf(" %d(%d)", r2 - 8, c); }
  printf("\n");
  return 0;
}
