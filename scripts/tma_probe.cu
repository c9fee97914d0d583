// Probe: which TMA load shapes does this GPU/driver accept? usage: tma_probe boxw_bytes boxrows x y [swz]
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include "../csrc/kernels/device_utils.cuh"
using namespace b2i::dev;
__global__ void probe(const __grid_constant__ CUtensorMap m, int x, int y, int bytes, uint32_t* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  __syncthreads();
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar, bytes);
    tma_load_2d(smem + 1024, &m, x, y, bar);
  }
  mbar_wait(bar, 0);
  uint32_t s = 0;
  for (int i = threadIdx.x; i < bytes / 4; i += blockDim.x) s += reinterpret_cast<uint32_t*>(smem + 1024)[i];
  atomicAdd(out, s);
}
int main(int argc, char** argv) {
  int boxw = atoi(argv[1]), boxr = atoi(argv[2]), x = atoi(argv[3]), y = atoi(argv[4]);
  int W = 3072, H = 256;
  uint8_t* d; cudaMalloc(&d, (size_t)W * H); cudaMemset(d, 1, (size_t)W * H);
  uint32_t* out; cudaMalloc(&out, 4); cudaMemset(out, 0, 4);
  void* p; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  auto enc = (PFN_cuTensorMapEncodeTiled_v12000)p;
  CUtensorMap m;
  cuuint64_t gd[2] = {(cuuint64_t)W / 4, (cuuint64_t)H}; cuuint64_t gs[1] = {(cuuint64_t)W};
  cuuint32_t box[2] = {(cuuint32_t)boxw / 4, (cuuint32_t)boxr}; cuuint32_t es[2] = {1, 1};
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d, gd, gs, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode=%d ", (int)r);
  int smem = 1024 + boxw * boxr + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe<<<1, 128, smem>>>(m, x, y, boxw * boxr, out);
  cudaError_t e = cudaDeviceSynchronize();
  uint32_t h = 0; cudaMemcpy(&h, out, 4, cudaMemcpyDeviceToHost);
  printf("box %dx%d at (%d,%d): %s sum=%u (expect<=%u)\n", boxw, boxr, x, y, cudaGetErrorString(e), h, (unsigned)(boxw/4*boxr)*0x01010101u);
  return 0;
}
