// Thread-block clusters sharing the overlapping rows of neighbouring tiles: the experiment.
//
// A stencil tile of T rows needs T + 2R input rows, so vertically neighbouring tiles read 2R rows
// twice. Two ways to load them:
//
//   plain    every CTA fetches its own (T + 2R)-row box with one TMA load (what the tile kernel does;
//            the second read of the shared rows is an L2 hit);
//   cluster  two CTAs of a (1, 2, 1) cluster own two stacked tiles. Each fetches the T rows only it
//            needs; the 2R rows both need are fetched ONCE by the upper CTA with a multicast TMA load
//            that lands in both CTAs' shared memory (same offset in each) and completes bytes on both
//            CTAs' mbarriers.
//
// The kernel is made load-bound on purpose (out[y] = in[y - R] ^ in[y + R], one LOP3 per 16 bytes) so
// that any saving in L2 -> SM traffic shows as time. b2i_cuda_cluster_probe allocates its own
// buffers with cudaMalloc on the current device, so the tensor maps cover this GPU's memory only:
// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from
// another GPU's memory hung the GPU.
//
// Reference: kernel.cu:195 launches one thread per pixel with no tiles at all; the north-star item this
// closes is "clusters sharing overlapping tiles" (DESIGN.md section 4 has the numbers).
#include <cuda_runtime.h>

#include <vector>

#include "../runtime/plan_internal.hpp"
#include "b200img.h"
#include "device_utils.cuh"

namespace b2i {
using namespace dev;

namespace {

constexpr int kCpWidth = 256;   // tile width in bytes
constexpr int kCpThreads = 128;

struct ClusterProbeParams {
  int32_t W3, H, T, R;
  uint64_t dst;
  int64_t dst_pitch;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 2-D tile global -> the same shared-memory offset of every CTA in `mask`; completes the bytes on the
// mbarrier at the same offset in each of them.
__device__ __forceinline__ void tma_load_2d_mcast(void* dst, const CUtensorMap* m, int x, int y, uint64_t* bar,
                                                  uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(x), "r"(y), "r"(smem_u32(bar)), "h"(mask)
      : "memory");
}

// CLUSTER = false: map_full has boxes of T + 2R rows. CLUSTER = true: map_priv has boxes of T rows,
// map_shared of 2R rows; blockIdx.y = 2 * pair + rank.
template <bool CLUSTER>
__global__ void __launch_bounds__(kCpThreads)
cluster_probe_kernel(const ClusterProbeParams P, const __grid_constant__ CUtensorMap map_full,
                     const __grid_constant__ CUtensorMap map_priv, const __grid_constant__ CUtensorMap map_shared) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  const int T = P.T, R = P.R;
  uint8_t* shared_rows = smem;                       // 2R rows (cluster variant), at the same offset in both CTAs
  uint8_t* priv_rows = smem + 2 * R * kCpWidth;      // T rows (cluster) / T + 2R rows from offset 0 (plain)
  const int x0 = blockIdx.x * kCpWidth;
  const int y0 = blockIdx.y * T;                     // first output row of this CTA's tile
  const uint32_t rank = CLUSTER ? cluster_ctarank() : 0;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    mbar_arrive_expect_tx(&bar, (uint32_t)((T + 2 * R) * kCpWidth));
  }
  if (CLUSTER) cluster_sync();   // both barriers are armed before the multicast load is issued
  else __syncthreads();
  if (threadIdx.x == 0) {
    if (!CLUSTER) {
      tma_load_2d(smem, &map_full, x0 / 4, y0 - R, &bar);
    } else if (rank == 0) {
      tma_load_2d(priv_rows, &map_priv, x0 / 4, y0 - R, &bar);                        // rows [y0 - R, y0 + T - R)
      tma_load_2d_mcast(shared_rows, &map_shared, x0 / 4, y0 + T - R, &bar, 0x3);    // rows [y0 + T - R, y0 + T + R): both CTAs
    } else {
      tma_load_2d(priv_rows, &map_priv, x0 / 4, y0 + R, &bar);                        // rows [y0 + R, y0 + T + R)
    }
  }
  mbar_wait(&bar, 0);
  // tile row j (input row y0 - R + j) -> shared-memory row
  auto row_ptr = [&](int j) -> const uint8_t* {
    if (!CLUSTER) return smem + j * kCpWidth;
    if (rank == 0) return j < T ? priv_rows + j * kCpWidth : shared_rows + (j - T) * kCpWidth;
    return j < 2 * R ? shared_rows + j * kCpWidth : priv_rows + (j - 2 * R) * kCpWidth;
  };
  // 16 threads x 16 bytes per row, 8 rows per pass
  const int c16 = threadIdx.x & 15, r0 = threadIdx.x >> 4;
  const int xb = x0 + c16 * 16;
  if (xb < P.W3) {
    for (int i = r0; i < T; i += kCpThreads / 16) {
      const int y = y0 + i;
      if (y >= P.H) break;
      const uint4 a = *reinterpret_cast<const uint4*>(row_ptr(i) + c16 * 16);            // input row y - R
      const uint4 b = *reinterpret_cast<const uint4*>(row_ptr(i + 2 * R) + c16 * 16);    // input row y + R
      *reinterpret_cast<uint4*>(P.dst + (uint64_t)((int64_t)y * P.dst_pitch + xb)) =
          make_uint4(a.x ^ b.x, a.y ^ b.y, a.z ^ b.z, a.w ^ b.w);
    }
  }
  if (CLUSTER) cluster_sync();   // neither CTA leaves while the other may still be written into / read from
}

}  // namespace
}  // namespace b2i

// Runs both variants on a W3-byte x H-row image it allocates itself; returns the time per launch of
// each (CUDA events around `iters` launches, after a warm-up) and the number of wrong output bytes.
extern "C" int b2i_cuda_cluster_probe(int W3, int H, int T, int R, int iters, float* us_plain, float* us_cluster,
                                      long long* mismatches) {
  using namespace b2i;
  B2I_CHECK_ARG(us_plain && us_cluster && mismatches, "null argument");
  B2I_CHECK_ARG(W3 >= kCpWidth && W3 % 16 == 0 && H >= 8 && T >= 8 && T % 8 == 0 && T <= 128 && R >= 1 && R <= 4 && iters >= 1,
                "bad probe shape W3=%d H=%d T=%d R=%d iters=%d", W3, H, T, R, iters);
  uint8_t *src = nullptr, *dst = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  const size_t bytes = (size_t)W3 * H;
  auto cleanup = [&]() {
    if (src) cudaFree(src);
    if (dst) cudaFree(dst);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
  };
#define CP_CUDA(x)                                                                               \
  do {                                                                                           \
    cudaError_t e_ = (x);                                                                        \
    if (e_ != cudaSuccess) {                                                                     \
      cleanup();                                                                                 \
      return b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "%s: %s", #x, cudaGetErrorString(e_)); \
    }                                                                                            \
  } while (0)
  CP_CUDA(cudaMalloc(&src, bytes));   // this GPU's memory: the only kind a multicast TMA load may read here
  CP_CUDA(cudaMalloc(&dst, bytes));
  std::vector<uint8_t> h(bytes), out(bytes);
  uint32_t s = 12345u;
  for (size_t i = 0; i < bytes; ++i) {
    s = s * 1664525u + 1013904223u;
    h[i] = (uint8_t)(s >> 24);
  }
  CP_CUDA(cudaMemcpy(src, h.data(), bytes, cudaMemcpyHostToDevice));
  CUtensorMap map_full, map_priv, map_shared;
  if (!encode_map(&map_full, (uint64_t)src, W3, W3, H, kCpWidth, T + 2 * R) ||
      !encode_map(&map_priv, (uint64_t)src, W3, W3, H, kCpWidth, T) ||
      !encode_map(&map_shared, (uint64_t)src, W3, W3, H, kCpWidth, 2 * R)) {
    cleanup();
    return b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "tensor map encoding failed");
  }
  ClusterProbeParams P{W3, H, T, R, (uint64_t)dst, W3};
  const int gx = (W3 + kCpWidth - 1) / kCpWidth, gy = ((H + T - 1) / T + 1) / 2 * 2;
  const int smem_bytes = (T + 2 * R) * kCpWidth;
  CP_CUDA(cudaFuncSetAttribute(cluster_probe_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  CP_CUDA(cudaFuncSetAttribute(cluster_probe_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  CP_CUDA(cudaEventCreate(&e0));
  CP_CUDA(cudaEventCreate(&e1));
  *mismatches = 0;
  for (int variant = 0; variant < 2; ++variant) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(gx, gy, 1);
    cfg.blockDim = dim3(kCpThreads, 1, 1);
    cfg.dynamicSmemBytes = smem_bytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = variant ? 2 : 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    auto launch = [&]() {
      return variant ? cudaLaunchKernelEx(&cfg, cluster_probe_kernel<true>, P, map_full, map_priv, map_shared)
                     : cudaLaunchKernelEx(&cfg, cluster_probe_kernel<false>, P, map_full, map_priv, map_shared);
    };
    CP_CUDA(cudaMemset(dst, 0, bytes));
    for (int i = 0; i < 3; ++i) CP_CUDA(launch());
    CP_CUDA(cudaDeviceSynchronize());
    CP_CUDA(cudaEventRecord(e0));
    for (int i = 0; i < iters; ++i) CP_CUDA(launch());
    CP_CUDA(cudaEventRecord(e1));
    CP_CUDA(cudaEventSynchronize(e1));
    float ms = 0;
    CP_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    *(variant ? us_cluster : us_plain) = ms * 1000.0f / iters;
    CP_CUDA(cudaMemcpy(out.data(), dst, bytes, cudaMemcpyDeviceToHost));
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W3; ++x) {
        const uint8_t a = y - R >= 0 ? h[(size_t)(y - R) * W3 + x] : 0, b = y + R < H ? h[(size_t)(y + R) * W3 + x] : 0;
        if (out[(size_t)y * W3 + x] != (uint8_t)(a ^ b)) ++*mismatches;
      }
  }
  cleanup();
#undef CP_CUDA
  return B2I_OK;
}
