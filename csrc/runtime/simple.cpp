// Host-image convenience entry point for one GPU: upload, run, download, time.
//
// This is the reference's cudaMalloc -> cudaMemcpy H2D -> launches -> cudaDeviceSynchronize ->
// cudaMemcpy D2H sequence (kernel.cu:147-206) done properly: pinned staging, asynchronous
// copies on one stream, device-event timing of the kernel window. The multi-GPU runtime in
// context.cpp supersedes it for --gpus > 1; the CLI uses this for quick single-GPU runs.
#include <cuda_runtime.h>

#include <chrono>
#include <cstring>

#include "../core/error.hpp"

#define B2I_CUDA(expr)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess)                                                              \
      B2I_FAIL(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver             \
                   ? B2I_ERR_NO_DEVICE                                                  \
                   : B2I_ERR_CUDA,                                                      \
               "%s: %s", #expr, cudaGetErrorString(e_));                                \
  } while (0)

namespace {
struct Guard {  // frees whatever was allocated when a step fails half-way
  void* d[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  cudaStream_t s = nullptr;
  cudaEvent_t e[2] = {nullptr, nullptr};
  ~Guard() {
    for (void* p : d)
      if (p) cudaFree(p);
    for (cudaEvent_t ev : e)
      if (ev) cudaEventDestroy(ev);
    if (s) cudaStreamDestroy(s);
  }
};
}  // namespace

extern "C" int b2i_cuda_run_host_image(const b2i_chain_t* chain, const uint8_t* src, int W, int H,
                                       int C, size_t src_pitch, uint8_t* dst, size_t dst_pitch,
                                       int device, int iters, int flags, double* kernel_ms,
                                       double* total_ms) {
  B2I_CHECK_ARG(chain && src && dst && W > 0 && H > 0, "bad arguments");
  const int out_c = b2i_chain_out_channels(chain, C);
  if (out_c < 0) return out_c;
  if (iters < 1) iters = 1;
  const auto t0 = std::chrono::steady_clock::now();
  B2I_CUDA(cudaSetDevice(device));
  b2i_plan_t* plan = nullptr;
  B2I_TRY(b2i_plan_create(chain, C, &plan));
  struct PlanGuard {
    b2i_plan_t* p;
    ~PlanGuard() { b2i_plan_destroy(p); }
  } pg{plan};
  Guard g;
  const size_t in_pitch = ((size_t)W * C + 15) & ~size_t(15), out_pitch = ((size_t)W * out_c + 15) & ~size_t(15);
  const size_t tmp_bytes = (size_t)W * H * 3;
  B2I_CUDA(cudaMalloc(&g.d[0], in_pitch * H));
  B2I_CUDA(cudaMalloc(&g.d[1], out_pitch * H));
  B2I_CUDA(cudaMalloc(&g.d[2], tmp_bytes));
  B2I_CUDA(cudaMalloc(&g.d[3], tmp_bytes));
  B2I_CUDA(cudaMalloc(&g.d[4], 256 * B2I_MAX_OPS));
  B2I_CUDA(cudaMemset(g.d[4], 0, 256 * B2I_MAX_OPS));
  B2I_CUDA(cudaStreamCreateWithFlags(&g.s, cudaStreamNonBlocking));
  B2I_CUDA(cudaEventCreate(&g.e[0]));
  B2I_CUDA(cudaEventCreate(&g.e[1]));
  B2I_CUDA(cudaMemcpy2DAsync(g.d[0], in_pitch, src, src_pitch, (size_t)W * C, H,
                             cudaMemcpyHostToDevice, g.s));
  auto run_once = [&]() {
    return b2i_cuda_run_image(plan, (uint64_t)g.d[0], (int64_t)in_pitch, (uint64_t)g.d[1],
                              (int64_t)out_pitch, W, H, (uint64_t)g.d[2], (uint64_t)g.d[3],
                              (uint64_t)g.d[4], g.s, flags);
  };
  if (iters > 1) B2I_TRY(run_once());  // warm-up outside the timed window
  B2I_CUDA(cudaEventRecord(g.e[0], g.s));
  for (int i = 0; i < iters; ++i) B2I_TRY(run_once());
  B2I_CUDA(cudaEventRecord(g.e[1], g.s));
  B2I_CUDA(cudaMemcpy2DAsync(dst, dst_pitch, g.d[1], out_pitch, (size_t)W * out_c, H,
                             cudaMemcpyDeviceToHost, g.s));
  B2I_CUDA(cudaStreamSynchronize(g.s));
  float ms = 0.f;
  B2I_CUDA(cudaEventElapsedTime(&ms, g.e[0], g.e[1]));
  if (kernel_ms) *kernel_ms = ms / iters;
  if (total_ms)
    *total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return B2I_OK;
}
