// b2i_cuda_run_image_noalloc: the whole-image run for callers that must not allocate.
//
// Who this is for. b2i_cuda_run_image needs two scratch images (and a small table block) as soon
// as a chain is split into more than one kernel launch, and it expects the caller to bring that
// memory. Some callers have no way to do so:
//   * code that runs inside another framework's execution callback (an inference-server
//     backend, a video-pipeline element, a plugin's enqueue hook) where the host framework owns
//     the device allocator and forbids allocation calls on the hot path;
//   * processes in which a caching allocator has already reserved practically all device memory,
//     so a late cudaMalloc fails or forces that allocator to release and fragment its pool;
//   * latency-bound loops that cannot afford the device-wide synchronisation cudaMalloc and
//     cudaFree imply.
// For them the scratch lives in a fixed array inside this library's device image. The price is
// stated up front: every GPU that loads libb200img.so carries B2I_NOALLOC_SCRATCH_BYTES
// (about 12 MiB) until the process exits, and images are limited to B2I_NOALLOC_MAX_PIXELS.
//
// There is nothing like this in the reference program: it calls cudaMalloc for every buffer it
// uses.
#include <cuda_runtime.h>

#include "../core/error.hpp"
#include "b200img.h"

namespace {

constexpr size_t kImageBytes = (size_t)B2I_NOALLOC_MAX_PIXELS * 3;  // one 3-channel scratch image
constexpr size_t kTableBytes = 256 * B2I_MAX_OPS;                   // LUT block of the per-op path
static_assert(2 * kImageBytes + kTableBytes == B2I_NOALLOC_SCRATCH_BYTES,
              "header constant and scratch layout disagree");

// The scratch itself. In host code this name is only a handle for the runtime's symbol table —
// NOT a device address; the address is looked up per device with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_scratch[B2I_NOALLOC_SCRATCH_BYTES];

// Device address of g_scratch on the current device, with that device current. Asked for on every
// call rather than remembered: the lookup allocates nothing once the module is loaded (the first
// call per device loads it), and a remembered address would dangle after cudaDeviceReset - which
// the reference itself calls (kernel.cu:241) - or any other teardown of the primary context.
int scratch_base(unsigned char** base) {
  void* p = nullptr;
  const cudaError_t e = cudaGetSymbolAddress(&p, g_scratch);
  if (e != cudaSuccess || !p) {
    int dev = -1;
    cudaGetDevice(&dev);
    B2I_FAIL(B2I_ERR_CUDA, "cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
  }
  *base = static_cast<unsigned char*>(p);
  return B2I_OK;
}

}  // namespace

// Keeps the array referenced from device code so that no toolchain drops it from the image.
__global__ void b2i_noalloc_scratch_anchor(unsigned long long* out) {
  if (out) *out = (unsigned long long)(size_t)&g_scratch[0];
}

// See include/b200img.h for the contract.
extern "C" int b2i_cuda_run_image_noalloc(b2i_plan_t* plan, uint64_t src, int64_t src_pitch,
                                          uint64_t dst, int64_t dst_pitch, int W, int H, void* stream,
                                          int flags) {
  B2I_CHECK_ARG(plan && src && dst, "null argument");
  B2I_CHECK_ARG(W > 0 && H > 0, "bad image size %dx%d", W, H);
  if ((long long)W * H > (long long)B2I_NOALLOC_MAX_PIXELS)
    B2I_FAIL(B2I_ERR_UNSUPPORTED,
             "image %dx%d has %lld pixels; b2i_cuda_run_image_noalloc is limited to %d pixels "
             "(its scratch is a fixed %zu-byte array) - use b2i_cuda_run_image with caller-provided "
             "scratch for larger images",
             W, H, (long long)W * H, B2I_NOALLOC_MAX_PIXELS, (size_t)B2I_NOALLOC_SCRATCH_BYTES);
  unsigned char* base = nullptr;
  B2I_TRY(scratch_base(&base));
  // Layout: [image 0][image 1][table block]. The two images are always written in full by one
  // launch before a later launch reads them, so whatever the previous call left there is never
  // observed; the table block is the one part that is reset here.
  cudaError_t e = cudaMemsetAsync(base + 2 * kImageBytes, 0, kTableBytes, (cudaStream_t)stream);
  if (e != cudaSuccess) B2I_FAIL(B2I_ERR_CUDA, "clearing the table block: %s", cudaGetErrorString(e));
  return b2i_cuda_run_image(plan, src, src_pitch, dst, dst_pitch, W, H, (uint64_t)(size_t)base,
                            (uint64_t)(size_t)(base + kImageBytes),
                            (uint64_t)(size_t)(base + 2 * kImageBytes), stream, flags);
}
