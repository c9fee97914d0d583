// Kernel launch with the programmatic-dependent-launch attribute (see pdl_wait() in device_utils.cuh).
#pragma once
#include <cuda_runtime.h>

#include <cstdlib>
#include <utility>

namespace b2i {

// B200IMG_NO_PDL=1 launches everything fully serialized.
inline bool pdl_enabled() {
  static const bool on = [] {
    const char* v = getenv("B200IMG_NO_PDL");
    return !(v && *v && *v != '0');
  }();
  return on;
}

template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), int grid, int block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

}  // namespace b2i
