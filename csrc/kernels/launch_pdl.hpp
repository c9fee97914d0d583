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

// Launches with multi-GPU hooks (halo pushes, neighbour flags) on bands of fewer rows than this stay
// ordinary launches: measured at N = 2 with gauss5 on 7680-pixel rows, the dependent launch takes a pass
// over 2160-row bands from 33.5 to 31.6 us and one over 1080-row bands from 21.6 to 20.3 us, but one over
// 540-row bands from 17.4 to 19.0 us.
// (B200IMG_PDL_MIN_ROWS_MG overrides the threshold.)
inline int pdl_min_rows_mg() {
  static const int rows = [] {
    const char* v = getenv("B200IMG_PDL_MIN_ROWS_MG");
    return v && *v ? atoi(v) : 800;
  }();
  return rows;
}

// allow = false: an ordinary, fully serialized launch.
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(bool allow, void (*kern)(KArgs...), int grid, int block, size_t smem, cudaStream_t stream,
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
  cfg.numAttrs = allow && pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

}  // namespace b2i
