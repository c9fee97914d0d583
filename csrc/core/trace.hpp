// Optional NVTX ranges (the reference has one wall-clock interval and nothing else).
// Off unless B200IMG_NVTX is set: a range is then pushed around every band launch and every
// multi-GPU run, visible in Nsight Systems / Nsight Compute timelines. Header-only NVTX3: no link
// dependency, and without a profiler attached the calls are no-ops.
#pragma once
#include <nvtx3/nvToolsExt.h>

#include <cstdlib>

namespace b2i {

inline bool nvtx_enabled() {
  static const bool on = [] {
    const char* v = getenv("B200IMG_NVTX");
    return v && *v && *v != '0';
  }();
  return on;
}

struct NvtxRange {
  bool active;
  explicit NvtxRange(const char* name) : active(nvtx_enabled()) {
    if (active) nvtxRangePushA(name);
  }
  ~NvtxRange() {
    if (active) nvtxRangePop();
  }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

}  // namespace b2i
