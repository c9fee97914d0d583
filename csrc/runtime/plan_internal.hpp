// Internals shared by the plan compiler (plan.cpp) and the launchers (launch_tile.cpp,
// launch_sweep.cpp). Not part of the public ABI.
#pragma once
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>

#include <vector>

#include "../core/error.hpp"
#include "../kernels/sweep3.hpp"
#include "../kernels/sweep5.hpp"
#include "../kernels/sweep_sobel.hpp"
#include "../kernels/stencil_mma.hpp"
#include "../kernels/tile_plan.hpp"

namespace b2i {
// kernels (csrc/kernels/*.cu)
cudaError_t launch_tile_chain(const TileParams& P, const CUtensorMap& src_map,
                              const CUtensorMap& dst_map, const CUtensorMap& push_up_map,
                              const CUtensorMap& push_dn_map, int grid, cudaStream_t stream);
cudaError_t tile_chain_occupancy(int smem_bytes, int* blocks_per_sm);
cudaError_t launch_pointwise(const PointwiseParams& P, int grid, cudaStream_t stream);
cudaError_t launch_sweep5(const Sweep5Params& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                          int grid, cudaStream_t stream, const CUtensorMap* src_map1 = nullptr,
                          const CUtensorMap* dst_map1 = nullptr);
cudaError_t launch_sweep_sobel(const SweepSobelParams& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                               int grid, cudaStream_t stream);
cudaError_t launch_sweep3(const SweepParams& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                          int grid, cudaStream_t stream);
cudaError_t launch_stencil_mma(const StencilMmaParams& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                               int grid, cudaStream_t stream);
int run_naive(const b2i_chain_t* chain, const uint8_t* src, int W, int H, int C, size_t src_pitch,
              uint8_t* dst, size_t dst_pitch, uint8_t* tmp0, uint8_t* tmp1, uint8_t* lut_scratch,
              cudaStream_t stream);

// One launch of a plan: a run of chain elements one kernel executes.
struct Segment {
  bool pointwise = false;
  int cin = 0, cout = 0, halo = 0;
  // tile segments
  std::vector<PhaseDesc> phases;
  int n_luts = 0;
  uint8_t luts[kMaxLuts][256];
  int raw_fix_border = -1;
  int chan_a = 0, chan_b = 0;  // widest plane each ping-pong buffer has to hold
  // the first phase's LUT as clamp((a v + b) >> s), if it is one (found once, at plan creation)
  int lut0_affine = 0;
  int32_t aff_a = 0, aff_b = 0, aff_s = 0;
  // a segment that is exactly one stencil (or the Sobel pair) with int8 weights and a border rule
  // other than REF: the tensor-core kernel (kernels/stencil_mma.cu) takes it
  int mma_ok = 0, mma_mode = 0, mma_R = 0, mma_shift = 0, mma_bias = 0, mma_border = 0;
  int8_t mma_w[2][5][5];
  // pointwise segments
  int pw_gray = 0, pw_expand = 0, pw_has_lut0 = 0, pw_has_lut1 = 0;
  uint8_t pw_lut0[256], pw_lut1[256];
};

// launch_tile.cpp
int env_int(const char* name, int dflt);
// The B200IMG_* switches are read once per process, at the first use of each call site: launches
// do not walk the environment (they are debugging / A-B switches, set before the process starts).
#define B2I_ENV(name, dflt) ([]() -> int { static const int v_ = ::b2i::env_int(name, dflt); return v_; }())
int align_up(int v, int a);
int device_sms(int dev, int* sms);
PFN_cuTensorMapEncodeTiled_v12000 tensor_map_encoder();
bool encode_map(CUtensorMap* m, uint64_t ptr, int64_t pitch, int64_t row_bytes, int rows, int box_w_bytes,
                int box_rows);
bool encode_map_sw128(CUtensorMap* m, uint64_t ptr, int64_t pitch, int64_t row_bytes, int rows, int box_rows);
void layout(const Segment& seg, int th, TileParams& P, int raw_stages = 2);
}  // namespace b2i

struct b2i_plan {
  b2i_chain_t chain;
  int cin = 0, cout = 0, halo = 0;
  std::vector<b2i::Segment> segs;
  // The same chain cut so that every stencil the sweep kernels take stands alone (empty if that is
  // the same cut). Two sweep launches beat one fused tile launch (8K gauss5,sobel: 290 us fused,
  // 89 + 62 us split), but only where the sweeps can run: b2i_cuda_run picks per launch.
  std::vector<b2i::Segment> segs_split;
  b2i_plan_info_t info;
};

namespace b2i {
// One launch of one segment: rows [y0, y1) of its result. The sweep launchers return 1 if they took
// the segment, 0 if it (or the buffers) does not qualify, -1 on a launch error.
int launch_tile_segment(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
                        const b2i_view_t& src, const b2i_view_t& dst, bool first, bool last);
int launch_pointwise_segment(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
                             const b2i_view_t& src, const b2i_view_t& dst);
int try_sweep(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1, const b2i_view_t& src,
              const b2i_view_t& dst, int sms, bool last);
int try_sweep5(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1, const b2i_view_t& src,
               const b2i_view_t& dst, int sms, bool first, bool last);
int try_sweep_sobel(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
                    const b2i_view_t& src, const b2i_view_t& dst, int sms, bool first, bool last);
int try_stencil_mma(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
                    const b2i_view_t& src, const b2i_view_t& dst, int sms, bool first, bool last);
// set by a sweep launcher that fused ROOT mode's completion flags into its kernel
extern thread_local bool g_wait_done_fused;
}  // namespace b2i

#define B2I_CUDA(expr)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess)                                                              \
      B2I_FAIL(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver             \
                   ? B2I_ERR_NO_DEVICE                                                  \
                   : B2I_ERR_CUDA,                                                      \
               "%s: %s", #expr, cudaGetErrorString(e_));                                \
  } while (0)
