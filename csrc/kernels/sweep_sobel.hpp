// Parameters of the Sobel sweep kernel (kernels/sweep_sobel.cu), shared with the plan compiler.
#pragma once
#include <stdint.h>

namespace b2i {

constexpr int kSSWarps = 8;        // warps per CTA, each an independent worker
constexpr int kSSCtasPerSm = 2;
constexpr int kSSStrip = 240;      // output pixels per warp row: 30 lanes x 8 pixels (lanes 0, 31 are halo)
constexpr int kSSInBytes = 800;    // TMA box width: 32 bytes left of the strip + 32 lanes x 24 bytes
constexpr int kSSOutBytes = 720;   // 240 pixels x 3 channels
constexpr int kSSChunk = 4;        // rows per TMA box; even, so the two row-sum sets swap roles statically
constexpr int kSSStages = 2;       // input boxes in flight per warp

struct SweepSobelParams {
  int32_t W, H;
  int32_t y0, y1;            // output rows of this launch
  int32_t src_y_begin;       // global row of row 0 of the source tensor map
  int32_t n_strips, seg_rows, n_segs, n_tasks;
  int32_t border;            // REFLECT101, REPLICATE or ZERO (REF goes to the tile kernel)
  int32_t has_lut1;          // LUT applied to the result bytes
  int32_t warp_smem, in_stage_bytes, out_buf_bytes;
  // multi-GPU hooks, as in sweep5.hpp
  int32_t push_rows, push_pitch;
  int32_t n_up_tasks, n_dn_tasks;
  int32_t bseg, bseg_last;   // > 0: short first / last segments (one TMA box; the last takes the remainder), as in sweep5.hpp
  uint32_t wait_epoch, signal_epoch;
  uint64_t push_up_ptr, push_dn_ptr;
  uint64_t wait_up_flag, wait_dn_flag, signal_up_flag, signal_dn_flag;
  uint64_t counters, status, timeout_ns;
  // completion flags (multi-GPU ROOT mode): see sweep_finish in device_utils.cuh
  uint64_t done_flag, wait_done_flags;
  uint32_t wait_done_mask, reserved0;
  // replicated output: result rows leave by 8-byte stores from registers to this address (row y0
  // of an NVLS multicast image) instead of the staging buffer + TMA store; 0 = off
  uint64_t plain_dst_ptr;
  int64_t plain_dst_pitch;
  uint8_t lut1[256];
};

}  // namespace b2i
