// Parameters of the sweep kernel (kernels/sweep3.cu), shared with the plan compiler.
#pragma once
#include <stdint.h>

namespace b2i {

constexpr int kSweepWarps = 8;       // warps per CTA, each an independent worker
// 2 x 8 warps per SM. 3 x 6 warps (96 registers) was measured too: 44.5 us vs 42.3 us on the 8K image.
constexpr int kSweepCtasPerSm = 2;
constexpr int kSweepStrip = 240;     // output pixels per warp row: 30 lanes x 8 (lanes 0, 31 are halo)
constexpr int kSweepInBytes = 800;   // TMA box width: 32 bytes left of the strip + 32 lanes x 24 bytes
constexpr int kSweepOutBytes = 720;  // 240 pixels x 3 channels
constexpr int kSweepChunk = 3;       // rows per TMA box (and per store box)
constexpr int kSweepStages = 3;      // input boxes in flight per warp

struct SweepParams {
  int32_t W, H;
  int32_t y0, y1;             // output rows of this launch
  int32_t src_y_begin;        // global row of row 0 of the source tensor map
  int32_t n_strips, seg_rows, n_tasks;
  int32_t gray;               // 1 = kernel.cu:39-42, 2 = cv::cvtColor
  int32_t rgb_order;
  int32_t has_lut0, has_lut1; // LUT after the gray conversion / after the stencil; has_lut0 == 2: the table is
                              // clamp((lut0_a * v + lut0_b) >> lut0_s), computed instead of looked up
  int32_t lut0_a, lut0_b, lut0_s;
  int32_t border, shift, bias, expand;
  int32_t warp_smem;          // bytes of dynamic shared memory per warp
  int32_t in_stage_bytes, out_buf_bytes;  // 128-byte aligned sizes of one input stage / one staging buffer
  // completion flags (multi-GPU ROOT mode): see sweep_finish in device_utils.cuh
  uint64_t done_flag, wait_done_flags, counters, status, timeout_ns;
  uint32_t wait_done_mask, signal_epoch;
  // replicated output: result rows leave by 8-byte stores from registers to this address (row y0
  // of an NVLS multicast image) instead of the staging buffer + TMA store; 0 = off
  uint64_t plain_dst_ptr;
  int64_t plain_dst_pitch;
  uint32_t wpk[3];            // 3x3 weights, one dp4a operand per stencil row
  uint8_t lut0[256], lut1[256];
};

}  // namespace b2i
