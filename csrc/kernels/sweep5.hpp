// Parameters of the separable sweep kernel (kernels/sweep5.cu), shared with the plan compiler.
#pragma once
#include <stdint.h>

namespace b2i {

constexpr int kS5Warps = 14;       // warps per CTA; 1 CTA per SM (shared memory bound): 4 + 4 + 3 + 3 on the schedulers (two CTAs of 7 measured 3 % slower)
constexpr int kS5CtasPerSm = 1;
constexpr int kS5Strip = 240;      // output pixels per warp row: 30 lanes x 8 pixels (lanes 0, 31 are halo)
constexpr int kS5InBytes = 800;    // TMA box width: 32 bytes left of the strip + 32 lanes x 24 bytes
constexpr int kS5OutBytes = 720;   // 240 pixels x 3 channels
constexpr int kS5Chunk = 5;        // rows per TMA box = period of the accumulator rotation
constexpr int kS5Stages = 2;       // input boxes in flight per warp

struct Sweep5Params {
  int32_t W, H;
  int32_t y0, y1;            // output rows of this launch
  int32_t src_y_begin;       // global row of row 0 of the source tensor map
  int32_t n_strips, seg_rows, n_tasks;
  int32_t border;            // REFLECT101, REPLICATE or ZERO (REF goes to the tile kernel)
  int32_t bias;              // result = (sum + bias) >> 8: the planner scales the vertical taps so that all taps sum to 256
  int32_t binom;             // the horizontal taps are 1 4 6 4 1 (or 1 2 1): adds and shifts instead of multiplies
  int32_t has_lut1;          // LUT applied to the result bytes
  int32_t warp_smem, in_stage_bytes, out_buf_bytes;
  int32_t wh[5], wv[5];      // horizontal / vertical taps (a 3-tap kernel has zeros at both ends)
  // multi-GPU hooks (all zero on a single GPU): the first / last push_rows result rows are also
  // stored into the neighbours' ghost rows, guarded by the same epoch flags as the tile kernel
  int32_t push_rows, push_pitch;
  int32_t n_segs, n_up_tasks, n_dn_tasks;  // tasks that count into the up / down signal
  uint32_t wait_epoch, signal_epoch;
  uint64_t push_up_ptr, push_dn_ptr;       // where row y0 / row y1 - push_rows lands at the neighbour
  uint64_t wait_up_flag, wait_dn_flag, signal_up_flag, signal_dn_flag;
  uint64_t counters, status, timeout_ns;
  // completion flags (multi-GPU ROOT mode): see sweep_finish in device_utils.cuh
  uint64_t done_flag, wait_done_flags;
  uint32_t wait_done_mask, reserved0;
  // replicated output: result rows leave by 8-byte stores from registers to this address (row y0
  // of an NVLS multicast image) instead of the staging buffer + TMA store; 0 = off
  uint64_t plain_dst_ptr;
  int64_t plain_dst_pitch;
  // resident multi-pass launch (n_passes > 1): the kernel itself walks the passes, ping-ponging
  // between two band buffers (tensor maps src_map/dst_map on even passes, src_map1/dst_map1 on odd
  // ones). A task of pass p starts once the (up to nine) tasks around it have finished pass p-1:
  // task_done[t] holds epoch_base + passes finished. Neighbour GPUs are waited for from pass 1 on
  // (wait_epoch + p) and signalled after every pass but the last (signal_epoch + p).
  int32_t n_passes;
  int32_t strip_flags;       // the neighbour flags are per-strip arrays (B2I_LAUNCH_STRIP_FLAGS)
  int32_t bseg, bseg_last;   // > 0: the first segment is bseg rows and the last bseg_last (short, so that halos leave
                             // early); the segments between are whole TMA boxes, the last one takes the remainder
  uint64_t task_done;                      // u32 per task in local memory, only ever growing
  uint64_t push_up_ptr_odd, push_dn_ptr_odd;  // the push targets of odd passes (the other buffer)
  uint8_t lut1[256];
};

}  // namespace b2i
