// Parameters of the tensor-core stencil kernel (kernels/stencil_mma.cu), shared with the launcher.
#pragma once
#include <stdint.h>

namespace b2i {

constexpr int kSmThreads = 192;      // warp 0: TMA producer, warp 1: MMA issuer, warps 2-5: epilogue
constexpr int kSmRows = 128;         // output rows per band = M of the MMA = TMEM lanes
constexpr int kSmSlab = 128;         // input bytes per slab row = one 128-byte swizzle row = 4 K-steps
constexpr int kSmKStep = 32;         // input bytes per MMA (K of an 8-bit tcgen05.mma)
constexpr int kSmWin = 64;           // result columns one K-step touches: [k0 - 16, k0 + 48)
constexpr int kSmSlabRowsAlloc = 136;  // 128 + 2 * 2 rows, rounded up to whole 8-row swizzle groups
constexpr int kSmSlabBytes = kSmSlabRowsAlloc * kSmSlab;  // 17408 = 17 * 1024
constexpr int kSmStages = 6;         // slabs in flight per SM (102 KB: HBM latency x per-SM bandwidth)
constexpr int kSmBBytes = kSmWin * kSmKStep;              // one band matrix: 64 result cols x 32 input cols
constexpr int kSmMaxDy = 5;
constexpr int kSmVariants = 4;       // band matrices: interior, left edge, right edge (last two K-steps)
constexpr int kSmOutBytes = kSmRows * 128;                // staged result chunk: 128 rows x 128 bytes
constexpr int kSmRing = 512;         // TMEM columns: a ring of four 128-column result chunks

enum SmMode : int32_t {
  SM_LINEAR = 0,   // out = clamp((sum + bias) >> shift): any k x k correlation with int8 weights
  SM_SOBEL = 1     // out = clamp(|gx| + |gy|): two accumulators, each half of the ring
};

struct StencilMmaParams {
  int32_t W3, H;             // row bytes (W * channels), image rows
  int32_t y0, y1;            // output rows of this launch
  int32_t src_y_begin;       // global row of row 0 of the source tensor map
  int32_t cs;                // bytes between horizontally neighbouring taps: the channel count
  int32_t R;                 // stencil radius (1 or 2)
  int32_t border;            // REFLECT101, REPLICATE or ZERO
  int32_t mode, shift, bias;
  int32_t fast8;             // shift == 8, non-negative taps summing to 256: byte 1 of the sum is the result, no clamp
  int32_t n_bands, n_slabs;  // bands of 128 rows x slabs of 128 bytes
  int32_t dbg;               // bit 0: (experiment) descriptor base_offset = dy for row-shifted operands; wrong on B200
  uint64_t dbg_ptr;          // 0, or 32 x u64 of device memory for a timeline of blocks 0 and 1 (experiments)
  int8_t w[2][kSmMaxDy][kSmMaxDy];  // [accumulator][dy][dx] weights (accumulator 1: Sobel's gy)
  int8_t pad_[6];
};

}  // namespace b2i
