// Device-executable form of a filter chain segment: what the tile-chain kernel interprets.
// Built on the host by runtime/plan.cpp, passed by value as a kernel parameter (< 4 KB).
//
// One segment = the part of a chain that one kernel launch executes with every intermediate
// image held in shared memory. The reference runs one launch per filter with the gray image
// bouncing through global memory (kernel.cu:192-195); here a tile of the source is loaded once
// (TMA), walks through the phases in shared memory and leaves as a tile of the result.
#pragma once
#include <stdint.h>

namespace b2i {

constexpr int kTileW = 256;            // output tile width in pixels
// Horizontal apron of a plane in pixels. 16, not 8: TMA wants the first byte of a box on a
// 16-byte boundary in global memory (a box starting 24 bytes left of a tile faults with "illegal
// instruction"; measured, scripts/tma_probe.cu), and 16 pixels are 16 or 48 bytes.
constexpr int kApron = 16;
constexpr int kMarginCols = 8;         // columns beyond the tile a non-final phase computes per side
constexpr int kPlaneW = kTileW + 2 * kApron;  // 288
constexpr int kMaxTileH = 64;
constexpr int kTileThreads = 256;
constexpr int kMaxPhases = 6;
constexpr int kMaxLuts = 4;
constexpr int kMaxSegHalo = kMarginCols;    // sum of stencil radii one segment can absorb
constexpr int kRowGroup = 4;           // rows a thread walks in a stencil phase
constexpr int kGhostRows = 8;          // ghost rows above/below a band buffer (>= kMaxSegHalo)

enum PhaseType : int32_t {
  PH_GRAY_REF = 1,  // 3 -> 1 (kernel.cu:39-42)
  PH_GRAY_CV = 2,   // 3 -> 1 (kern.cpp:73)
  PH_COPY = 3,      // C -> C, optional LUT
  PH_EXPAND = 4,    // 1 -> 3, optional LUT (kernel.cu:210 when it is not fused into a stencil)
  PH_STENCIL = 5,   // k x k integer correlation (kernel.cu:64-94, kern.cpp:75)
  PH_SOBEL = 6,     // |gx| + |gy|
  PH_SEPH = 7,      // separable stencil, horizontal pass: u8 plane -> plane of 16-bit lanes
  PH_SEPV = 8       // separable stencil, vertical pass: 16-bit lanes -> u8, rounded and shifted
};

enum BufId : int32_t { BUF_RAW = 0, BUF_A = 1, BUF_B = 2, BUF_STAGE = 3 };

struct PhaseDesc {
  int32_t type;
  int32_t cin, cout;      // channels of the input plane / of what is written
  int32_t src, dst;       // BufId
  int32_t margin;         // rows around the tile this phase must still produce
  int32_t margin_x;       // pixels left/right of the tile this phase must still produce
  int32_t lut;            // LUT slot applied to the phase's result bytes, or -1
  int32_t expand;         // stencil phases: write each result byte three times (gray -> BGR)
  int32_t border;         // b2i_border of this stencil
  int32_t fix_border;     // border rule to paint onto out-of-image cells of dst, or -1
  int32_t ksize, shift, bias;
  int32_t ndp;            // dp4a windows per stencil row
  int32_t dp_off[3];      // byte offset of each window relative to the output byte
  uint32_t wpk[5][3];     // packed int8 weights, [row][window]
  uint32_t wpk2[3][3];    // Sobel: the gy weights
  int32_t sep[5];         // PH_SEPH / PH_SEPV: the 1-D weights of this pass
  // filled per launch (they depend on the tile height): work decomposition of the phase, so that
  // the kernel does no runtime integer division (item / units == (item * units_magic) >> 20)
  int32_t units, units_magic, n_items, drow, dcol;
};

struct BufDesc {
  int32_t off;    // byte offset in dynamic shared memory
  int32_t pitch;  // bytes per row
  int32_t xorg;   // plane column of byte 0 (0 for planes, kApron for the staging tile)
  int32_t yorg;   // plane row of row 0 (0 for planes, halo for the staging tile)
};

struct TileParams {
  // geometry (global image coordinates)
  int32_t W, H, cin, cout;
  int32_t y0, y1;            // rows this launch produces
  int32_t th;                // tile height
  int32_t tiles_x, tiles_y, n_tiles;
  int32_t halo;              // segment halo R: plane row 0 is tile row -R
  int32_t n_phases;
  int32_t rgb_order;
  int32_t debug;             // bit 0: skip prefetch.tensormap
  int32_t raw_fix_border;    // border rule to paint onto the raw tile (first phase is a stencil), or -1
  // source / destination views for the non-TMA path, and tensor-map row origins
  int32_t use_tma_load, use_tma_store;
  int32_t src_y_begin, src_y_end;  // global rows addressable through src_ptr (row src_y_begin at src_ptr)
  int32_t dst_y_begin, dst_y_end;
  int64_t src_pitch, dst_pitch;
  uint64_t src_ptr, dst_ptr;
  // halo push (multi-GPU resident chains): rows [y0, y0+push_rows) also go to the upper
  // neighbour's lower ghost rows, rows [y1-push_rows, y1) to the lower neighbour's upper ghosts
  int32_t push_rows;
  int32_t push_up_y_begin, push_dn_y_begin;  // global row that lands on row 0 of the push view
  uint64_t push_up_ptr, push_dn_ptr;
  int64_t push_pitch;
  // cross-GPU flags (0 = unused)
  uint64_t wait_up_flag, wait_dn_flag;      // local flags to reach `wait_epoch` before ghost rows are read
  uint64_t signal_up_flag, signal_dn_flag;  // neighbours' flags to set to `signal_epoch` after the pushes
  uint64_t counters;                        // 4 x u32 scratch in local memory, zero before launch
  uint64_t status;                          // u32 in local memory: nonzero = a wait timed out
  uint32_t wait_epoch, signal_epoch;
  uint64_t timeout_ns;
  // shared memory layout
  BufDesc buf[4];
  int32_t raw_stages;        // 2: the next tile is prefetched while this one is filtered; 1: no prefetch, less smem
  int32_t raw_stage_bytes;   // stage s of BUF_RAW lives at buf[0].off + s*raw_stage_bytes
  int32_t raw_rows, raw_tx_bytes;
  int32_t lut_off, bar_off, smem_bytes;
  PhaseDesc ph[kMaxPhases];
  uint8_t luts[kMaxLuts][256];
};

// Pointwise-only chains: [lut0] -> [gray] -> [lut1] -> [expand], streamed with 16-byte accesses.
struct PointwiseParams {
  int32_t W, cin, cout;
  int32_t rows;               // rows to process
  int32_t gray;               // 0 none, 1 ref, 2 cv
  int32_t expand;
  int32_t rgb_order;
  int32_t has_lut0, has_lut1;
  int32_t vec_ok;             // 16-byte path usable (pitches and bases aligned)
  int64_t src_pitch, dst_pitch;
  uint64_t src_ptr, dst_ptr;  // row 0 of the rows to process
  uint8_t lut0[256], lut1[256];
};

}  // namespace b2i
