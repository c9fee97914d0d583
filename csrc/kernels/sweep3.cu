// The sweep kernel: the reference's whole pipeline (kernel.cu:192-195,210 / kern.cpp:73-77) as a
// warp-synchronous column sweep with everything between load and store in registers.
//
//   gray (kernel.cu:39-42 or cv) -> [LUT: contrast] -> 3x3 stencil on gray -> [LUT] -> [gray->BGR]
//
// The generic tile-chain kernel (tile_chain.cu) runs this shape at ~48 % of the HBM roofline: it
// is bound by instruction issue, not by memory (ncu, profiles/). Two thirds of its instructions
// are shared-memory round trips, index arithmetic and block barriers between phases. This kernel
// removes all three for the shape the reference actually runs:
//   * a warp owns a strip of 240 output pixels (30 lanes x 8 pixels; lanes 0 and 31 only feed
//     their neighbours) and walks down a segment of rows;
//   * rows arrive by TMA (3-row boxes, 3 stages per warp, one mbarrier per stage) and each lane
//     converts its 24 bytes to 8 gray pixels exactly once;
//   * horizontal neighbours come from two warp shuffles per row, vertical neighbours from three
//     rotating accumulator sets: a gray row is used for the three output rows it touches the
//     moment it exists and is never stored;
//   * finished rows are expanded to BGR in registers, staged per warp and leave by TMA store.
// No __syncthreads, no inter-phase planes, 15-16 instructions per pixel instead of 37.
// Source and destination may be peer memory (scatter / gather fused), exactly as in tile_chain.cu.
#include "b200img.h"
#include "device_utils.cuh"
#include "launch_pdl.hpp"
#include "sweep_hooks.cuh"
#include "sweep3.hpp"

#include <type_traits>

namespace b2i {
using namespace dev;

namespace {

constexpr int kWarps = kSweepWarps;
constexpr int CH = kSweepChunk;   // rows per TMA box
constexpr int NST = kSweepStages;  // input stages per warp

__device__ __forceinline__ uint32_t lut4s(const uint8_t* lut, uint32_t w) {
  const uint32_t a = lut[w & 0xFF], b = lut[(w >> 8) & 0xFF], c = lut[(w >> 16) & 0xFF], d = lut[w >> 24];
  return pack4_lo(a, b, c, d);
}

// Bytes [o, o+4) of the 16-byte window {left, own0, own1, right}.
template <int O>
__device__ __forceinline__ uint32_t win(const uint32_t (&a)[4]) {
  if constexpr ((O & 3) == 0) return a[O >> 2];
  else return __funnelshift_r(a[O >> 2], a[(O >> 2) + 1], (O & 3) * 8);
}

__device__ __forceinline__ void taps_init(const uint32_t (&v)[8], uint32_t w, int (&acc)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = dp4a_us(v[i], w, 0);
}
// acc[i] += dot(gray[x_i-1 .. x_i+1], w) for the lane's 8 pixels.
__device__ __forceinline__ void taps(const uint32_t (&v)[8], uint32_t w, int (&acc)[8]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = dp4a_us(v[i], w, acc[i]);
}

template <int GRAY>
__device__ __forceinline__ void gray8(const uint8_t* s, const GrayRefCoef& gk, int rgb, uint32_t (&g)[8]) {
  const uint2 v0 = *reinterpret_cast<const uint2*>(s), v1 = *reinterpret_cast<const uint2*>(s + 8),
              v2 = *reinterpret_cast<const uint2*>(s + 16);
  const uint32_t w0 = v0.x, w1 = v0.y, w2 = v1.x, w3 = v1.y, w4 = v2.x, w5 = v2.y;
  if (GRAY == 1) {
    g[0] = gray_ref_f23(byte_to_f23<0>(w0), byte_to_f23<1>(w0), byte_to_f23<2>(w0), gk);
    g[1] = gray_ref_f23(byte_to_f23<3>(w0), byte_to_f23<0>(w1), byte_to_f23<1>(w1), gk);
    g[2] = gray_ref_f23(byte_to_f23<2>(w1), byte_to_f23<3>(w1), byte_to_f23<0>(w2), gk);
    g[3] = gray_ref_f23(byte_to_f23<1>(w2), byte_to_f23<2>(w2), byte_to_f23<3>(w2), gk);
    g[4] = gray_ref_f23(byte_to_f23<0>(w3), byte_to_f23<1>(w3), byte_to_f23<2>(w3), gk);
    g[5] = gray_ref_f23(byte_to_f23<3>(w3), byte_to_f23<0>(w4), byte_to_f23<1>(w4), gk);
    g[6] = gray_ref_f23(byte_to_f23<2>(w4), byte_to_f23<3>(w4), byte_to_f23<0>(w5), gk);
    g[7] = gray_ref_f23(byte_to_f23<1>(w5), byte_to_f23<2>(w5), byte_to_f23<3>(w5), gk);
  } else {
    g[0] = gray_cv_px(w0 & 0xFF, (w0 >> 8) & 0xFF, (w0 >> 16) & 0xFF, rgb);
    g[1] = gray_cv_px(w0 >> 24, w1 & 0xFF, (w1 >> 8) & 0xFF, rgb);
    g[2] = gray_cv_px((w1 >> 16) & 0xFF, w1 >> 24, w2 & 0xFF, rgb);
    g[3] = gray_cv_px((w2 >> 8) & 0xFF, (w2 >> 16) & 0xFF, w2 >> 24, rgb);
    g[4] = gray_cv_px(w3 & 0xFF, (w3 >> 8) & 0xFF, (w3 >> 16) & 0xFF, rgb);
    g[5] = gray_cv_px(w3 >> 24, w4 & 0xFF, (w4 >> 8) & 0xFF, rgb);
    g[6] = gray_cv_px((w4 >> 16) & 0xFF, w4 >> 24, w5 & 0xFF, rgb);
    g[7] = gray_cv_px((w5 >> 8) & 0xFF, (w5 >> 16) & 0xFF, w5 >> 24, rgb);
  }
}

// PLAIN: result rows leave by 8-byte stores from registers (NVLS multicast output) instead of the
// staging buffer + TMA store; a separate instantiation, so the single-GPU kernel pays nothing.
// SPEC: the reference programs' own shapes are compiled in -- 1 = kernel.cu (border=ref), 2 = kern.cpp
// (border=reflect101), both without rescale, without a LUT after the stencil, with the gray -> BGR
// expand: the warp-uniform tests of those launch parameters, ~10 % of the row loop's instructions,
// disappear. 0 = everything else, decided at run time.
template <int GRAY, int LUT0, bool PLAIN, int SPEC>
__global__ void __launch_bounds__(kWarps * 32, kSweepCtasPerSm)
sweep3_kernel(const __grid_constant__ SweepParams P, const __grid_constant__ CUtensorMap src_map,
              const __grid_constant__ CUtensorMap dst_map) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(256) uint8_t lut0[256];
  __shared__ __align__(256) uint8_t lut1[256];
  __shared__ __align__(8) uint64_t bars[kWarps * NST];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();   // the next launch in the stream may set itself up behind this one
  for (int i = threadIdx.x; i < 64; i += kWarps * 32) {
    reinterpret_cast<uint32_t*>(lut0)[i] = reinterpret_cast<const uint32_t*>(P.lut0)[i];
    reinterpret_cast<uint32_t*>(lut1)[i] = reinterpret_cast<const uint32_t*>(P.lut1)[i];
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < kWarps * NST; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
    tma_prefetch_desc(&src_map);
    if (!PLAIN) tma_prefetch_desc(&dst_map);
  }
  __syncthreads();  // the only block-wide barrier of the kernel
  pdl_wait();       // the launch before this one has completed: its output may be this one's input

  const bool expand = SPEC != 0 || P.expand;
  const int out_pitch = expand ? kSweepOutBytes : kSweepStrip;  // staged bytes per output row
  uint8_t* in_ring = smem + warp * P.warp_smem;        // NST stages of CH rows x kSweepInBytes (128 B aligned)
  uint8_t* out_ring = in_ring + NST * P.in_stage_bytes;  // 2 staging buffers of CH rows x out_pitch
  uint64_t* bar = &bars[warp * NST];
  const bool halo_lane = lane == 0 || lane == 31;
  const GrayRefCoef gk = gray_ref_coef(P.rgb_order);
  const int rgb = P.rgb_order;
  const uint32_t w0 = P.wpk[0], w1 = P.wpk[1], w2 = P.wpk[2];
  // gray8<1> returns the float bits 0x4B740000 + gray: folded into the affine table's constant
  const uint32_t aff_a = (uint32_t)P.lut0_a, aff_c = (uint32_t)P.lut0_b - (GRAY == 1 ? aff_a * 0x4B740000u : 0u);
  const int aff_s = P.lut0_s;
  const bool ref = SPEC == 1 || (SPEC == 0 && P.border == B2I_BORDER_REF);
  const bool reflect = SPEC == 2 || (SPEC == 0 && P.border == B2I_BORDER_REFLECT101);
  const bool scale = SPEC == 0 && (P.shift | P.bias) != 0;
  const bool lut1_on = SPEC == 0 && P.has_lut1;

  uint32_t n_in = 0;   // input chunks consumed by this warp so far (stage = n_in % NST)
  uint32_t n_out = 0;  // output chunks stored so far (buffer = n_out & 1)
  const int total_warps = gridDim.x * kWarps;
  // Fewer tasks than warps (small bands): warp w of CTA c is global warp w * gridDim.x + c, so every SM gets
  // its share of the busy warps instead of some SMs a full CTA and others none. Otherwise CTA by CTA.
  const int global_warp = P.n_tasks < total_warps ? warp * (int)gridDim.x + (int)blockIdx.x : (int)blockIdx.x * kWarps + warp;
  for (int task = global_warp; task < P.n_tasks; task += total_warps) {
    const int seg = task / P.n_strips, strip = task - seg * P.n_strips;
    const int ya = P.y0 + seg * P.seg_rows, yb = min(P.y1, ya + P.seg_rows);  // output rows [ya, yb)
    const int steps = yb - ya + 2;                   // gray rows ya-1 .. yb
    const int n_chunks = (steps + CH - 1) / CH;
    const int xs = strip * kSweepStrip;              // first output pixel of the strip
    const int x0 = xs - 8 + lane * 8;                // first pixel this lane holds (lane 0: halo)
    const int xw_in = (xs * 3 - 32) >> 2;            // TMA x (u32 elements): 32 bytes left of the strip
    const int xw_out = (xs * (expand ? 3 : 1)) >> 2;

    // ---- input pipeline: the first NST chunks are in flight before any is consumed --------
    if (lane == 0) {
      for (int c = 0; c < min(NST, n_chunks); ++c) {
        const uint32_t st = (n_in + c) % NST;
        mbar_arrive_expect_tx(&bar[st], CH * kSweepInBytes);
        tma_load_2d(in_ring + st * P.in_stage_bytes, &src_map, xw_in, ya - 1 + c * CH - P.src_y_begin, &bar[st]);
      }
    }

    // REF border: which of my 8 pixels are filtered at all (x in [2, W-2]); rows checked below
    uint32_t m0 = 0xFFFFFFFFu, m1 = 0xFFFFFFFFu;
    if (ref) {
      const int lo = 2 - x0, hi = P.W - 2 - x0;
      if (lo > 0 || hi < 7) {
        m0 = m1 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if (i >= lo && i <= hi) m0 |= 0xFFu << (8 * i);
          if (i + 4 >= lo && i + 4 <= hi) m1 |= 0xFFu << (8 * i);
        }
      }
    }
    const bool left_edge = x0 == 0, right_edge = x0 + 8 == P.W, beyond = x0 < 0 || x0 >= P.W;
    const bool folds = SPEC == 2 || (SPEC == 0 && (reflect || P.border == B2I_BORDER_REPLICATE));
    const bool edge_fix = folds && (xs == 0 || xs + kSweepStrip >= P.W);         // warp-uniform
    const bool vert_fix = folds && (ya <= 1 || yb >= P.H - 1);                   // warp-uniform

    // Three accumulator sets, X[j] collecting the output row whose index (relative to ya-1) is
    // j mod 3. The row loop is unrolled by the box height so that the roles (bottom / middle /
    // top taps), the input-stage bookkeeping and the store schedule are all static per position.
    int X0[8], X1[8], X2[8];
    uint32_t prev[4] = {0, 0, 0, 0};  // window words of the previous gray row
#pragma unroll
    for (int i = 0; i < 8; ++i) X0[i] = X1[i] = X2[i] = 0;

    // one gray row: convert, exchange edge pixels, apply to the three output rows it touches,
    // emit the output row that is now complete
    // MID (compile-time): a box from the middle of the segment (not one of the first two, not one of
    // the last two): every row is inside the image and inside the filtered rows, an output row is
    // always due, and a store box ends exactly at its third row. The tests for all of that are
    // warp-uniform but cost ~8 % of the row loop's instructions; the middle boxes run without them.
    auto step = [&](auto qc, auto midc, int c, const uint8_t* stage_base, uint32_t ob_prev, uint32_t ob_cur,
                    int (&A)[8], int (&B)[8], int (&Cc)[8]) {
      constexpr int q = decltype(qc)::value;
      constexpr bool MID = decltype(midc)::value;
      const int u = c * CH + q;
      const int y = ya - 1 + u;
      uint32_t a[4];
      {
        // branch-free: rows outside the image arrive zero-filled from TMA and are converted like
        // any other, then replaced by 0 (they contribute nothing; borders are folded in below)
        uint32_t g[8];
        gray8<GRAY>(stage_base + q * kSweepInBytes + 8 + lane * 24, gk, rgb, g);
        const bool valid = (MID || (unsigned)y < (unsigned)P.H) && !beyond;
        if constexpr (LUT0 == 2) {
          // the table is clamp((a v + b) >> s) (a contrast stretch, kernel.cu:56 / kern.cpp:74): one
          // multiply-add and a shift per pixel, clamped by the saturating pack; no shared-memory
          // lookups in the row loop. The constant absorbs the float bits gray8 leaves above the value.
          int t[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) t[i] = (int)(g[i] * aff_a + aff_c) >> aff_s;
          a[1] = valid ? sat_pack4(t[0], t[1], t[2], t[3]) : 0u;
          a[2] = valid ? sat_pack4(t[4], t[5], t[6], t[7]) : 0u;
        } else {
          if constexpr (LUT0 == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) g[i] = lut0[g[i] & 0xFF];
          }
          a[1] = valid ? pack4_lo(g[0], g[1], g[2], g[3]) : 0u;
          a[2] = valid ? pack4_lo(g[4], g[5], g[6], g[7]) : 0u;
        }
      }
      a[0] = __shfl_up_sync(0xFFFFFFFFu, a[2], 1);
      a[3] = __shfl_down_sync(0xFFFFFFFFu, a[1], 1);
      if (edge_fix) {
        if (reflect) {  // pixel -1 := pixel 1, pixel W := pixel W-2
          if (left_edge) a[0] = __byte_perm(a[0], a[1], 0x5210);
          if (right_edge) a[3] = __byte_perm(a[3], a[2], 0x3216);
        } else {                                  // replicate: pixel -1 := pixel 0, pixel W := pixel W-1
          if (left_edge) a[0] = __byte_perm(a[0], a[1], 0x4210);
          if (right_edge) a[3] = __byte_perm(a[3], a[2], 0x3217);
        }
      }
      uint32_t v[8];
      v[0] = win<3>(a); v[1] = win<4>(a); v[2] = win<5>(a); v[3] = win<6>(a);
      v[4] = win<7>(a); v[5] = win<8>(a); v[6] = win<9>(a); v[7] = win<10>(a);
      taps(v, w2, A);   // bottom taps of output row y-1
      taps(v, w1, B);   // middle taps of output row y
      taps_init(v, w0, Cc);  // top taps of output row y+1: this set starts here
      if (!MID && vert_fix) {   // row -1 := row 1 / row 0, row H := row H-2 / row H-1 (first / last segment only)
        if (reflect) {
          if (y == 1) taps(v, w0, A);
          if (y == P.H - 1) {
            uint32_t pv[8];
            pv[0] = win<3>(prev); pv[1] = win<4>(prev); pv[2] = win<5>(prev); pv[3] = win<6>(prev);
            pv[4] = win<7>(prev); pv[5] = win<8>(prev); pv[6] = win<9>(prev); pv[7] = win<10>(prev);
            taps(pv, w2, B);
          }
        } else {
          if (y == 0) taps(v, w0, B);
          if (y == P.H - 1) taps(v, w2, B);
        }
      }
      // ---- output row y-1 (index e = u-2 of the task) is complete ----------------------------
      if (MID || u >= 2) {
        constexpr int er = (q + 1) % 3;             // row of the store box
        const uint32_t ob = er == 0 ? ob_cur : ob_prev;
        if (er == 0) {                               // first row of a box: its previous store has read it
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
        }
        if (scale) {
#pragma unroll
          for (int i = 0; i < 8; ++i) A[i] = (A[i] + P.bias) >> P.shift;
        }
        uint32_t r0 = sat_pack4(A[0], A[1], A[2], A[3]);
        uint32_t r1 = sat_pack4(A[4], A[5], A[6], A[7]);
        if (ref) {
          const int oy = y - 1;
          const bool row_ok = MID || (oy >= 2 && oy <= P.H - 2);
          const uint32_t mm0 = row_ok ? m0 : 0u, mm1 = row_ok ? m1 : 0u;
          r0 = (r0 & mm0) | (prev[1] & ~mm0);  // prev = gray row y-1: the pass-through value
          r1 = (r1 & mm1) | (prev[2] & ~mm1);
        }
        if (lut1_on) {
          r0 = lut4s(lut1, r0);
          r1 = lut4s(lut1, r1);
        }
        if (!halo_lane) {
          // predicated stores: the halo lanes (0, 31) stage nothing. (They used to store into a spare
          // slot; that slot sat on the banks of two other lanes and made every staging store a 3- or
          // 4-way bank conflict: ncu, profiles/ncu/prof_r2_sweep3_kernel_cu.json.)
          uint8_t* d = out_ring + ob * P.out_buf_bytes + er * out_pitch + (lane - 1) * (expand ? 24 : 8);
          if (expand) {  // kernel.cu:210 in registers: 8 gray -> 24 bytes
            uint2* qd = reinterpret_cast<uint2*>(d);
            qd[0] = make_uint2(__byte_perm(r0, 0, 0x1000), __byte_perm(r0, 0, 0x2211));
            qd[1] = make_uint2(__byte_perm(r0, 0, 0x3332), __byte_perm(r1, 0, 0x1000));
            qd[2] = make_uint2(__byte_perm(r1, 0, 0x2211), __byte_perm(r1, 0, 0x3332));
          } else {
            *reinterpret_cast<uint2*>(d) = make_uint2(r0, r1);
          }
        }
        if (PLAIN && (er == 2 || (!MID && u == steps - 1))) {
          // NVLS: the box leaves in coalesced 16-byte stores to the multicast image, one store per
          // byte, and the switch writes every GPU's copy
          const int ec = (u - 2) / CH, px = expand ? 3 : 1;
          mc_store_box(out_ring + ob * P.out_buf_bytes, er + 1, out_pitch, min(out_pitch, (P.W - xs) * px),
                       P.plain_dst_ptr + (int64_t)(ya + ec * CH - P.y0) * P.plain_dst_pitch + (int64_t)xs * px,
                       P.plain_dst_pitch, lane);
        }
        if (!PLAIN && (er == 2 || (!MID && u == steps - 1))) {  // a box of finished rows leaves by TMA
          const int ec = (u - 2) / CH;
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&dst_map, xw_out, ya + ec * CH - P.y0, out_ring + ob * P.out_buf_bytes);
            tma_store_commit();
          }
        }
      }
      prev[0] = a[0]; prev[1] = a[1]; prev[2] = a[2]; prev[3] = a[3];
    };

    for (int c = 0; c < n_chunks; ++c) {
      const uint32_t st = (n_in + c) % NST;
      const uint8_t* stage_base = in_ring + st * P.in_stage_bytes;
      // output box c-1 is finished during rows 0,1 of input box c; box c starts at row 2
      const uint32_t ob_prev = (n_out + c + 1) & 1, ob_cur = (n_out + c) & 1;
      const int rows_here = min(CH, steps - c * CH);
      mbar_wait(&bar[st], ((n_in + c) / NST) & 1);
      if (SPEC != 1 && c >= 2 && c + 3 <= n_chunks) {   // measured: kernel.cu's own shape runs 5 % slower with it
        step(std::integral_constant<int, 0>{}, std::true_type{}, c, stage_base, ob_prev, ob_cur, X2, X0, X1);
        step(std::integral_constant<int, 1>{}, std::true_type{}, c, stage_base, ob_prev, ob_cur, X0, X1, X2);
        step(std::integral_constant<int, 2>{}, std::true_type{}, c, stage_base, ob_prev, ob_cur, X1, X2, X0);
      } else {
        step(std::integral_constant<int, 0>{}, std::false_type{}, c, stage_base, ob_prev, ob_cur, X2, X0, X1);
        if (rows_here > 1) step(std::integral_constant<int, 1>{}, std::false_type{}, c, stage_base, ob_prev, ob_cur, X0, X1, X2);
        if (rows_here > 2) step(std::integral_constant<int, 2>{}, std::false_type{}, c, stage_base, ob_prev, ob_cur, X1, X2, X0);
      }
      // every lane has read the box: the stage can take box c + NST
      __syncwarp();
      if (lane == 0 && c + NST < n_chunks) {
        mbar_arrive_expect_tx(&bar[st], CH * kSweepInBytes);
        tma_load_2d(in_ring + st * P.in_stage_bytes, &src_map, xw_in, ya - 1 + (c + NST) * CH - P.src_y_begin,
                    &bar[st]);
      }
    }
    n_in += n_chunks;
    n_out += (yb - ya + CH - 1) / CH;
  }
  sweep_kernel_end(P, global_warp, total_warps, lane);
}

}  // namespace

cudaError_t launch_sweep3(const SweepParams& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                          int grid, cudaStream_t stream) {
  static bool configured[64][24] = {{false}};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const int smem = P.warp_smem * kSweepWarps;
  const bool shape = !P.plain_dst_ptr && !(P.shift | P.bias) && !P.has_lut1 && P.expand;
  const int spec = !shape ? 0 : P.border == B2I_BORDER_REF ? 1 : P.border == B2I_BORDER_REFLECT101 ? 2 : 0;
  const int mode = P.plain_dst_ptr ? 1 : spec ? 1 + spec : 0;  // 0 generic, 1 NVLS stores, 2 kernel.cu shape, 3 kern.cpp shape
  const int variant = mode * 6 + (P.gray == 1 ? 0 : 3) + P.has_lut0;  // has_lut0: 0 none, 1 table, 2 affine
  using Kern = void (*)(SweepParams, CUtensorMap, CUtensorMap);
#define B2I_S3_ROW(PLAIN, SPEC)                                                                       \
  sweep3_kernel<1, 0, PLAIN, SPEC>, sweep3_kernel<1, 1, PLAIN, SPEC>, sweep3_kernel<1, 2, PLAIN, SPEC>, \
      sweep3_kernel<2, 0, PLAIN, SPEC>, sweep3_kernel<2, 1, PLAIN, SPEC>, sweep3_kernel<2, 2, PLAIN, SPEC>
  static const Kern table[24] = {B2I_S3_ROW(false, 0), B2I_S3_ROW(true, 0), B2I_S3_ROW(false, 1), B2I_S3_ROW(false, 2)};
#undef B2I_S3_ROW
  const Kern kern = table[variant];
  if (dev < 64 && !configured[dev][variant]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
    configured[dev][variant] = true;
  }
  return launch_pdl(true, kern, grid, kSweepWarps * 32, smem, stream, P, src_map, dst_map);
}

}  // namespace b2i
