// The Sobel sweep kernel: out = clamp(|gx| + |gy|) per channel on interleaved 3-channel data (the
// second half of BASELINE.json config 4, "blur -> Sobel"), as a warp-synchronous column sweep with
// the skeleton of sweep5.cu: a warp owns a strip of 240 pixels and walks down a segment of rows
// that arrive by TMA, results leave by TMA store, nothing but the warp's own rows is shared.
//
// Both Sobel kernels are separable, gx = [1 2 1]^T x [-1 0 1] and gy = [-1 0 1]^T x [1 2 1], so a
// row is first reduced horizontally, in registers, to a difference D and a sum S per byte (taps at
// +-3 bytes come from the lane's own 24 bytes and two warp shuffles). Bytes are widened to the two
// 16-bit lanes of a word (even / odd bytes). D is signed: the lanes are kept as one 32-bit integer
// hi * 65536 + lo with a signed lo, which 32-bit adds and multiplies keep exact; the bias 1024 added
// with the last term makes both lanes non-negative again before they are looked at.
//   gx(y-1) = D(y-2) + 2 D(y-1) + D(y)   -> one running partial sum and the previous D
//   gy(y-1) = S(y) - S(y-2)              -> the two previous S rows, swapping roles every row
// The mirrored / replicated row above the image is read from the box that holds its source row; the
// one below is folded into the state once per strip. The row loop itself carries no border rules.
#include "b200img.h"
#include "device_utils.cuh"
#include "launch_pdl.hpp"
#include "sweep_hooks.cuh"
#include "sweep_sobel.hpp"

#include <type_traits>

namespace b2i {
using namespace dev;

namespace {

constexpr int CH = kSSChunk;
constexpr int NST = kSSStages;
constexpr uint32_t kBias = 0x04000400u;  // 1024 per lane: |gx|, |gy| <= 1020

__device__ __forceinline__ uint32_t lut4s(const uint8_t* lut, uint32_t w) {
  const uint32_t a = lut[w & 0xFF], b = lut[(w >> 8) & 0xFF], c = lut[(w >> 16) & 0xFF], d = lut[w >> 24];
  return pack4_lo(a, b, c, d);
}

// The lane pair (x[J], x[J+2]) of the 32-byte window x = {left5, own0..own5, right0}: the two bytes a
// tap at byte offset J contributes to the even (or, at J+1, odd) lanes of a word. Written per residue
// of J so that every pair is extracted once per row and shared by the taps that use it.
template <int J>
__device__ __forceinline__ uint32_t pair_at(const uint32_t (&a)[8]) {
  static_assert(J >= 0 && J + 2 < 32, "pair outside the row registers");
  if constexpr ((J & 3) == 0) return __byte_perm(a[J >> 2], 0u, 0x4240);
  else if constexpr ((J & 3) == 1) return __byte_perm(a[J >> 2], 0u, 0x4341);
  else {
    const uint32_t h = __funnelshift_r(a[J >> 2], a[(J >> 2) + 1], 16);  // bytes 4k+2 .. 4k+5
    return __byte_perm(h, 0u, (J & 3) == 2 ? 0x4240 : 0x4341);
  }
}

// |v - 1024| + 1024 per lane, for lanes in [4, 2044]
__device__ __forceinline__ uint32_t abs_biased(uint32_t v) { return __vmaxu2(v, 0x08000800u - v); }

// MG: the multi-GPU hooks (neighbour flags, halo pushes, NVLS multicast output) are compiled in.
// The single-GPU instantiation carries none of them.
template <bool MG>
__global__ void __launch_bounds__(kSSWarps * 32, kSSCtasPerSm)
sweep_sobel_kernel(const __grid_constant__ SweepSobelParams P, const __grid_constant__ CUtensorMap src_map,
                   const __grid_constant__ CUtensorMap dst_map) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(256) uint8_t lut1[256];
  __shared__ __align__(8) uint64_t bars[kSSWarps * NST];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();   // the next launch in the stream may set itself up behind this one
  for (int i = threadIdx.x; i < 64; i += kSSWarps * 32)
    reinterpret_cast<uint32_t*>(lut1)[i] = reinterpret_cast<const uint32_t*>(P.lut1)[i];
  if (threadIdx.x == 0) {
    for (int i = 0; i < kSSWarps * NST; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
    tma_prefetch_desc(&src_map);
    if (!(MG && P.plain_dst_ptr)) tma_prefetch_desc(&dst_map);
  }
  __syncthreads();  // the only block-wide barrier of the kernel
  pdl_wait();       // the launch before this one has completed: its output may be this one's input

  uint8_t* in_ring = smem + warp * P.warp_smem;          // NST stages of CH rows x kSSInBytes
  uint8_t* out_ring = in_ring + NST * P.in_stage_bytes;  // 2 staging buffers of CH rows x kSSOutBytes
  uint64_t* bar = &bars[warp * NST];
  const bool halo_lane = lane == 0 || lane == 31;
  const bool reflect = P.border == B2I_BORDER_REFLECT101, replicate = P.border == B2I_BORDER_REPLICATE;

  uint32_t n_in = 0, n_out = 0;  // boxes consumed / stored by this warp so far
  const int total_warps = gridDim.x * kSSWarps;
  // Fewer tasks than warps (small bands): warp w of CTA c is global warp w * gridDim.x + c, so every SM gets
  // its share of the busy warps instead of some SMs a full CTA and others none. Otherwise CTA by CTA.
  const int global_warp = !MG && P.n_tasks < total_warps ? warp * (int)gridDim.x + (int)blockIdx.x : (int)blockIdx.x * kSSWarps + warp;
  for (int task = global_warp; task < P.n_tasks; task += total_warps) {
    // boundary segments first: their rows feed the neighbours' next pass
    const int so = task / P.n_strips, strip = task - so * P.n_strips;
    const int seg = so == 0 ? 0 : so == 1 ? P.n_segs - 1 : so - 1;
    int ya, yb;                                       // output rows [ya, yb)
    if (MG && P.bseg > 0) {   // short first / last segments: their halo rows are on the way early in the pass
      ya = seg == 0 ? P.y0 : seg == P.n_segs - 1 ? P.y1 - P.bseg_last : P.y0 + P.bseg + (seg - 1) * P.seg_rows;
      yb = seg == 0 ? P.y0 + P.bseg : seg == P.n_segs - 1 ? P.y1 : min(P.y1 - P.bseg_last, ya + P.seg_rows);
    } else {
      ya = P.y0 + seg * P.seg_rows;
      yb = min(P.y1, ya + P.seg_rows);
    }
    const int steps = yb - ya + 2;                   // input rows ya-1 .. yb
    const int n_chunks = (steps + CH - 1) / CH;
    const int xs = strip * kSSStrip;
    const int x0 = xs - 8 + lane * 8;                // first pixel this lane holds
    const int xw_in = (xs * 3 - 32) >> 2, xw_out = (xs * 3) >> 2;
    const bool sig_up = MG && P.signal_up_flag && ya < P.y0 + P.push_rows;
    const bool sig_dn = MG && P.signal_dn_flag && yb > P.y1 - P.push_rows;
    if (lane == 0) {
      // a task that reads ghost rows, or writes the neighbour's, first sees the neighbour's flag
      if (MG) sweep_wait_neighbours(P, ya, yb, 1, sig_up, sig_dn);
      for (int c = 0; c < min(NST, n_chunks); ++c) {
        const uint32_t st = (n_in + c) % NST;
        mbar_arrive_expect_tx(&bar[st], CH * kSSInBytes);
        tma_load_2d(in_ring + st * P.in_stage_bytes, &src_map, xw_in, ya - 1 + c * CH - P.src_y_begin, &bar[st]);
      }
    }
    const bool left_edge = x0 == 0, right_edge = x0 + 8 == P.W, beyond = x0 < 0 || x0 >= P.W;
    const bool edge_fix = (reflect || replicate) && (xs == 0 || xs + kSSStrip >= P.W);  // warp-uniform
    const bool vert_fix = (reflect || replicate) && (ya <= 1 || yb >= P.H - 1);        // warp-uniform

    // per (word, even/odd) lane pair: running gx partial, previous D, the two previous S rows
    uint32_t PA[12], DP[12], SA[12], SB[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) PA[i] = DP[i] = SA[i] = SB[i] = 0;

    // one input row y; emits output row y-1. S1 holds S(y-1), S2 holds S(y-2) and takes S(y).
    // SLOW (compile-time): the task also pushes halo rows or stores to a multicast image; every
    // other task of a multi-GPU launch runs the single-GPU row loop (as in sweep5.cu).
    auto step = [&](auto qc, auto slowc, int c, const uint8_t* stage_base, uint32_t ob_prev, uint32_t ob_cur,
                    uint32_t (&S1)[12], uint32_t (&S2)[12]) {
      constexpr int q = decltype(qc)::value;
      constexpr bool SLOW = MG && decltype(slowc)::value;
      const int u = c * CH + q;
      const int y = ya - 1 + u;
      uint32_t a[8];
      {
        // The row above the image is row 1 (reflect101) or row 0 (replicate): both sit in the same box.
        // Rows outside the image otherwise (zero-filled by TMA) and lanes beyond the right edge give 0.
        const bool above = vert_fix && y < 0;
        const uint8_t* s = stage_base + (above ? (reflect ? 2 : 1) : q) * kSSInBytes + 8 + lane * 24;
        const uint2 v0 = *reinterpret_cast<const uint2*>(s), v1 = *reinterpret_cast<const uint2*>(s + 8),
                    v2 = *reinterpret_cast<const uint2*>(s + 16);
        const bool valid = ((unsigned)y < (unsigned)P.H || above) && !beyond;
        a[1] = valid ? v0.x : 0u; a[2] = valid ? v0.y : 0u; a[3] = valid ? v1.x : 0u;
        a[4] = valid ? v1.y : 0u; a[5] = valid ? v2.x : 0u; a[6] = valid ? v2.y : 0u;
      }
      a[0] = __shfl_up_sync(0xFFFFFFFFu, a[6], 1);
      a[7] = __shfl_down_sync(0xFFFFFFFFu, a[1], 1);
      if (edge_fix) {
        if (reflect) {  // pixel -1 := pixel 1; pixel W := pixel W-2
          if (left_edge) a[0] = __byte_perm(a[1], a[2], 0x5433);
          if (right_edge) a[7] = __byte_perm(a[5], a[6], 0x4432);
        } else {        // replicate: pixel -1 := pixel 0; pixel W := pixel W-1
          if (left_edge) a[0] = __byte_perm(a[1], 0u, 0x2100);
          if (right_edge) a[7] = __byte_perm(a[6], 0u, 0x3321);
        }
      }
      // The row below the image never arrives (this step sees zeros, d = s = 0): its mirror image is
      // folded into the state instead, once per strip, so that the row loop carries no border rules.
      //   reflect101: D(H) := D(H-2), S(H) := S(H-2)  ->  gx = 2 PA - 2 D(H-1) - bias, gy = 0
      //   replicate:  D(H) := D(H-1), S(H) := S(H-1)  ->  gx = PA + D(H-1),           gy = S(H-1) - S(H-2)
      if (vert_fix && y == P.H) {
#pragma unroll
        for (int i = 0; i < 12; ++i) {
          if (reflect) {
            PA[i] = 2 * PA[i] - 2 * DP[i] - kBias;
            S2[i] = 0;
          } else {
            PA[i] = PA[i] + DP[i];
            S2[i] = S2[i] - S1[i];
          }
        }
      }
      constexpr int er = (q + 2) % CH;  // row of the store box
      const uint32_t ob = er < 2 ? ob_cur : ob_prev;
      if (u >= 2 && u < steps && er == 0) {  // first row of a box: its previous store has read the buffer
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        __syncwarp();
      }
      uint32_t r[6];
      auto word = [&](auto mc) {
        constexpr int M = decltype(mc)::value;
        constexpr int B = 4 + 4 * M;
        const uint32_t pl[2] = {pair_at<B - 3>(a), pair_at<B - 2>(a)}, pc[2] = {pair_at<B>(a), pair_at<B + 1>(a)},
                       pr[2] = {pair_at<B + 3>(a), pair_at<B + 4>(a)};
        uint32_t res[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int i = 2 * M + h;
          const uint32_t l = pl[h], cc = pc[h], rr = pr[h];
          const uint32_t d = rr - l, s = l + rr + 2 * cc;
          const uint32_t gx = PA[i] + d;
          const uint32_t gy = s - S2[i] + kBias;
          PA[i] = DP[i] + kBias + 2 * d;
          DP[i] = d;
          S2[i] = s;
          res[h] = __vminu2(abs_biased(gx) + abs_biased(gy), 0x08FF08FFu);  // low byte = clamp(|gx|+|gy|)
        }
        r[M] = __byte_perm(res[0], res[1], 0x6240);
      };
      word(std::integral_constant<int, 0>{});
      word(std::integral_constant<int, 1>{});
      word(std::integral_constant<int, 2>{});
      word(std::integral_constant<int, 3>{});
      word(std::integral_constant<int, 4>{});
      word(std::integral_constant<int, 5>{});
      // ---- output row y-1 (index e = u-2 of the task) ---------------------------------------------
      if (u >= 2 && u < steps) {
        if (P.has_lut1) {
#pragma unroll
          for (int m = 0; m < 6; ++m) r[m] = lut4s(lut1, r[m]);
        }
        if (!halo_lane) {   // predicated: the halo lanes stage nothing (a spare slot for them cost bank conflicts)
          uint2* d = reinterpret_cast<uint2*>(out_ring + ob * P.out_buf_bytes + er * kSSOutBytes + (lane - 1) * 24);
          d[0] = make_uint2(r[0], r[1]);
          d[1] = make_uint2(r[2], r[3]);
          d[2] = make_uint2(r[4], r[5]);
        }
        if (SLOW && P.push_rows && !halo_lane && !beyond)  // halo rows also go into the neighbours' ghost rows
          sweep_push_row(P, y - 1, (int64_t)xs * 3 + (lane - 1) * 24, r);
        if (SLOW && P.plain_dst_ptr && (er == CH - 1 || u == steps - 1)) {
          // NVLS: the box leaves in coalesced 16-byte stores to the multicast image, one store per
          // byte, and the switch writes every GPU's copy
          const int ec = (u - 2) / CH;
          mc_store_box(out_ring + ob * P.out_buf_bytes, er + 1, kSSOutBytes, min(kSSOutBytes, (P.W - xs) * 3),
                       P.plain_dst_ptr + (int64_t)(ya + ec * CH - P.y0) * P.plain_dst_pitch + (int64_t)xs * 3,
                       P.plain_dst_pitch, lane);
        }
        if (!(SLOW && P.plain_dst_ptr) && (er == CH - 1 || u == steps - 1)) {  // a box of finished rows leaves by TMA
          const int ec = (u - 2) / CH;
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&dst_map, xw_out, ya + ec * CH - P.y0, out_ring + ob * P.out_buf_bytes);
            tma_store_commit();
          }
        }
      }
    };

    auto run_chunks = [&](auto slowc) {
    for (int c = 0; c < n_chunks; ++c) {
      const uint32_t st = (n_in + c) % NST;
      const uint8_t* stage_base = in_ring + st * P.in_stage_bytes;
      // output box c-1 is finished during rows 0..1 of input box c; box c starts at row 2
      const uint32_t ob_prev = (n_out + c + 1) & 1, ob_cur = (n_out + c) & 1;
      mbar_wait(&bar[st], ((n_in + c) / NST) & 1);
      // All rows of the box run, also those of the last box that lie beyond the segment: they are
      // computed from whatever the box holds and never emitted. Skipping them costs a shuffle of
      // the 36 state registers at every join (measured here: 68.5 -> 65.4 us; the same change made
      // sweep3.cu and sweep5.cu slower, so those keep their conditional rows).
      step(std::integral_constant<int, 0>{}, slowc, c, stage_base, ob_prev, ob_cur, SA, SB);
      step(std::integral_constant<int, 1>{}, slowc, c, stage_base, ob_prev, ob_cur, SB, SA);
      step(std::integral_constant<int, 2>{}, slowc, c, stage_base, ob_prev, ob_cur, SA, SB);
      step(std::integral_constant<int, 3>{}, slowc, c, stage_base, ob_prev, ob_cur, SB, SA);
      __syncwarp();  // every lane has read the box: the stage can take box c + NST
      if (lane == 0 && c + NST < n_chunks) {
        mbar_arrive_expect_tx(&bar[st], CH * kSSInBytes);
        tma_load_2d(in_ring + st * P.in_stage_bytes, &src_map, xw_in, ya - 1 + (c + NST) * CH - P.src_y_begin,
                    &bar[st]);
      }
    }
    };
    if (MG && ((P.push_rows && (ya < P.y0 + P.push_rows || yb > P.y1 - P.push_rows)) || P.plain_dst_ptr))
      run_chunks(std::true_type{});
    else
      run_chunks(std::false_type{});
    n_in += n_chunks;
    n_out += (yb - ya + CH - 1) / CH;
    if (sig_up || sig_dn) sweep_signal_neighbours(P, sig_up, sig_dn, lane);
  }
  sweep_kernel_end(P, global_warp, total_warps, lane);
}

}  // namespace

cudaError_t launch_sweep_sobel(const SweepSobelParams& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                               int grid, cudaStream_t stream) {
  static bool configured[64][2] = {{false}};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const bool mg = P.push_rows || P.wait_up_flag || P.wait_dn_flag || P.signal_up_flag || P.signal_dn_flag ||
                  P.plain_dst_ptr;
  auto kern = mg ? sweep_sobel_kernel<true> : sweep_sobel_kernel<false>;
  if (dev < 64 && !configured[dev][mg]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return e;
    configured[dev][mg] = true;
  }
  // (small bands with multi-GPU hooks: plain launch, see launch_sweep5)
  return launch_pdl(!mg || P.y1 - P.y0 >= pdl_min_rows_mg(), kern, grid, kSSWarps * 32, (size_t)P.warp_smem * kSSWarps, stream, P, src_map,
                    dst_map);
}
}  // namespace b2i
