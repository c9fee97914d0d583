// The separable sweep kernel: a 5x5 (or 3x3) separable stencil on interleaved 3-channel data — the
// 5x5 Gaussian of BASELINE.json configs 3-5 — as a warp-synchronous column sweep, the way sweep3.cu
// runs the reference pipeline.
//
// The tile-chain kernel runs the same filter as two shared-memory phases at 0.24 of the HBM
// roofline: half of its time is per-tile overhead (barriers, phase prologues) and the horizontal
// pass is recomputed for the tile's vertical apron. Here:
//   * a warp owns a strip of 240 pixels (30 lanes x 8 pixels x 3 bytes; lanes 0 and 31 only feed
//     their neighbours) and walks down a segment of rows that arrive by TMA (5-row boxes);
//   * the horizontal pass runs once per row in registers: taps at +-3 and +-6 bytes come from the
//     lane's own 24 bytes and four warp shuffles, bytes are widened to 16-bit lanes (even / odd
//     bytes of a word) and summed with packed integer multiply-adds;
//   * the vertical pass keeps five rotating accumulator sets in registers; a horizontal row is
//     added to the five output rows it touches and never stored. The row loop is unrolled by five
//     so the roles of the sets are static;
//   * image borders cost nothing per pixel: a mirrored or replicated row is the same data as a
//     real row, so it only changes the five (warp-uniform) multipliers of that step;
//   * finished rows are rounded, narrowed, staged per warp and leave by TMA store.
// Exact: every lane stays below 65536 (normalised non-negative kernels, checked by the planner).
#include "b200img.h"
#include "device_utils.cuh"
#include "launch_pdl.hpp"
#include "sweep_hooks.cuh"
#include "sweep5.hpp"

#include <type_traits>

namespace b2i {
using namespace dev;

namespace {

constexpr int CH = kS5Chunk;
constexpr int NST = kS5Stages;

__device__ __forceinline__ uint32_t lut4q(const uint8_t* lut, uint32_t w) {
  const uint32_t a = lut[w & 0xFF], b = lut[(w >> 8) & 0xFF], c = lut[(w >> 16) & 0xFF], d = lut[w >> 24];
  return pack4_lo(a, b, c, d);
}

// The lane pair (x[J], x[J+2]) of the 40-byte window x = {left4, left5, own0..own5, right0, right1}:
// the two bytes a horizontal tap at byte offset J contributes to the even (or, at J+1, odd) lanes
// of a word. Written per residue of J so that every pair is extracted once per row: taps 3 bytes
// apart at even and odd lanes of neighbouring words keep asking for the same pairs, and the
// compiler merges identical expressions (34 extractions per row instead of 60, 9 funnel shifts
// instead of 24).
template <int J>
__device__ __forceinline__ uint32_t pair_at(const uint32_t (&a)[10]) {
  static_assert(J >= 0 && J + 2 < 40, "pair outside the row registers");
  if constexpr ((J & 3) == 0) return __byte_perm(a[J >> 2], 0u, 0x4240);
  else if constexpr ((J & 3) == 1) return __byte_perm(a[J >> 2], 0u, 0x4341);
  else {
    const uint32_t h = __funnelshift_r(a[J >> 2], a[(J >> 2) + 1], 16);  // bytes 4k+2 .. 4k+5
    return __byte_perm(h, 0u, (J & 3) == 2 ? 0x4240 : 0x4341);
  }
}

// Horizontal taps for own word M (bytes 4M..4M+3 of the lane): even-lane and odd-lane sums.
// FIVE: 5 taps (else the 3 in the middle); BINOM: the taps are 1 4 6 4 1 resp. 1 2 1 (the Gaussians),
// which turns most multiplies into adds.
template <int M, bool FIVE, bool BINOM>
__device__ __forceinline__ void hword(const uint32_t (&a)[10], const int (&w)[5], uint32_t& e, uint32_t& o) {
  constexpr int B = 8 + 4 * M;
  const uint32_t e1 = pair_at<B - 3>(a), e2 = pair_at<B>(a), e3 = pair_at<B + 3>(a);
  const uint32_t o1 = pair_at<B - 2>(a), o2 = pair_at<B + 1>(a), o3 = pair_at<B + 4>(a);
  if constexpr (FIVE) {
    const uint32_t e0 = pair_at<B - 6>(a), e4 = pair_at<B + 6>(a);
    const uint32_t o0 = pair_at<B - 5>(a), o4 = pair_at<B + 7>(a);
    if constexpr (BINOM) {
      e = e2 * 6u + (e0 + e4) + ((e1 + e3) << 2);
      o = o2 * 6u + (o0 + o4) + ((o1 + o3) << 2);
    } else {
      e = e0 * w[0] + e1 * w[1] + e2 * w[2] + e3 * w[3] + e4 * w[4];
      o = o0 * w[0] + o1 * w[1] + o2 * w[2] + o3 * w[3] + o4 * w[4];
    }
  } else if constexpr (BINOM) {
    e = e1 + e3 + (e2 << 1);
    o = o1 + o3 + (o2 << 1);
  } else {
    e = e1 * w[1] + e2 * w[2] + e3 * w[3];
    o = o1 * w[1] + o2 * w[2] + o3 * w[3];
  }
}

// acc (+)= h * w over the 12 lane words of a row; a set starts from the rounding bias.
template <bool INIT>
__device__ __forceinline__ void vacc(uint32_t (&acc)[12], const uint32_t (&h)[12], uint32_t w, uint32_t bias2) {
#pragma unroll
  for (int i = 0; i < 12; ++i) acc[i] = (INIT ? bias2 : acc[i]) + h[i] * w;
}

// MG: the multi-GPU hooks (neighbour flags, halo pushes, NVLS multicast output) are compiled in.
// The single-GPU instantiation carries none of them.
template <bool MG, bool FIVE, bool BINOM>
__global__ void __launch_bounds__(kS5Warps * 32, kS5CtasPerSm)
sweep5_kernel(const __grid_constant__ Sweep5Params P, const __grid_constant__ CUtensorMap src_map,
              const __grid_constant__ CUtensorMap dst_map, const __grid_constant__ CUtensorMap src_map1,
              const __grid_constant__ CUtensorMap dst_map1) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(256) uint8_t lut1[256];
  __shared__ __align__(8) uint64_t bars[kS5Warps * NST];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();   // the next launch in the stream may set itself up behind this one
  for (int i = threadIdx.x; i < 64; i += kS5Warps * 32)
    reinterpret_cast<uint32_t*>(lut1)[i] = reinterpret_cast<const uint32_t*>(P.lut1)[i];
  if (threadIdx.x == 0) {
    for (int i = 0; i < kS5Warps * NST; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
    tma_prefetch_desc(&src_map);
    if (!(MG && P.plain_dst_ptr)) tma_prefetch_desc(&dst_map);
  }
  __syncthreads();  // the only block-wide barrier of the kernel
  pdl_wait();       // the launch before this one has completed: its output may be this one's input

  uint8_t* in_ring = smem + warp * P.warp_smem;          // NST stages of CH rows x kS5InBytes
  uint8_t* out_ring = in_ring + NST * P.in_stage_bytes;  // 2 staging buffers of CH rows x kS5OutBytes
  uint64_t* bar = &bars[warp * NST];
  const bool halo_lane = lane == 0 || lane == 31;
  int wh[5], wv[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    wh[i] = P.wh[i];
    wv[i] = P.wv[i];
  }
  const uint32_t bias2 = (uint32_t)P.bias * 0x00010001u;
  const bool reflect = P.border == B2I_BORDER_REFLECT101, replicate = P.border == B2I_BORDER_REPLICATE;

  uint32_t n_in = 0, n_out = 0;  // boxes consumed / stored by this warp so far
  const int total_warps = gridDim.x * kS5Warps;
  // Tasks are dealt to warps CTA by CTA. With fewer tasks than warps (small bands) they are dealt across the
  // CTAs instead, so that every SM gets its share of the busy warps rather than some SMs a full CTA and others
  // none. With neighbours to push to, the boundary tasks (the first 2 n_strips, short, on every neighbour's
  // critical path) keep CTAs of their own and only the tasks between them are dealt across the other CTAs.
  int global_warp = (int)blockIdx.x * kS5Warps + warp;
  if (P.n_tasks < total_warps) {
    if (!MG) {   // (launches that only wait or signal measured slower dealt across: 12.6 -> 17.5 us on a 540-row band)
      global_warp = warp * (int)gridDim.x + (int)blockIdx.x;
    } else if (P.bseg > 0) {
      const int nb = 2 * P.n_strips, cb = (nb + kS5Warps - 1) / kS5Warps, rest = (int)gridDim.x - cb;
      if (rest > 0 && rest * kS5Warps >= P.n_tasks - nb) {
        if ((int)blockIdx.x < cb) global_warp = global_warp < nb ? global_warp : P.n_tasks;
        else global_warp = nb + warp * rest + ((int)blockIdx.x - cb);
      }
    }
  }
  // Resident multi-pass launches walk the passes here; every other launch is one pass.
  const int n_passes = MG ? max(1, P.n_passes) : 1;
  for (int pass = 0; pass < n_passes; ++pass) {
  const bool odd = MG && (pass & 1);
  const CUtensorMap* const smap = odd ? &src_map1 : &src_map;
  const CUtensorMap* const dmap = odd ? &dst_map1 : &dst_map;
  const bool multi = MG && n_passes > 1;
  const bool wait_nb = MG && (!multi || pass > 0);             // ghost rows of pass 0 come from the scatter
  const bool push_nb = MG && (!multi || pass + 1 < n_passes);  // nobody reads the last pass's halo
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
    const int steps = yb - ya + 4;                   // input rows ya-2 .. yb+1
    const int n_chunks = (steps + CH - 1) / CH;
    const int xs = strip * kS5Strip;
    const int x0 = xs - 8 + lane * 8;                // first pixel this lane holds
    const int xw_in = (xs * 3 - 32) >> 2, xw_out = (xs * 3) >> 2;
    // A task that reads ghost rows, or writes the neighbour's, first sees the neighbour's flag of
    // the previous pass: its rows have landed here and it has finished reading what we overwrite.
    // (near_*: within push_rows of an end of the launch. Those tasks wait; they also signal if there is a
    // flag to release. A launch may wait without signalling or the reverse: parallel/dist.py, exchange cycles.)
    const bool near_up = ya < P.y0 + P.push_rows, near_dn = yb > P.y1 - P.push_rows;
    const bool sig_up = push_nb && P.signal_up_flag && near_up;
    const bool sig_dn = push_nb && P.signal_dn_flag && near_dn;
    if (multi && pass > 0) {
      // the tasks around this one (strips s-1..s+1 x segments seg-2..seg+2) have finished the pass
      // before: their rows are what this task reads, and they have read what it overwrites
      if (lane < 15) {   // two segments up and down: a short remainder segment may lie in between
        const int ds = lane % 3 - 1, dg = lane / 3 - 2;
        const int s2 = strip + ds, g2 = seg + dg;
        if (s2 >= 0 && s2 < P.n_strips && g2 >= 0 && g2 < P.n_segs) {
          const int so2 = g2 == 0 ? 0 : g2 == P.n_segs - 1 ? 1 : g2 + 1;
          if (!wait_flag_ge_gpu(reinterpret_cast<const uint32_t*>(P.task_done) + so2 * P.n_strips + s2,
                                P.wait_epoch + (uint32_t)pass, P.timeout_ns, reinterpret_cast<const uint32_t*>(P.status)))
            atomicCAS(reinterpret_cast<uint32_t*>(P.status), 0u, 4u);
        }
      }
      __syncwarp();
      if (lane == 0) asm volatile("fence.proxy.async;" ::: "memory");  // the flags (generic proxy) before the TMA loads
    }
    if (MG && P.strip_flags && wait_nb) {
      // per-strip neighbour flags: the three strips above (lanes 0-2) / below (lanes 3-5) this one
      if (lane < 6) {
        const bool up = lane < 3;
        const int s2 = strip + lane % 3 - 1;
        const uint64_t base = up ? P.wait_up_flag : P.wait_dn_flag;
        const bool need = up ? (ya - 2 < P.y0 || near_up) : (yb + 2 > P.y1 || near_dn);
        if (base && need && s2 >= 0 && s2 < P.n_strips &&
            !wait_flag_ge(reinterpret_cast<const uint32_t*>(base) + s2, P.wait_epoch + (uint32_t)pass, P.timeout_ns,
                          reinterpret_cast<const uint32_t*>(P.status)))
          sweep_abort(P, up ? 1u : 2u, P.signal_epoch + (uint32_t)pass, P.n_strips);
      }
      __syncwarp();
      if (lane == 0) asm volatile("fence.proxy.async;" ::: "memory");
    }
    if (lane == 0) {
      if (wait_nb && !(MG && P.strip_flags)) sweep_wait_neighbours(P, ya, yb, 2, near_up, near_dn, (uint32_t)pass);
      for (int c = 0; c < min(NST, n_chunks); ++c) {
        const uint32_t st = (n_in + c) % NST;
        mbar_arrive_expect_tx(&bar[st], CH * kS5InBytes);
        tma_load_2d(in_ring + st * P.in_stage_bytes, smap, xw_in, ya - 2 + c * CH - P.src_y_begin, &bar[st]);
      }
    }
    const bool left_edge = x0 == 0, right_edge = x0 + 8 == P.W, beyond = x0 < 0 || x0 >= P.W;
    const bool edge_fix = (reflect || replicate) && (xs == 0 || xs + kS5Strip >= P.W);  // warp-uniform
    const bool vert_fix = (reflect || replicate) && (ya <= 2 || yb >= P.H - 2);        // warp-uniform

    uint32_t X0[12], X1[12], X2[12], X3[12], X4[12];  // output row o collects in X[o mod 5]
#pragma unroll
    for (int i = 0; i < 12; ++i) X0[i] = X1[i] = X2[i] = X3[i] = X4[i] = 0;

    // one input row: horizontal pass in registers, then into the five output rows it touches
    // SLOW (compile-time): the task also pushes halo rows to a neighbour or stores to a multicast
    // image. Only the first and last segments of a band do the former, so every other task of a
    // multi-GPU launch runs the same row loop as the single-GPU kernel.
    auto step = [&](auto qc, auto slowc, int c, const uint8_t* stage_base, uint32_t ob_prev, uint32_t ob_cur,
                    uint32_t (&K0)[12], uint32_t (&K1)[12], uint32_t (&K2)[12], uint32_t (&K3)[12],
                    uint32_t (&K4)[12]) {
      constexpr int q = decltype(qc)::value;
      constexpr bool SLOW = MG && decltype(slowc)::value;
      const int u = c * CH + q;
      const int y = ya - 2 + u;
      uint32_t a[10];
      {
        const uint8_t* s = stage_base + q * kS5InBytes + 8 + lane * 24;
        const uint2 v0 = *reinterpret_cast<const uint2*>(s), v1 = *reinterpret_cast<const uint2*>(s + 8),
                    v2 = *reinterpret_cast<const uint2*>(s + 16);
        // rows outside the image (zero-filled by TMA) and lanes beyond the right edge contribute 0;
        // mirrored / replicated rows are folded into the multipliers below
        const bool valid = (unsigned)y < (unsigned)P.H && !beyond;
        a[2] = valid ? v0.x : 0u; a[3] = valid ? v0.y : 0u; a[4] = valid ? v1.x : 0u;
        a[5] = valid ? v1.y : 0u; a[6] = valid ? v2.x : 0u; a[7] = valid ? v2.y : 0u;
      }
      a[0] = __shfl_up_sync(0xFFFFFFFFu, a[6], 1);
      a[1] = __shfl_up_sync(0xFFFFFFFFu, a[7], 1);
      a[8] = __shfl_down_sync(0xFFFFFFFFu, a[2], 1);
      a[9] = __shfl_down_sync(0xFFFFFFFFu, a[3], 1);
      if (edge_fix) {
        if (reflect) {  // pixels -1,-2 := pixels 1,2; pixels W,W+1 := pixels W-2,W-3
          if (left_edge) {
            a[1] = __byte_perm(__byte_perm(a[2], a[3], 0x5433), a[4], 0x3214);
            a[0] = __byte_perm(a[3], 0u, 0x3200);
          }
          if (right_edge) {
            a[8] = __byte_perm(__byte_perm(a[6], a[7], 0x4432), a[5], 0x7210);
            a[9] = __byte_perm(a[6], 0u, 0x1010);
          }
        } else {        // replicate: pixels -1,-2 := pixel 0; pixels W,W+1 := pixel W-1
          if (left_edge) {
            a[1] = __byte_perm(a[2], 0u, 0x2102);
            a[0] = __byte_perm(a[2], 0u, 0x1000);
          }
          if (right_edge) {
            a[8] = __byte_perm(a[7], 0u, 0x1321);
            a[9] = __byte_perm(a[7], 0u, 0x3232);
          }
        }
      }
      uint32_t h[12];
      hword<0, FIVE, BINOM>(a, wh, h[0], h[1]);
      hword<1, FIVE, BINOM>(a, wh, h[2], h[3]);
      hword<2, FIVE, BINOM>(a, wh, h[4], h[5]);
      hword<3, FIVE, BINOM>(a, wh, h[6], h[7]);
      hword<4, FIVE, BINOM>(a, wh, h[8], h[9]);
      hword<5, FIVE, BINOM>(a, wh, h[10], h[11]);
      // multipliers of this row for the output rows y+2 (k0) .. y-2 (k4). A row that also stands
      // for a mirrored / replicated row outside the image simply weighs more (warp-uniform).
      int k0 = wv[0], k1 = wv[1], k2 = wv[2], k3 = wv[3], k4 = wv[4];
      if (vert_fix) {
        if (reflect) {
          if (y == 1) { k2 += wv[0]; k3 += wv[1]; }
          if (y == 2) { k4 += wv[0]; }
          if (y == P.H - 2) { k1 += wv[3]; k2 += wv[4]; }
          if (y == P.H - 3) { k0 += wv[4]; }
        } else {
          if (y == 0) { k1 += wv[0]; k2 += wv[1] + wv[0]; }
          if (y == P.H - 1) { k2 += wv[3] + wv[4]; k3 += wv[4]; }
        }
      }
      if constexpr (FIVE) {
        vacc<true>(K0, h, (uint32_t)k0, bias2);
        vacc<false>(K1, h, (uint32_t)k1, bias2);
        vacc<false>(K2, h, (uint32_t)k2, bias2);
        vacc<false>(K3, h, (uint32_t)k3, bias2);
        vacc<false>(K4, h, (uint32_t)k4, bias2);
      } else {  // 3 taps: the outer multipliers are zero; the set of row y+1 starts here, K4 is long complete
        vacc<true>(K1, h, (uint32_t)k1, bias2);
        vacc<false>(K2, h, (uint32_t)k2, bias2);
        vacc<false>(K3, h, (uint32_t)k3, bias2);
      }
      // ---- output row y-2 (index e = u-4 of the task) is complete --------------------------------
      if (u >= 4) {
        constexpr int er = (q + 1) % 5;  // row of the store box
        const uint32_t ob = er == 0 ? ob_cur : ob_prev;
        if (er == 0) {                   // first row of a box: its previous store has read it
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
        }
        uint32_t r[6];
#pragma unroll
        for (int m = 0; m < 6; ++m) {
          r[m] = __byte_perm(K4[2 * m], K4[2 * m + 1], 0x7351);  // >> 8 (the planner scales the taps to 256) and narrow
        }
        if (P.has_lut1) {
#pragma unroll
          for (int m = 0; m < 6; ++m) r[m] = lut4q(lut1, r[m]);
        }
        if (!halo_lane) {   // predicated: the halo lanes stage nothing (a spare slot for them cost bank conflicts)
          uint2* d = reinterpret_cast<uint2*>(out_ring + ob * P.out_buf_bytes + er * kS5OutBytes + (lane - 1) * 24);
          d[0] = make_uint2(r[0], r[1]);
          d[1] = make_uint2(r[2], r[3]);
          d[2] = make_uint2(r[4], r[5]);
        }
        if (SLOW && push_nb && P.push_rows && !halo_lane && !beyond)  // halo rows also go into the neighbours' ghost rows
          sweep_push_row(P, y - 2, (int64_t)xs * 3 + (lane - 1) * 24, r, odd);
        if (SLOW && P.plain_dst_ptr && (er == 4 || u == steps - 1)) {
          // NVLS: the box leaves in coalesced 16-byte stores to the multicast image, one store per
          // byte, and the switch writes every GPU's copy
          const int ec = (u - 4) / CH;
          mc_store_box(out_ring + ob * P.out_buf_bytes, er + 1, kS5OutBytes, min(kS5OutBytes, (P.W - xs) * 3),
                       P.plain_dst_ptr + (int64_t)(ya + ec * CH - P.y0) * P.plain_dst_pitch + (int64_t)xs * 3,
                       P.plain_dst_pitch, lane);
        }
        if (!(SLOW && P.plain_dst_ptr) && (er == 4 || u == steps - 1)) {  // a box of finished rows leaves by TMA
          const int ec = (u - 4) / CH;
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(dmap, xw_out, ya + ec * CH - P.y0, out_ring + ob * P.out_buf_bytes);
            tma_store_commit();
          }
        }
      }
    };

    auto run_chunks = [&](auto slowc) {
    for (int c = 0; c < n_chunks; ++c) {
      const uint32_t st = (n_in + c) % NST;
      const uint8_t* stage_base = in_ring + st * P.in_stage_bytes;
      // output box c-1 is finished during rows 0..3 of input box c; box c starts at row 4
      const uint32_t ob_prev = (n_out + c + 1) & 1, ob_cur = (n_out + c) & 1;
      const int rows_here = min(CH, steps - c * CH);
      mbar_wait(&bar[st], ((n_in + c) / NST) & 1, MG ? reinterpret_cast<uint32_t*>(P.status) : nullptr);
      step(std::integral_constant<int, 0>{}, slowc, c, stage_base, ob_prev, ob_cur, X2, X1, X0, X4, X3);
      if (rows_here > 1) step(std::integral_constant<int, 1>{}, slowc, c, stage_base, ob_prev, ob_cur, X3, X2, X1, X0, X4);
      if (rows_here > 2) step(std::integral_constant<int, 2>{}, slowc, c, stage_base, ob_prev, ob_cur, X4, X3, X2, X1, X0);
      if (rows_here > 3) step(std::integral_constant<int, 3>{}, slowc, c, stage_base, ob_prev, ob_cur, X0, X4, X3, X2, X1);
      if (rows_here > 4) step(std::integral_constant<int, 4>{}, slowc, c, stage_base, ob_prev, ob_cur, X1, X0, X4, X3, X2);
      __syncwarp();  // every lane has read the box: the stage can take box c + NST
      if (lane == 0 && c + NST < n_chunks) {
        mbar_arrive_expect_tx(&bar[st], CH * kS5InBytes);
        tma_load_2d(in_ring + st * P.in_stage_bytes, smap, xw_in, ya - 2 + (c + NST) * CH - P.src_y_begin,
                    &bar[st]);
      }
    }
    };
    if (MG && ((push_nb && P.push_rows && (ya < P.y0 + P.push_rows || yb > P.y1 - P.push_rows)) || P.plain_dst_ptr))
      run_chunks(std::true_type{});
    else
      run_chunks(std::false_type{});
    n_in += n_chunks;
    n_out += (yb - ya + CH - 1) / CH;
    if (MG && P.strip_flags) {
      if (sig_up || sig_dn) {   // this strip's halo rows are at the neighbour: release its flag for this strip
        __threadfence_system();
        __syncwarp();
        if (lane == 0) {
          if (sig_up) st_release_sys(reinterpret_cast<uint32_t*>(P.signal_up_flag) + strip, P.signal_epoch + (uint32_t)pass);
          if (sig_dn) st_release_sys(reinterpret_cast<uint32_t*>(P.signal_dn_flag) + strip, P.signal_epoch + (uint32_t)pass);
        }
      }
    } else if (sig_up || sig_dn) {
      sweep_signal_neighbours(P, sig_up, sig_dn, lane, (uint32_t)pass);
    }
    if (multi && pass + 1 < n_passes) {
      // this task's rows of the pass are in memory: publish it to the tasks around it
      __syncwarp();
      if (lane == 0) {
        tma_store_wait_all();
        asm volatile("fence.proxy.async;" ::: "memory");
        st_release_gpu(reinterpret_cast<uint32_t*>(P.task_done) + task, P.wait_epoch + (uint32_t)pass + 1);
      }
    }
  }
  }
  sweep_kernel_end(P, global_warp, total_warps, lane);
}

}  // namespace

// src_map1 / dst_map1: the other direction of a resident multi-pass launch (P.n_passes > 1), which
// is launched cooperatively: its tasks wait for one another, so every CTA has to be resident.
cudaError_t launch_sweep5(const Sweep5Params& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                          int grid, cudaStream_t stream, const CUtensorMap* src_map1, const CUtensorMap* dst_map1) {
  static bool configured[64][8] = {{false}};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const bool mg = P.push_rows || P.wait_up_flag || P.wait_dn_flag || P.signal_up_flag || P.signal_dn_flag ||
                  P.plain_dst_ptr || P.n_passes > 1;
  const bool five = (P.wh[0] | P.wh[4]) != 0;
  const int variant = (mg ? 4 : 0) + (five ? 2 : 0) + (P.binom ? 1 : 0);
  using Kern = void (*)(Sweep5Params, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap);
  static const Kern table[8] = {
      sweep5_kernel<false, false, false>, sweep5_kernel<false, false, true>, sweep5_kernel<false, true, false>,
      sweep5_kernel<false, true, true>,   sweep5_kernel<true, false, false>, sweep5_kernel<true, false, true>,
      sweep5_kernel<true, true, false>,   sweep5_kernel<true, true, true>};
  const Kern kern = table[variant];
  if (dev < 64 && !configured[dev][variant]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, P.warp_smem * kS5Warps);
    if (e != cudaSuccess) return e;
    configured[dev][variant] = true;
  }
  if (P.n_passes > 1) {
    if (!src_map1 || !dst_map1) return cudaErrorInvalidValue;
    void* args[5] = {(void*)&P, (void*)&src_map, (void*)&dst_map, (void*)src_map1, (void*)dst_map1};
    return cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(kS5Warps * 32), args,
                                       (size_t)P.warp_smem * kS5Warps, stream);
  }
  // Measured on 540-row bands (7680 wide, gauss5): alone on a GPU, dealing the tasks across the CTAs takes a pass
  // from 14.6 to 12.5 us and the dependent launch to 10.2; with neighbours to push to and wait for, both cost time
  // (17.4 -> 20.0 us): launches with multi-GPU hooks keep the CTA-by-CTA order, and the plain launch on small bands.
  const bool pushes = P.push_up_ptr || P.push_dn_ptr;
  return launch_pdl(!pushes || P.y1 - P.y0 >= pdl_min_rows_mg(), kern, grid, kS5Warps * 32, (size_t)P.warp_smem * kS5Warps, stream, P, src_map,
                    dst_map, src_map, dst_map);
}
}  // namespace b2i
