// Launchers of the sweep kernels (kernels/sweep3.cu, sweep5.cu, sweep_sobel.cu): eligibility,
// parameters, tensor maps, segment heights, multi-GPU hooks.
#include <algorithm>
#include <cstring>

#include "plan_internal.hpp"

namespace b2i {

// Completion flags of ROOT mode, fused into the end of a sweep kernel (device_utils.cuh:
// sweep_finish). Tells b2i_cuda_run that no separate wait kernel is needed.
thread_local bool g_wait_done_fused = false;
template <class Params>
static void sweep_finish_fill(Params& P, const b2i_launch_t& L) {
  P.done_flag = L.done_flag;
  P.wait_done_flags = L.wait_done_flags;
  P.wait_done_mask = L.wait_done_flags ? L.wait_done_mask : 0;
  P.signal_epoch = L.signal_epoch;
  P.counters = L.counters;
  P.status = L.status;
  P.timeout_ns = L.timeout_ns ? L.timeout_ns : 2000000000ull;
  g_wait_done_fused = true;
}

static int sweep_seg_rows(int n_strips, int rows, int resident, int chunk, int lo, int hi, int halo_rows);
// CTAs of a sweep launch: full CTAs when there are enough tasks, and never fewer CTAs than SMs when the
// tasks allow it (the kernels deal tasks to warps CTA by CTA, so a small band keeps every SM partly busy
// instead of some SMs full and others idle: 7680 x 540 gauss5 14.6 -> 12.6 us per pass).
static int sweep_grid(int n_tasks, int warps, int sms, int ctas_per_sm) {
  const int full = (n_tasks + warps - 1) / warps;
  return std::max(1, std::min(sms * ctas_per_sm, std::max(full, std::min(n_tasks, sms))));
}

// The warp-sweep kernel (kernels/sweep3.cu) takes segments of the reference's own shape:
// gray (+LUT) -> one 3x3 stencil on the gray plane (+LUT, +expand). Returns 1 if it launched,
// 0 if the segment or the buffers do not qualify (the tile-chain kernel then runs it).
int try_sweep(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
              const b2i_view_t& src, const b2i_view_t& dst, int sms, bool last) {
  if (seg.phases.size() != 2 || B2I_ENV("B200IMG_NO_SWEEP", 0) || (L.flags & B2I_LAUNCH_NO_TMA) ||
      B2I_ENV("B200IMG_NO_TMA", 0))
    return 0;
  const PhaseDesc& g = seg.phases[0];
  const PhaseDesc& st = seg.phases[1];
  if ((g.type != PH_GRAY_REF && g.type != PH_GRAY_CV) || st.type != PH_STENCIL || st.ksize != 3 ||
      st.cin != 1 || seg.cin != 3)
    return 0;
  if ((L.push_rows && (L.push_up.ptr || L.push_dn.ptr)) || L.wait_up_flag || L.wait_dn_flag ||
      L.signal_up_flag || L.signal_dn_flag)
    return 0;
  const bool finish = last && (L.done_flag || L.wait_done_mask);
  if (finish && !(L.counters && L.status)) return 0;
  // Widths that are not a multiple of 8 pixels: fine under the reference's border rule (kernel.cu:83:
  // the last columns pass through, nothing is mirrored at a lane boundary) when the rows have room for
  // the tensor maps' 4-byte granularity, i.e. in pitched buffers. Mirrored / replicated borders fold
  // at lane boundaries and keep the multiple-of-8 rule (the tensor-core kernel takes those).
  const int64_t in_row = ((int64_t)L.W * 3 + 3) & ~3LL, out_row_b = ((int64_t)L.W * seg.cout + 3) & ~3LL;
  if (L.W & 7) {
    if (st.border != B2I_BORDER_REF || src.pitch < in_row || dst.pitch < out_row_b) return 0;
  }
  if (L.W < 16 || y1 - y0 < 1) return 0;
  SweepParams P;
  memset(&P, 0, sizeof P);
  P.W = L.W;
  P.H = L.H;
  P.y0 = y0;
  P.y1 = y1;
  P.src_y_begin = src.y_begin;
  P.gray = g.type == PH_GRAY_REF ? 1 : 2;
  P.rgb_order = plan->chain.order == B2I_ORDER_RGB;
  P.has_lut0 = g.lut >= 0;
  P.has_lut1 = st.lut >= 0;
  if (g.lut >= 0) memcpy(P.lut0, seg.luts[g.lut], 256);
  if (st.lut >= 0) memcpy(P.lut1, seg.luts[st.lut], 256);
  if (g.lut >= 0 && seg.lut0_affine && !B2I_ENV("B200IMG_NO_AFFINE_LUT", 0)) {
    P.has_lut0 = 2;
    P.lut0_a = seg.aff_a;
    P.lut0_b = seg.aff_b;
    P.lut0_s = seg.aff_s;
  }
  P.border = st.border;
  P.shift = st.shift;
  P.bias = st.bias;
  P.expand = st.expand;
  for (int k = 0; k < 3; ++k) P.wpk[k] = st.wpk[k][0];
  const int out_row = st.expand ? kSweepOutBytes : kSweepStrip;
  P.in_stage_bytes = align_up(kSweepChunk * kSweepInBytes, 128);
  P.out_buf_bytes = align_up(kSweepChunk * out_row, 128);
  P.warp_smem = kSweepStages * P.in_stage_bytes + 2 * P.out_buf_bytes + 128;  // + dummy slot for the halo lanes
  CUtensorMap src_map, dst_map;
  const int src_rows = std::min(src.y_end, L.H) - src.y_begin;
  if (!encode_map(&src_map, src.ptr, src.pitch, in_row, src_rows, kSweepInBytes, kSweepChunk)) return 0;
  const uint64_t dptr = dst.ptr + (uint64_t)((int64_t)(y0 - dst.y_begin) * dst.pitch);
  const bool plain = last && (L.flags & B2I_LAUNCH_DST_MCAST);
  if (plain) {  // NVLS multicast image: 16-byte stores from the staging buffer, no tensor map on that address
    if (((dptr | (uint64_t)dst.pitch) & 15) || (((int64_t)L.W * seg.cout) & 15)) return 0;
    memset(&dst_map, 0, sizeof dst_map);
    P.plain_dst_ptr = dptr;
    P.plain_dst_pitch = dst.pitch;
  } else if (!encode_map(&dst_map, dptr, dst.pitch, out_row_b, y1 - y0, out_row, kSweepChunk)) {
    return 0;
  }
  // segment height: a multiple of the box height, chosen so that the tasks fill the resident
  // warps evenly and the two halo rows per segment stay cheap
  P.n_strips = (L.W + kSweepStrip - 1) / kSweepStrip;
  const int resident = sms * kSweepCtasPerSm * kSweepWarps;
  P.seg_rows = sweep_seg_rows(P.n_strips, y1 - y0, resident, kSweepChunk, 6, 96, 2);
  P.n_tasks = P.n_strips * ((y1 - y0 + P.seg_rows - 1) / P.seg_rows);
  const int grid = sweep_grid(P.n_tasks, kSweepWarps, sms, kSweepCtasPerSm);
  plan->info.last_tile_h = -P.seg_rows;  // negative: rows per sweep segment
  plan->info.last_smem_bytes = P.warp_smem * kSweepWarps;
  plan->info.last_grid = grid;
  plan->info.last_blocks_per_sm = kSweepCtasPerSm;
  plan->info.last_tma_load = plan->info.last_tma_store = 1;
  plan->info.last_launches++;
  if (finish) sweep_finish_fill(P, L);
  cudaError_t e = launch_sweep3(P, src_map, dst_map, grid, (cudaStream_t)L.stream);
  if (e != cudaSuccess) {
    b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "sweep3 launch: %s", cudaGetErrorString(e));
    return -1;
  }
  return 1;
}

// Multi-GPU hooks of the sweep kernels (sweep5.cu, sweep_sobel.cu), with the tile kernel's meaning:
// waits guard the first segment's ghost-row reads, halo pushes and signals ride on the last
// segment's result rows. Returns false if the launch asks for something the sweeps do not do (the
// tile kernel then runs it, or reports the error).
static bool sweep_hooks_ok(const b2i_launch_t& L, int y0, int y1, bool first, bool last, bool* pushes, bool* flags) {
  *pushes = last && L.push_rows > 0 && (L.push_up.ptr || L.push_dn.ptr);
  *flags = (first && (L.wait_up_flag || L.wait_dn_flag)) || (last && (L.signal_up_flag || L.signal_dn_flag));
  if (*pushes) {
    if (L.push_rows > y1 - y0 || L.push_rows > kGhostRows) return false;
    if (L.push_up.ptr && L.push_dn.ptr && L.push_up.pitch != L.push_dn.pitch) return false;
    const b2i_view_t& v = L.push_up.ptr ? L.push_up : L.push_dn;
    if ((v.pitch & 7) || ((L.push_up.ptr | L.push_dn.ptr) & 7)) return false;
    if (L.push_up.ptr && (L.push_up.y_begin > y0 || L.push_up.y_end < y0 + L.push_rows)) return false;
    if (L.push_dn.ptr && (L.push_dn.y_begin > y1 - L.push_rows || L.push_dn.y_end < y1)) return false;
  }
  if ((*flags || (last && (L.done_flag || L.wait_done_mask))) && !(L.counters && L.status)) return false;
  return true;
}


template <class Params>
static void sweep_hooks_fill(Params& P, const b2i_launch_t& L, int y0, int y1, bool first, bool last, bool pushes) {
  P.push_rows = last ? L.push_rows : 0;
  if (pushes) {
    P.push_pitch = (int32_t)(L.push_up.ptr ? L.push_up.pitch : L.push_dn.pitch);
    if (L.push_up.ptr)
      P.push_up_ptr = L.push_up.ptr + (uint64_t)((int64_t)(y0 - L.push_up.y_begin) * L.push_up.pitch);
    if (L.push_dn.ptr)
      P.push_dn_ptr = L.push_dn.ptr + (uint64_t)((int64_t)(y1 - L.push_rows - L.push_dn.y_begin) * L.push_dn.pitch);
  }
  if (first) {
    P.wait_up_flag = L.wait_up_flag;
    P.wait_dn_flag = L.wait_dn_flag;
    P.wait_epoch = L.wait_epoch;
  }
  if (last) {
    P.signal_up_flag = L.signal_up_flag;
    P.signal_dn_flag = L.signal_dn_flag;
    P.signal_epoch = L.signal_epoch;
  }
  // segments whose rows reach into the first / last push_rows rows of the band (exact: the last
  // of these tasks releases the neighbour's flag)
  P.n_up_tasks = P.n_strips * ((P.push_rows + P.seg_rows - 1) / P.seg_rows);
  P.n_dn_tasks = P.n_strips * (P.n_segs - (y1 - y0 - P.push_rows) / P.seg_rows);
  P.counters = L.counters;
  P.status = L.status;
  P.timeout_ns = L.timeout_ns ? L.timeout_ns : 2000000000ull;
}

// Rows per sweep segment: a multiple of the box height that maximises rows per warp-time. A task
// costs its rows, the halo rows it reads again, and a fixed part (first TMA round trip, pipeline
// fill, store drain: about six rows' worth, measured on small bands with scripts/resident_probe.py).
// Big images end up with one or two waves of tall segments; small bands (a resident pass at N = 8
// is 540 rows per GPU) with short ones, where the run time is one task's latency.
static int sweep_seg_rows(int n_strips, int rows, int resident, int chunk, int lo, int hi, int halo_rows) {
  double best = -1;
  int pick = lo;
  for (int sr = lo; sr <= hi; sr += chunk) {
    const long long tasks = (long long)n_strips * ((rows + sr - 1) / sr);
    const long long waves = (tasks + resident - 1) / resident;
    const double eff = (double)tasks / (double)(waves * resident) * sr / (sr + (double)halo_rows + 6.0);
    if (eff > best + 1e-9) {
      best = eff;
      pick = sr;
    }
  }
  return pick;
}

// The separable sweep kernel (kernels/sweep5.cu) takes segments that are exactly one separable
// stencil on 3-channel data (+LUT): the 5x5 / 3x3 Gaussians. Returns 1 if it launched, 0 if the
// segment or the buffers do not qualify, -1 on a launch error.
int try_sweep5(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
               const b2i_view_t& src, const b2i_view_t& dst, int sms, bool first, bool last) {
  if (seg.phases.size() != 2 || B2I_ENV("B200IMG_NO_SWEEP", 0) || B2I_ENV("B200IMG_NO_SWEEP5", 0) ||
      (L.flags & B2I_LAUNCH_NO_TMA) || B2I_ENV("B200IMG_NO_TMA", 0))
    return 0;
  const PhaseDesc& ph = seg.phases[0];
  const PhaseDesc& pv = seg.phases[1];
  if (ph.type != PH_SEPH || pv.type != PH_SEPV || seg.cin != 3 || seg.cout != 3 || pv.expand) return 0;
  if ((L.W & 7) || L.W < 16 || L.H < 8 || y1 - y0 < 1) return 0;
  bool pushes = false, flags = false;
  if (!sweep_hooks_ok(L, y0, y1, first, last, &pushes, &flags)) return 0;
  Sweep5Params P;
  memset(&P, 0, sizeof P);
  P.W = L.W;
  P.H = L.H;
  P.y0 = y0;
  P.y1 = y1;
  P.src_y_begin = src.y_begin;
  P.border = ph.border;
  // the kernel narrows with a fixed >> 8: scale the vertical taps (and the bias) so that all taps
  // sum to 256; lanes stay below 65536 (255 * 256 + bias)
  const int up = 8 - pv.shift;
  if (up < 0) return 0;
  P.bias = pv.bias << up;
  P.has_lut1 = pv.lut >= 0;
  if (pv.lut >= 0) memcpy(P.lut1, seg.luts[pv.lut], 256);
  const int off = ph.ksize == 5 ? 0 : 1;  // a 3-tap kernel sits in the middle of the five slots
  for (int i = 0; i < ph.ksize; ++i) {
    P.wh[i + off] = ph.sep[i];
    P.wv[i + off] = pv.sep[i] << up;
  }
  static const int b5[5] = {1, 4, 6, 4, 1}, b3[5] = {0, 1, 2, 1, 0};
  P.binom = !memcmp(P.wh, b5, sizeof b5) || !memcmp(P.wh, b3, sizeof b3);
  P.in_stage_bytes = align_up(kS5Chunk * kS5InBytes, 128);
  P.out_buf_bytes = align_up(kS5Chunk * kS5OutBytes, 128);
  P.warp_smem = kS5Stages * P.in_stage_bytes + 2 * P.out_buf_bytes + 128;
  CUtensorMap src_map, dst_map;
  const int src_rows = std::min(src.y_end, L.H) - src.y_begin;
  if (!encode_map(&src_map, src.ptr, src.pitch, (int64_t)L.W * 3, src_rows, kS5InBytes, kS5Chunk)) return 0;
  const uint64_t dptr = dst.ptr + (uint64_t)((int64_t)(y0 - dst.y_begin) * dst.pitch);
  const bool plain = last && (L.flags & B2I_LAUNCH_DST_MCAST);
  if (plain) {  // NVLS multicast image: 16-byte stores from the staging buffer, no tensor map on that address
    if (((dptr | (uint64_t)dst.pitch) & 15) || (((int64_t)L.W * seg.cout) & 15)) return 0;
    memset(&dst_map, 0, sizeof dst_map);
    P.plain_dst_ptr = dptr;
    P.plain_dst_pitch = dst.pitch;
  } else if (!encode_map(&dst_map, dptr, dst.pitch, (int64_t)L.W * 3, y1 - y0, kS5OutBytes, kS5Chunk)) {
    return 0;
  }
  P.n_strips = (L.W + kS5Strip - 1) / kS5Strip;
  const int resident = sms * kS5CtasPerSm * kS5Warps;
  if (pushes && y1 - y0 >= 6 * kS5Chunk && L.push_rows <= 2 * kS5Chunk && !B2I_ENV("B200IMG_NO_BSEG", 0)) {
    // a launch that pushes halo rows: first and last segments of one box (two when more than five rows
    // leave: exchange cycles), which run first and are short, so the rows and the flag are with the
    // neighbour while the other tasks still work
    // (every segment but the last is whole boxes: a ragged box would spill into the rows below it)
    P.bseg = (L.push_rows + kS5Chunk - 1) / kS5Chunk * kS5Chunk;
    const int mid = (y1 - y0 - 2 * P.bseg) / kS5Chunk * kS5Chunk;
    P.bseg_last = y1 - y0 - P.bseg - mid;
    P.seg_rows = sweep_seg_rows(P.n_strips, mid, std::max(kS5Warps, resident - 2 * P.n_strips), kS5Chunk, 5, 120, 4);
    P.n_segs = 2 + (mid + P.seg_rows - 1) / P.seg_rows;
  } else {
    P.seg_rows = sweep_seg_rows(P.n_strips, y1 - y0, resident, kS5Chunk, 5, 120, 4);
    P.n_segs = (y1 - y0 + P.seg_rows - 1) / P.seg_rows;
  }
  P.n_tasks = P.n_strips * P.n_segs;
  if (pushes || flags) sweep_hooks_fill(P, L, y0, y1, first, last, pushes);
  if (P.bseg) P.n_up_tasks = P.n_dn_tasks = P.n_strips;   // the short segments hold all the pushed rows
  const int grid = sweep_grid(P.n_tasks, kS5Warps, sms, kS5CtasPerSm);
  plan->info.last_tile_h = -P.seg_rows;
  plan->info.last_smem_bytes = P.warp_smem * kS5Warps;
  plan->info.last_grid = grid;
  plan->info.last_blocks_per_sm = kS5CtasPerSm;
  plan->info.last_tma_load = plan->info.last_tma_store = 1;
  plan->info.last_launches++;
  if (last && (L.done_flag || L.wait_done_mask)) sweep_finish_fill(P, L);
  cudaError_t e = launch_sweep5(P, src_map, dst_map, grid, (cudaStream_t)L.stream);
  if (e != cudaSuccess) {
    b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "sweep5 launch: %s", cudaGetErrorString(e));
    return -1;
  }
  return 1;
}

// The Sobel sweep kernel (kernels/sweep_sobel.cu) takes segments that are exactly one Sobel
// magnitude on 3-channel data (+LUT). Same return values as try_sweep5.
int try_sweep_sobel(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
                    const b2i_view_t& src, const b2i_view_t& dst, int sms, bool first, bool last) {
  if (seg.phases.size() != 1 || B2I_ENV("B200IMG_NO_SWEEP", 0) || B2I_ENV("B200IMG_NO_SWEEP_SOBEL", 0) ||
      (L.flags & B2I_LAUNCH_NO_TMA) || B2I_ENV("B200IMG_NO_TMA", 0))
    return 0;
  const PhaseDesc& ph = seg.phases[0];
  if (ph.type != PH_SOBEL || seg.cin != 3 || seg.cout != 3 || ph.expand || ph.border == B2I_BORDER_REF) return 0;
  if ((L.W & 7) || L.W < 16 || L.H < 8 || y1 - y0 < 1) return 0;
  bool pushes = false, flags = false;
  if (!sweep_hooks_ok(L, y0, y1, first, last, &pushes, &flags)) return 0;
  SweepSobelParams P;
  memset(&P, 0, sizeof P);
  P.W = L.W;
  P.H = L.H;
  P.y0 = y0;
  P.y1 = y1;
  P.src_y_begin = src.y_begin;
  P.border = ph.border;
  P.has_lut1 = ph.lut >= 0;
  if (ph.lut >= 0) memcpy(P.lut1, seg.luts[ph.lut], 256);
  P.in_stage_bytes = align_up(kSSChunk * kSSInBytes, 128);
  P.out_buf_bytes = align_up(kSSChunk * kSSOutBytes, 128);
  P.warp_smem = kSSStages * P.in_stage_bytes + 2 * P.out_buf_bytes + 128;
  CUtensorMap src_map, dst_map;
  const int src_rows = std::min(src.y_end, L.H) - src.y_begin;
  if (!encode_map(&src_map, src.ptr, src.pitch, (int64_t)L.W * 3, src_rows, kSSInBytes, kSSChunk)) return 0;
  const uint64_t dptr = dst.ptr + (uint64_t)((int64_t)(y0 - dst.y_begin) * dst.pitch);
  const bool plain = last && (L.flags & B2I_LAUNCH_DST_MCAST);
  if (plain) {  // NVLS multicast image: 16-byte stores from the staging buffer, no tensor map on that address
    if (((dptr | (uint64_t)dst.pitch) & 15) || (((int64_t)L.W * seg.cout) & 15)) return 0;
    memset(&dst_map, 0, sizeof dst_map);
    P.plain_dst_ptr = dptr;
    P.plain_dst_pitch = dst.pitch;
  } else if (!encode_map(&dst_map, dptr, dst.pitch, (int64_t)L.W * 3, y1 - y0, kSSOutBytes, kSSChunk)) {
    return 0;
  }
  P.n_strips = (L.W + kSSStrip - 1) / kSSStrip;
  const int resident = sms * kSSCtasPerSm * kSSWarps;
  if (pushes && y1 - y0 >= 4 * kSSChunk && L.push_rows <= kSSChunk && !B2I_ENV("B200IMG_NO_BSEG", 0)) {
    // one-box first and last segments when halo rows are pushed (see try_sweep5)
    P.bseg = kSSChunk;
    const int mid = (y1 - y0 - 2 * kSSChunk) / kSSChunk * kSSChunk;
    P.bseg_last = y1 - y0 - P.bseg - mid;
    P.seg_rows = sweep_seg_rows(P.n_strips, mid, std::max(kSSWarps, resident - 2 * P.n_strips), kSSChunk, 4, 120, 2);
    P.n_segs = 2 + (mid + P.seg_rows - 1) / P.seg_rows;
  } else {
    P.seg_rows = sweep_seg_rows(P.n_strips, y1 - y0, resident, kSSChunk, 4, 120, 2);
    P.n_segs = (y1 - y0 + P.seg_rows - 1) / P.seg_rows;
  }
  P.n_tasks = P.n_strips * P.n_segs;
  if (pushes || flags) sweep_hooks_fill(P, L, y0, y1, first, last, pushes);
  if (P.bseg) P.n_up_tasks = P.n_dn_tasks = P.n_strips;
  const int grid = sweep_grid(P.n_tasks, kSSWarps, sms, kSSCtasPerSm);
  plan->info.last_tile_h = -P.seg_rows;
  plan->info.last_smem_bytes = P.warp_smem * kSSWarps;
  plan->info.last_grid = grid;
  plan->info.last_blocks_per_sm = kSSCtasPerSm;
  plan->info.last_tma_load = plan->info.last_tma_store = 1;
  plan->info.last_launches++;
  if (last && (L.done_flag || L.wait_done_mask)) sweep_finish_fill(P, L);
  cudaError_t e = launch_sweep_sobel(P, src_map, dst_map, grid, (cudaStream_t)L.stream);
  if (e != cudaSuccess) {
    b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "sweep_sobel launch: %s", cudaGetErrorString(e));
    return -1;
  }
  return 1;
}

}  // namespace b2i

// ---- resident multi-pass launch -------------------------------------------------------------------
// `passes` passes of a one-segment separable stencil (the Gaussians) over a resident band in ONE
// cooperative launch of sweep5_kernel<MG>: the kernel ping-pongs between the two band buffers that
// `even` (buffer 0 -> 1) and `odd` (1 -> 0) describe, orders its own tasks with one flag per task
// and its neighbours with the epoch flags (kernels/sweep5.hpp). What stood here before was one
// launch per pass: at 8 GPUs an 8K band is 540 rows, and launch, pipeline fill and drain cost more
// than the rows. The reference has neither passes nor halos (kernel.cu:195 embosses bands in
// isolation between two MPI_Barriers, kernel.cu:127,133).
extern "C" int b2i_cuda_run_resident_fused(b2i_plan_t* plan, const b2i_launch_t* even, const b2i_launch_t* odd,
                                           int passes, uint32_t epoch_base, uint64_t task_flags,
                                           int n_task_flags) {
  using namespace b2i;
  B2I_CHECK_ARG(plan && even && odd && passes >= 1 && task_flags, "bad arguments");
  const b2i_launch_t& L = *even;
  const std::vector<Segment>& segs = plan->segs_split.empty() ? plan->segs : plan->segs_split;
  if (segs.size() != 1 || segs[0].phases.size() != 2 || segs[0].phases[0].type != PH_SEPH ||
      segs[0].phases[1].type != PH_SEPV || segs[0].cin != 3 || segs[0].cout != 3 || segs[0].phases[1].expand ||
      segs[0].phases[1].lut >= 0)
    B2I_FAIL(B2I_ERR_UNSUPPORTED, "the fused resident launch takes one separable 3-channel stencil");
  const Segment& seg = segs[0];
  const PhaseDesc& ph = seg.phases[0];
  const PhaseDesc& pv = seg.phases[1];
  const int y0 = L.y0, y1 = L.y1;
  if ((L.W & 7) || L.W < 16 || L.H < 8 || y1 - y0 < 1 || odd->y0 != y0 || odd->y1 != y1 ||
      !tensor_map_encoder())
    B2I_FAIL(B2I_ERR_UNSUPPORTED, "geometry the sweep kernel does not take");
  bool pushes = false, flags = false;
  if (!sweep_hooks_ok(L, y0, y1, true, true, &pushes, &flags) || !(L.counters && L.status))
    B2I_FAIL(B2I_ERR_UNSUPPORTED, "multi-GPU hooks the sweep kernel does not take");
  int dev = 0, sms = 0;
  B2I_CUDA(cudaGetDevice(&dev));
  B2I_TRY(device_sms(dev, &sms));
  Sweep5Params P;
  memset(&P, 0, sizeof P);
  P.W = L.W;
  P.H = L.H;
  P.y0 = y0;
  P.y1 = y1;
  P.src_y_begin = L.src.y_begin;
  P.border = ph.border;
  const int up = 8 - pv.shift;
  if (up < 0) B2I_FAIL(B2I_ERR_UNSUPPORTED, "shift beyond 8");
  P.bias = pv.bias << up;
  const int off = ph.ksize == 5 ? 0 : 1;
  for (int i = 0; i < ph.ksize; ++i) {
    P.wh[i + off] = ph.sep[i];
    P.wv[i + off] = pv.sep[i] << up;
  }
  static const int b5[5] = {1, 4, 6, 4, 1}, b3[5] = {0, 1, 2, 1, 0};
  P.binom = !memcmp(P.wh, b5, sizeof b5) || !memcmp(P.wh, b3, sizeof b3);
  P.in_stage_bytes = align_up(kS5Chunk * kS5InBytes, 128);
  P.out_buf_bytes = align_up(kS5Chunk * kS5OutBytes, 128);
  P.warp_smem = kS5Stages * P.in_stage_bytes + 2 * P.out_buf_bytes + 128;
  CUtensorMap sm0, dm0, sm1, dm1;
  auto maps = [&](const b2i_launch_t& X, CUtensorMap* sm, CUtensorMap* dm) {
    if (X.src.y_begin != L.src.y_begin || X.src.y_end != L.src.y_end) return false;
    const int src_rows = std::min(X.src.y_end, L.H) - X.src.y_begin;
    if (!encode_map(sm, X.src.ptr, X.src.pitch, (int64_t)L.W * 3, src_rows, kS5InBytes, kS5Chunk)) return false;
    const uint64_t dptr = X.dst.ptr + (uint64_t)((int64_t)(y0 - X.dst.y_begin) * X.dst.pitch);
    return encode_map(dm, dptr, X.dst.pitch, (int64_t)L.W * 3, y1 - y0, kS5OutBytes, kS5Chunk);
  };
  if (!maps(*even, &sm0, &dm0) || !maps(*odd, &sm1, &dm1))
    B2I_FAIL(B2I_ERR_UNSUPPORTED, "band buffers the tensor maps do not take");
  P.n_strips = (L.W + kS5Strip - 1) / kS5Strip;
  const int resident = sms * kS5CtasPerSm * kS5Warps;
  // With neighbours, the first and the last segment are one box (5 rows) high: those tasks run first
  // and are short, so their halo rows and flags are under way while the other tasks still work, and
  // the wait for the neighbour's rows of the pass before is over before it could hold anything up.
  P.strip_flags = (L.flags & B2I_LAUNCH_STRIP_FLAGS) ? 1 : 0;
  if (P.strip_flags && P.n_strips > B2I_MAX_STRIPS) B2I_FAIL(B2I_ERR_UNSUPPORTED, "image wider than the strip flags");
  if (pushes && passes > 1 && y1 - y0 >= 4 * kS5Chunk && !B2I_ENV("B200IMG_NO_BSEG", 0)) {
    // (every segment but the last is whole boxes: a ragged box would spill into the rows below it)
    P.bseg = kS5Chunk;
    const int mid = (y1 - y0 - 2 * kS5Chunk) / kS5Chunk * kS5Chunk;
    P.bseg_last = y1 - y0 - P.bseg - mid;
    P.seg_rows = sweep_seg_rows(P.n_strips, mid, std::max(kS5Warps, resident - 2 * P.n_strips), kS5Chunk, 5, 120, 4);
    P.n_segs = 2 + (mid + P.seg_rows - 1) / P.seg_rows;
  } else {
    P.seg_rows = sweep_seg_rows(P.n_strips, y1 - y0, resident, kS5Chunk, 5, 120, 4);
    P.n_segs = (y1 - y0 + P.seg_rows - 1) / P.seg_rows;
  }
  P.n_tasks = P.n_strips * P.n_segs;
  if (P.n_tasks + 8 > n_task_flags)
    B2I_FAIL(B2I_ERR_ARG, "%d task flags given, %d needed", n_task_flags, P.n_tasks + 8);
  sweep_hooks_fill(P, L, y0, y1, true, true, pushes);
  if (pushes) {   // odd passes push into the other buffer of the neighbours
    if (odd->push_up.ptr)
      P.push_up_ptr_odd = odd->push_up.ptr + (uint64_t)((int64_t)(y0 - odd->push_up.y_begin) * odd->push_up.pitch);
    if (odd->push_dn.ptr)
      P.push_dn_ptr_odd =
          odd->push_dn.ptr + (uint64_t)((int64_t)(y1 - L.push_rows - odd->push_dn.y_begin) * odd->push_dn.pitch);
  }
  if (P.bseg) P.n_up_tasks = P.n_dn_tasks = P.n_strips;   // the short segments hold all the pushed rows
  if (passes == 1) {   // a single pass reads what the scatter left and nobody reads its halo
    P.wait_up_flag = P.wait_dn_flag = P.signal_up_flag = P.signal_dn_flag = 0;
    P.push_up_ptr = P.push_dn_ptr = 0;
    P.push_rows = 0;
  }
  P.n_passes = passes;
  P.wait_epoch = epoch_base;
  P.signal_epoch = epoch_base + 1;
  P.task_done = task_flags;
  P.counters = task_flags + 4ull * (uint64_t)P.n_tasks;   // eight zeroed words behind the task flags
  const int grid = sweep_grid(P.n_tasks, kS5Warps, sms, kS5CtasPerSm);
  plan->info.last_tile_h = -P.seg_rows;
  plan->info.last_smem_bytes = P.warp_smem * kS5Warps;
  plan->info.last_grid = grid;
  plan->info.last_blocks_per_sm = kS5CtasPerSm;
  plan->info.last_tma_load = plan->info.last_tma_store = 1;
  plan->info.last_launches = 1;
  cudaError_t e = passes > 1 ? launch_sweep5(P, sm0, dm0, grid, (cudaStream_t)L.stream, &sm1, &dm1)
                             : launch_sweep5(P, sm0, dm0, grid, (cudaStream_t)L.stream);
  if (e != cudaSuccess) B2I_FAIL(B2I_ERR_CUDA, "fused resident launch: %s", cudaGetErrorString(e));
  return B2I_OK;
}
