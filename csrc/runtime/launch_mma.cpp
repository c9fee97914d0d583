// Launcher of the tensor-core stencil kernel (kernels/stencil_mma.cu): eligibility, tensor maps
// (128-byte-swizzled boxes), work split.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "plan_internal.hpp"

namespace b2i {

// Takes segments that are exactly one k x k stencil (or the Sobel pair) on 1- or 3-channel rows
// with a border rule other than REF. Returns 1 if it launched, 0 if the segment or the buffers do
// not qualify (the sweep or tile kernels then run it), -1 on a launch error.
int try_stencil_mma(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
                    const b2i_view_t& src, const b2i_view_t& dst, int sms, bool first, bool last) {
  (void)first;
  if (!seg.mma_ok || B2I_ENV("B200IMG_NO_MMA", 0) || (L.flags & B2I_LAUNCH_NO_TMA) || B2I_ENV("B200IMG_NO_TMA", 0))
    return 0;
  // multi-GPU hooks are not in this kernel (yet): the sweeps carry them
  if (L.push_rows || L.wait_up_flag || L.wait_dn_flag || L.signal_up_flag || L.signal_dn_flag ||
      (last && (L.done_flag || L.wait_done_mask || (L.flags & B2I_LAUNCH_DST_MCAST))))
    return 0;
  const int64_t W3 = (int64_t)L.W * seg.cin;
  if (W3 < 2 * kSmSlab || L.H < 2 * seg.mma_R + 2 || y1 - y0 < 1 || W3 > (1 << 30)) return 0;
  StencilMmaParams P;
  memset(&P, 0, sizeof P);
  P.W3 = (int32_t)W3;
  P.H = L.H;
  P.y0 = y0;
  P.y1 = y1;
  P.src_y_begin = src.y_begin;
  P.cs = seg.cin;
  P.R = seg.mma_R;
  P.border = seg.mma_border;
  P.mode = seg.mma_mode;
  P.shift = seg.mma_shift;
  P.bias = seg.mma_bias;
  memcpy(P.w, seg.mma_w, sizeof P.w);
  int sum = 0;
  bool nonneg = true;
  for (int j = 0; j < 2 * P.R + 1; ++j)
    for (int i = 0; i < 2 * P.R + 1; ++i) {
      sum += P.w[0][j][i];
      nonneg = nonneg && P.w[0][j][i] >= 0;
    }
  P.fast8 = P.mode == SM_LINEAR && P.shift == 8 && nonneg && sum == 256 && P.bias >= 0 && P.bias < 256;
  P.n_bands = (y1 - y0 + kSmRows - 1) / kSmRows;
  P.n_slabs = (int)((W3 + kSmSlab - 1) / kSmSlab);
  P.dbg = B2I_ENV("B200IMG_MMA_DBG", 0);
  if (const char* dp = getenv("B200IMG_MMA_DBGPTR")) P.dbg_ptr = strtoull(dp, nullptr, 0);
  CUtensorMap src_map, dst_map;
  const int src_rows = std::min(src.y_end, L.H) - src.y_begin;
  if (!encode_map_sw128(&src_map, src.ptr, src.pitch, W3, src_rows, kSmRows + 2 * P.R)) return 0;
  const uint64_t dptr = dst.ptr + (uint64_t)((int64_t)(y0 - dst.y_begin) * dst.pitch);
  if (!encode_map_sw128(&dst_map, dptr, dst.pitch, W3, y1 - y0, kSmRows)) return 0;
  const long long total = (long long)P.n_bands * P.n_slabs;
  const int grid = (int)std::max(1LL, std::min<long long>(total, sms));
  plan->info.last_tile_h = kSmRows;
  plan->info.last_smem_bytes = kSmStages * kSmSlabBytes + kSmVariants * 6 * kSmBBytes + 2 * kSmOutBytes + 1024;
  plan->info.last_grid = grid;
  plan->info.last_blocks_per_sm = 1;
  plan->info.last_tma_load = plan->info.last_tma_store = 1;
  plan->info.last_launches++;
  cudaError_t e = launch_stencil_mma(P, src_map, dst_map, grid, (cudaStream_t)L.stream);
  if (e != cudaSuccess) {
    b2i::fail(B2I_ERR_CUDA, __FILE__, __LINE__, "stencil_mma launch: %s", cudaGetErrorString(e));
    return -1;
  }
  return 1;
}

}  // namespace b2i
