// Tile-chain and pointwise launchers: tensor maps, shared-memory layout, tile height, grid.
#include <algorithm>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "plan_internal.hpp"

namespace b2i {

PFN_cuTensorMapEncodeTiled_v12000 tensor_map_encoder() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
  });
  return fn;
}

// 2-D byte image as a tensor of 32-bit elements: dim0 = row_bytes/4, dim1 = rows.
bool encode_map(CUtensorMap* m, uint64_t ptr, int64_t pitch, int64_t row_bytes, int rows,
                int box_w_bytes, int box_rows) {
  auto enc = tensor_map_encoder();
  if (!enc || rows <= 0) return false;
  if ((ptr & 15) || (pitch & 15) || (row_bytes & 3) || pitch <= 0) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)(row_bytes / 4), (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)pitch};
  cuuint32_t box[2] = {(cuuint32_t)(box_w_bytes / 4), (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  if (box[0] > 256 || box[1] > 256) return false;
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, reinterpret_cast<void*>(ptr), gdim, gstr, box,
             estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// 2-D byte image as a tensor of bytes with 128-byte-swizzled boxes of 128 bytes x box_rows: the
// shared-memory layout the tensor cores read K-major operands from (kernels/stencil_mma.cu). Byte
// elements: any row length clips correctly, only the base and the pitch must be 16-byte aligned.
bool encode_map_sw128(CUtensorMap* m, uint64_t ptr, int64_t pitch, int64_t row_bytes, int rows, int box_rows) {
  auto enc = tensor_map_encoder();
  if (!enc || rows <= 0 || row_bytes <= 0) return false;
  if ((ptr & 15) || (pitch & 15) || pitch <= 0 || box_rows > 256) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)pitch};
  cuuint32_t box[2] = {128u, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, reinterpret_cast<void*>(ptr), gdim, gstr, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int align_up(int v, int a) { return (v + a - 1) / a * a; }

// Shared-memory layout of a tile segment for tile height th. Every buffer starts 128-byte
// aligned behind a 128-byte gap: stencil threads read 8 bytes to the left of a strip.
void layout(const Segment& seg, int th, TileParams& P, int raw_stages) {
  const int R = seg.halo, rows_alloc = th + 2 * R + kRowGroup;
  int off = 0;
  auto place = [&](int bytes) {
    off = align_up(off, 128) + 128;
    const int o = off;
    off += bytes;
    return o;
  };
  P.th = th;
  P.halo = R;
  P.raw_rows = th + 2 * R;
  const int raw_pitch = kPlaneW * seg.cin;
  P.raw_tx_bytes = raw_pitch * P.raw_rows;
  P.raw_stage_bytes = align_up(raw_pitch * rows_alloc, 128) + 128;
  P.raw_stages = raw_stages;
  P.buf[BUF_RAW] = {place(raw_stages * P.raw_stage_bytes), raw_pitch, 0, 0};
  P.buf[BUF_A] = {seg.chan_a ? place(kPlaneW * seg.chan_a * rows_alloc) : 0, kPlaneW * seg.chan_a, 0, 0};
  P.buf[BUF_B] = {seg.chan_b ? place(kPlaneW * seg.chan_b * rows_alloc) : 0, kPlaneW * seg.chan_b, 0, 0};
  P.buf[BUF_STAGE] = {place(kTileW * seg.cout * th), kTileW * seg.cout, kApron, R};
  P.lut_off = place(kMaxLuts * 256);
  P.bar_off = place(64);
  P.smem_bytes = align_up(off, 128);
}

int device_sms(int dev, int* sms) {
  static std::atomic<int> cache[64];   // zero-initialised; launches may come from several host threads
  if (dev >= 0 && dev < 64) {
    const int v = cache[dev].load(std::memory_order_relaxed);
    if (v) {
      *sms = v;
      return B2I_OK;
    }
  }
  B2I_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
  if (dev >= 0 && dev < 64) cache[dev].store(*sms, std::memory_order_relaxed);
  return B2I_OK;
}

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

// One tile-kernel launch: rows [y0,y1) of this segment's result.
int launch_tile_segment(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
                        const b2i_view_t& src, const b2i_view_t& dst, bool first, bool last) {
  int dev = 0, sms = 0;
  B2I_CUDA(cudaGetDevice(&dev));
  B2I_TRY(device_sms(dev, &sms));
  {
    B2I_CHECK_ARG(src.ptr && dst.ptr, "null image view");
    const int need_a = std::max(0, y0 - seg.halo), need_b = std::min(L.H, y1 + seg.halo);
    B2I_CHECK_ARG(src.y_begin <= need_a && src.y_end >= need_b && dst.y_begin <= y0 && dst.y_end >= y1,
                  "views do not cover the band [%d,%d) and its halo", y0, y1);
    // Measured on 8K RGB (scripts/kbench.py): gauss5 54.7 us on the separable sweep, 55.1 on the
    // tensor-core kernel; gauss3 44.1 vs 47.6; Sobel 55.1 vs 75. So 3-channel Gaussians and Sobel go
    // to the sweeps, and the tensor-core kernel takes what they cannot: one-channel planes, widths
    // that are not multiples of 8 pixels, and every stencil that is not separable (the tile kernel
    // needs 130 us and more for those). B200IMG_PREFER_MMA=1 turns the order around.
    static const int prefer_mma = B2I_ENV("B200IMG_PREFER_MMA", 0);
    int sw = try_sweep(plan, seg, L, y0, y1, src, dst, sms, last);
    if (sw == 0 && prefer_mma) sw = try_stencil_mma(plan, seg, L, y0, y1, src, dst, sms, first, last);
    if (sw == 0) sw = try_sweep5(plan, seg, L, y0, y1, src, dst, sms, first, last);
    if (sw == 0) sw = try_sweep_sobel(plan, seg, L, y0, y1, src, dst, sms, first, last);
    if (sw == 0) sw = try_stencil_mma(plan, seg, L, y0, y1, src, dst, sms, first, last);
    if (sw < 0) return B2I_ERR_CUDA;
    if (sw > 0) return B2I_OK;
  }
  TileParams P;
  memset(&P, 0, sizeof P);
  // tile height: tallest that still lets two CTAs share an SM; smaller for small images so that
  // every SM gets tiles
  int th = L.tile_h > 0 ? L.tile_h : B2I_ENV("B200IMG_TILE_H", 0);
  int raw_stages = B2I_ENV("B200IMG_RAW_STAGES", 2);
  const int tiles_x = (L.W + kTileW - 1) / kTileW;
  if (th <= 0) {
    // Measured on 8K images (scripts/tile_h_sweep.py): the per-tile cost (barriers between
    // phases, pipeline fill) is large, so tall tiles win. Two CTAs per SM with th >= 16 beat
    // one CTA with th = 32, which beats anything with th = 8.
    th = 0;
    for (int c : {32, 24, 16}) {
      for (int st : {2, 1}) {  // without the prefetch stage a taller tile may still fit twice per SM
        layout(seg, c, P, st);
        if (P.smem_bytes <= 113 * 1024) {
          th = c;
          raw_stages = st;
          break;
        }
      }
      if (th) break;
    }
    if (!th)
      for (int c : {32, 24, 16}) {
        layout(seg, c, P);
        if (P.smem_bytes <= 227 * 1024) {
          th = c;
          break;
        }
      }
    if (!th) th = 8;
    // small images: shorter tiles so that every SM gets work
    while (th > 8 && tiles_x * ((y1 - y0 + th - 1) / th) < 2 * sms) th -= 8;
  }
  B2I_CHECK_ARG(th >= 4 && th <= kMaxTileH, "tile height %d out of range", th);
  layout(seg, th, P, raw_stages == 1 ? 1 : 2);
  if (P.smem_bytes > 227 * 1024)
    B2I_FAIL(B2I_ERR_UNSUPPORTED, "segment needs %d bytes of shared memory", P.smem_bytes);
  P.W = L.W;
  P.H = L.H;
  P.cin = seg.cin;
  P.cout = seg.cout;
  P.y0 = y0;
  P.y1 = y1;
  P.tiles_x = tiles_x;
  P.tiles_y = (y1 - y0 + th - 1) / th;
  P.n_tiles = P.tiles_x * P.tiles_y;
  P.n_phases = (int)seg.phases.size();
  P.rgb_order = plan->chain.order == B2I_ORDER_RGB;
  P.raw_fix_border = seg.raw_fix_border;
  P.debug = B2I_ENV("B200IMG_DEBUG", 0);
  for (int k = 0; k < P.n_phases; ++k) {
    PhaseDesc& ph = P.ph[k];
    ph = seg.phases[k];
    // work decomposition (see tile_chain.cu): units per row, rows or row groups, items
    const int mc = ph.margin_x > 0 ? kMarginCols : 0, rows = th + 2 * ph.margin;
    int units, lines;
    switch (ph.type) {
      case PH_GRAY_REF:
      case PH_GRAY_CV:
        units = (kTileW + 2 * mc) >> 3;  // 8-pixel items
        lines = rows;
        break;
      case PH_COPY:
      case PH_EXPAND:
        units = (kTileW + 2 * mc) >> 2;  // 4-pixel items
        lines = rows;
        break;
      case PH_SEPH:
        units = ((kTileW + 2 * mc) * ph.cin) >> 3;  // 8-byte strips
        lines = rows;
        break;
      case PH_SOBEL:
        units = ((kTileW + 2 * mc) * ph.cin) >> 3;
        lines = (rows + 1) / 2;  // row groups of 2
        break;
      default:  // PH_STENCIL, PH_SEPV: 8-byte strips x row groups of kRowGroup
        units = ((kTileW + 2 * mc) * ph.cin) >> 3;
        lines = (rows + kRowGroup - 1) / kRowGroup;
        break;
    }
    ph.units = units;
    ph.units_magic = ((1 << 20) + units - 1) / units;
    ph.n_items = units * lines;
    ph.drow = kTileThreads / units;
    ph.dcol = kTileThreads - ph.drow * units;
  }
  memcpy(P.luts, seg.luts, sizeof P.luts);
  // views
  B2I_CHECK_ARG(src.ptr && dst.ptr, "null image view");
  const int need_a = std::max(0, y0 - seg.halo), need_b = std::min(L.H, y1 + seg.halo);
  B2I_CHECK_ARG(src.y_begin <= need_a && src.y_end >= need_b,
                "source view rows [%d,%d) do not cover the rows [%d,%d) this band reads",
                src.y_begin, src.y_end, need_a, need_b);
  B2I_CHECK_ARG(dst.y_begin <= y0 && dst.y_end >= y1, "destination view rows [%d,%d) do not "
                "cover the band [%d,%d)", dst.y_begin, dst.y_end, y0, y1);
  P.src_ptr = src.ptr;
  P.src_pitch = src.pitch;
  P.src_y_begin = src.y_begin;
  P.src_y_end = std::min(src.y_end, L.H);
  P.dst_ptr = dst.ptr + (uint64_t)((int64_t)(y0 - dst.y_begin) * dst.pitch);
  P.dst_pitch = dst.pitch;
  P.dst_y_begin = y0;
  P.dst_y_end = y1;
  const bool no_tma = (L.flags & B2I_LAUNCH_NO_TMA) || B2I_ENV("B200IMG_NO_TMA", 0);
  CUtensorMap src_map, dst_map, up_map, dn_map;
  memset(&src_map, 0, sizeof src_map);
  memset(&dst_map, 0, sizeof dst_map);
  memset(&up_map, 0, sizeof up_map);
  memset(&dn_map, 0, sizeof dn_map);
  const bool no_tma_load = no_tma || B2I_ENV("B200IMG_NO_TMA_LOAD", 0);
  const bool no_tma_store = no_tma || B2I_ENV("B200IMG_NO_TMA_STORE", 0) || (last && (L.flags & B2I_LAUNCH_DST_MCAST));
  P.use_tma_load = !no_tma_load && encode_map(&src_map, P.src_ptr, P.src_pitch, (int64_t)L.W * seg.cin,
                                         P.src_y_end - P.src_y_begin, kPlaneW * seg.cin, P.raw_rows);
  P.use_tma_store = !no_tma_store && encode_map(&dst_map, P.dst_ptr, P.dst_pitch, (int64_t)L.W * seg.cout,
                                          y1 - y0, kTileW * seg.cout, th);
  // multi-GPU hooks: waits guard the first segment's ghost-row reads, pushes and signals ride
  // on the last segment's stores
  if (first) {
    P.wait_up_flag = L.wait_up_flag;
    P.wait_dn_flag = L.wait_dn_flag;
    P.wait_epoch = L.wait_epoch;
  }
  if (last) {
    P.signal_up_flag = L.signal_up_flag;
    P.signal_dn_flag = L.signal_dn_flag;
    P.signal_epoch = L.signal_epoch;
    if (L.push_rows > 0 && (L.push_up.ptr || L.push_dn.ptr)) {
      B2I_CHECK_ARG(L.push_rows <= y1 - y0 && L.push_rows <= kGhostRows,
                    "push_rows %d exceeds the band height or the %d ghost rows", L.push_rows,
                    kGhostRows);
      P.push_rows = L.push_rows;
      P.push_pitch = L.push_up.ptr ? L.push_up.pitch : L.push_dn.pitch;
      B2I_CHECK_ARG(!(L.push_up.ptr && L.push_dn.ptr) || L.push_up.pitch == L.push_dn.pitch,
                    "push views must share one pitch");
      if (L.push_up.ptr) {  // view: rows [y0, y0+push_rows) of the result, ptr at row y_begin
        B2I_CHECK_ARG(L.push_up.y_begin <= y0 && L.push_up.y_end >= y0 + L.push_rows,
                      "push_up view does not cover rows [%d,%d)", y0, y0 + L.push_rows);
        P.push_up_ptr = L.push_up.ptr + (uint64_t)((int64_t)(y0 - L.push_up.y_begin) * L.push_up.pitch);
        P.push_up_y_begin = y0;
      }
      if (L.push_dn.ptr) {
        const int a = y1 - L.push_rows;
        B2I_CHECK_ARG(L.push_dn.y_begin <= a && L.push_dn.y_end >= y1,
                      "push_dn view does not cover rows [%d,%d)", a, y1);
        P.push_dn_ptr = L.push_dn.ptr + (uint64_t)((int64_t)(a - L.push_dn.y_begin) * L.push_dn.pitch);
        P.push_dn_y_begin = a;
      }
      if (P.use_tma_store) {
        bool ok = true;
        if (P.push_up_ptr)
          ok = ok && encode_map(&up_map, P.push_up_ptr, P.push_pitch, (int64_t)L.W * seg.cout,
                                P.push_rows, kTileW * seg.cout, 1);
        if (P.push_dn_ptr)
          ok = ok && encode_map(&dn_map, P.push_dn_ptr, P.push_pitch, (int64_t)L.W * seg.cout,
                                P.push_rows, kTileW * seg.cout, 1);
        if (!ok) {  // fall back to plain stores for everything
          P.use_tma_store = 0;
        }
      }
    }
  }
  if (P.wait_up_flag || P.wait_dn_flag || P.signal_up_flag || P.signal_dn_flag) {
    B2I_CHECK_ARG(L.counters && L.status, "flag use needs the counters and status scratch words");
    P.counters = L.counters;
    P.status = L.status;
  }
  P.timeout_ns = L.timeout_ns ? L.timeout_ns : 2000000000ull;
  // persistent grid: resident CTAs per SM x SMs, never more than there are tiles
  static std::mutex mu;
  static int occ_cache_smem[8] = {0}, occ_cache_bps[8] = {0};
  int bps = 0;
  {
    std::lock_guard<std::mutex> g(mu);
    for (int i = 0; i < 8; ++i)
      if (occ_cache_smem[i] == P.smem_bytes + dev * (1 << 20)) bps = occ_cache_bps[i];
    if (!bps) {
      B2I_CUDA(tile_chain_occupancy(P.smem_bytes, &bps));
      static int next = 0;
      occ_cache_smem[next & 7] = P.smem_bytes + dev * (1 << 20);
      occ_cache_bps[next & 7] = bps;
      ++next;
    }
  }
  if (bps < 1) B2I_FAIL(B2I_ERR_UNSUPPORTED, "tile kernel does not fit on an SM (%d B smem)", P.smem_bytes);
  const int grid = std::max(1, std::min(P.n_tiles, sms * bps));
  plan->info.last_tile_h = th;
  plan->info.last_smem_bytes = P.smem_bytes;
  plan->info.last_grid = grid;
  plan->info.last_blocks_per_sm = bps;
  plan->info.last_tma_load = P.use_tma_load;
  plan->info.last_tma_store = P.use_tma_store;
  plan->info.last_launches++;
  B2I_CUDA(launch_tile_chain(P, src_map, dst_map, up_map, dn_map, grid, (cudaStream_t)L.stream));
  return B2I_OK;
}

int launch_pointwise_segment(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0,
                             int y1, const b2i_view_t& src, const b2i_view_t& dst) {
  int dev = 0, sms = 0;
  B2I_CUDA(cudaGetDevice(&dev));
  B2I_TRY(device_sms(dev, &sms));
  B2I_CHECK_ARG(src.ptr && dst.ptr, "null image view");
  B2I_CHECK_ARG(src.y_begin <= y0 && src.y_end >= y1 && dst.y_begin <= y0 && dst.y_end >= y1,
                "views do not cover the band [%d,%d)", y0, y1);
  PointwiseParams P;
  memset(&P, 0, sizeof P);
  P.W = L.W;
  P.cin = seg.cin;
  P.cout = seg.cout;
  P.rows = y1 - y0;
  P.gray = seg.pw_gray;
  P.expand = seg.pw_expand;
  P.rgb_order = plan->chain.order == B2I_ORDER_RGB;
  auto lut_kind = [](int has, const uint8_t* lut) {  // 0 none, 1 table, 2 the table is 255 - v
    if (!has) return 0;
    bool ident = true, inv = true;
    for (int v = 0; v < 256; ++v) {
      ident = ident && lut[v] == v;
      inv = inv && lut[v] == 255 - v;
    }
    return ident ? 0 : inv ? 2 : 1;
  };
  P.has_lut0 = lut_kind(seg.pw_has_lut0, seg.pw_lut0);
  P.has_lut1 = lut_kind(seg.pw_has_lut1, seg.pw_lut1);
  memcpy(P.lut0, seg.pw_lut0, 256);
  memcpy(P.lut1, seg.pw_lut1, 256);
  P.src_pitch = src.pitch;
  P.dst_pitch = dst.pitch;
  P.src_ptr = src.ptr + (uint64_t)((int64_t)(y0 - src.y_begin) * src.pitch);
  P.dst_ptr = dst.ptr + (uint64_t)((int64_t)(y0 - dst.y_begin) * dst.pitch);
  P.vec_ok = !((P.src_ptr | P.dst_ptr | (uint64_t)P.src_pitch | (uint64_t)P.dst_pitch) & 15) &&
             !B2I_ENV("B200IMG_NO_VEC", 0);
  const long long units = (long long)((L.W + 511) / 512) * P.rows;  // one warp per 512-pixel run
  const long long want = (units + 7) / 8;
  const int grid = (int)std::max(1LL, std::min(want, (long long)sms * 8));
  plan->info.last_grid = grid;
  plan->info.last_tile_h = 0;
  plan->info.last_smem_bytes = 512;
  plan->info.last_tma_load = plan->info.last_tma_store = 0;
  plan->info.last_launches++;
  B2I_CUDA(launch_pointwise(P, grid, (cudaStream_t)L.stream));
  return B2I_OK;
}

}  // namespace b2i
