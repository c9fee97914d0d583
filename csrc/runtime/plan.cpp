// Chain -> plan compiler and the band-launch entry point.
//
// The reference has no such layer: its pipeline is three literal kernel launches
// (kernel.cu:192-195). Here a chain is compiled once into segments; each segment is one launch
// of either the streaming pointwise kernel (kernels/pointwise.cu) or the tile-chain kernel
// (kernels/tile_chain.cu). Compilation folds every run of u8->u8 pointwise ops into one LUT,
// attaches LUTs and the gray->BGR expand to the neighbouring stencil, lowers stencil weights to
// packed dp4a operands and lays out shared memory for a tile height chosen per launch.
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "../core/error.hpp"
#include "../kernels/sweep3.hpp"
#include "../kernels/sweep5.hpp"
#include "../kernels/sweep_sobel.hpp"
#include "../kernels/tile_plan.hpp"

namespace b2i {
cudaError_t launch_tile_chain(const TileParams& P, const CUtensorMap& src_map,
                              const CUtensorMap& dst_map, const CUtensorMap& push_up_map,
                              const CUtensorMap& push_dn_map, int grid, cudaStream_t stream);
cudaError_t tile_chain_occupancy(int smem_bytes, int* blocks_per_sm);
cudaError_t launch_pointwise(const PointwiseParams& P, int grid, cudaStream_t stream);
cudaError_t launch_sweep5(const Sweep5Params& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                          int grid, cudaStream_t stream);
cudaError_t launch_sweep_sobel(const SweepSobelParams& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                               int grid, cudaStream_t stream);
cudaError_t launch_sweep3(const SweepParams& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                          int grid, cudaStream_t stream);
int run_naive(const b2i_chain_t* chain, const uint8_t* src, int W, int H, int C, size_t src_pitch,
              uint8_t* dst, size_t dst_pitch, uint8_t* tmp0, uint8_t* tmp1, uint8_t* lut_scratch,
              cudaStream_t stream);
}  // namespace b2i

#define B2I_CUDA(expr)                                                                  \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess)                                                              \
      B2I_FAIL(e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver             \
                   ? B2I_ERR_NO_DEVICE                                                  \
                   : B2I_ERR_CUDA,                                                      \
               "%s: %s", #expr, cudaGetErrorString(e_));                                \
  } while (0)

using namespace b2i;

namespace {

int env_int(const char* name, int dflt);

struct Segment {
  bool pointwise = false;
  int cin = 0, cout = 0, halo = 0;
  // tile segments
  std::vector<PhaseDesc> phases;
  int n_luts = 0;
  uint8_t luts[kMaxLuts][256];
  int raw_fix_border = -1;
  int chan_a = 0, chan_b = 0;  // widest plane each ping-pong buffer has to hold
  // the first phase's LUT as clamp((a v + b) >> s), if it is one (found once, at plan creation)
  int lut0_affine = 0;
  int32_t aff_a = 0, aff_b = 0, aff_s = 0;
  // pointwise segments
  int pw_gray = 0, pw_expand = 0, pw_has_lut0 = 0, pw_has_lut1 = 0;
  uint8_t pw_lut0[256], pw_lut1[256];
};

// A chain element after LUT folding.
struct Elem {
  int kind;  // b2i_opkind
  const b2i_op_t* op;
  uint8_t lut[256];
};

void compose_lut(uint8_t* first_then, const uint8_t* second) {  // first_then := second o first
  for (int v = 0; v < 256; ++v) first_then[v] = second[first_then[v]];
}

uint32_t pack_w(int a, int b, int c, int d) {
  return (uint32_t)(uint8_t)(int8_t)a | ((uint32_t)(uint8_t)(int8_t)b << 8) |
         ((uint32_t)(uint8_t)(int8_t)c << 16) | ((uint32_t)(uint8_t)(int8_t)d << 24);
}

// dp4a operand layout per (channels, ksize); must match DpOff/DpCount in tile_chain.cu.
void pack_rows(int C, int K, const int* w /*K*K*/, uint32_t out[5][3]) {
  memset(out, 0, sizeof(uint32_t) * 15);
  for (int j = 0; j < K; ++j) {
    const int* r = w + j * K;
    if (C == 1 && K == 3) {
      out[j][0] = pack_w(r[0], r[1], r[2], 0);           // window at -1
    } else if (C == 1 && K == 5) {
      out[j][0] = pack_w(r[0], r[1], r[2], r[3]);        // window at -2
      out[j][1] = pack_w(r[4], 0, 0, 0);                 // window at +2
    } else if (C == 3 && K == 3) {
      out[j][0] = pack_w(r[0], 0, 0, r[1]);              // window at -3: taps -3, 0
      out[j][1] = pack_w(0, 0, 0, r[2]);                 // window at  0: tap +3
    } else {
      out[j][0] = pack_w(r[0], 0, 0, r[1]);              // window at -6: taps -6, -3
      out[j][1] = pack_w(r[2], 0, 0, r[3]);              // window at  0: taps 0, +3
      out[j][2] = pack_w(0, 0, 0, r[4]);                 // window at +3: tap +6
    }
  }
}

// w (k x k, row-major) == outer(wv, wh) with non-negative integers, sum(wv)*sum(wh) == 1<<shift:
// the kernel can run as two 1-D passes on 16-bit lanes (tile_chain.cu, PH_SEPH / PH_SEPV).
bool separable(const b2i_op_t& op, int* wv, int* wh) {
  const int k = op.ksize;
  if (op.kind != B2I_OP_STENCIL || op.border == B2I_BORDER_REF || op.shift > 8) return false;
  int j0 = -1, i0 = -1;
  for (int j = 0; j < k && j0 < 0; ++j)
    for (int i = 0; i < k; ++i)
      if (op.weights[j * k + i] != 0) {
        j0 = j;
        i0 = i;
        break;
      }
  if (j0 < 0) return false;
  auto gcd = [](int a, int b) {
    a = a < 0 ? -a : a;
    b = b < 0 ? -b : b;
    while (b) {
      int t = a % b;
      a = b;
      b = t;
    }
    return a;
  };
  int g = 0;
  for (int i = 0; i < k; ++i) g = gcd(g, op.weights[j0 * k + i]);
  if (g == 0) return false;
  int sv = 0, sh = 0;
  for (int i = 0; i < k; ++i) {
    wh[i] = op.weights[j0 * k + i] / g;
    if (wh[i] < 0) return false;
    sh += wh[i];
  }
  if (wh[i0] == 0) return false;
  for (int j = 0; j < k; ++j) {
    if (op.weights[j * k + i0] % wh[i0]) return false;
    wv[j] = op.weights[j * k + i0] / wh[i0];
    if (wv[j] < 0) return false;
    sv += wv[j];
  }
  for (int j = 0; j < k; ++j)
    for (int i = 0; i < k; ++i)
      if (wv[j] * wh[i] != op.weights[j * k + i]) return false;
  if (sh > 257 || sv * sh != (1 << op.shift) || op.bias < 0 || op.bias >= (1 << op.shift)) return false;
  return true;
}

}  // namespace

struct b2i_plan {
  b2i_chain_t chain;
  int cin = 0, cout = 0, halo = 0;
  std::vector<Segment> segs;
  // The same chain cut so that every stencil the sweep kernels take stands alone (empty if that is
  // the same cut). Two sweep launches beat one fused tile launch (8K gauss5,sobel: 290 us fused,
  // 89 + 62 us split), but only where the sweeps can run: b2i_cuda_run picks per launch.
  std::vector<Segment> segs_split;
  b2i_plan_info_t info;
};

namespace {

// Builds the tile-kernel phases of one segment from its elements.
int build_tile_segment(const std::vector<Elem>& el, int cin, Segment& seg) {
  seg.pointwise = false;
  seg.cin = cin;
  seg.phases.clear();
  seg.n_luts = 0;
  int c = cin;
  auto add_lut = [&](PhaseDesc& ph, const uint8_t* lut) -> int {
    if (ph.lut >= 0) {
      compose_lut(seg.luts[ph.lut], lut);
      return B2I_OK;
    }
    if (seg.n_luts >= kMaxLuts) return B2I_ERR_UNSUPPORTED;
    memcpy(seg.luts[seg.n_luts], lut, 256);
    ph.lut = seg.n_luts++;
    return B2I_OK;
  };
  for (size_t i = 0; i < el.size(); ++i) {
    const Elem& e = el[i];
    PhaseDesc ph;
    memset(&ph, 0, sizeof ph);
    ph.lut = -1;
    ph.fix_border = -1;
    ph.cin = c;
    switch (e.kind) {
      case B2I_OP_LUT:
        if (!seg.phases.empty()) {  // rides on the previous phase's result
          if (add_lut(seg.phases.back(), e.lut) != B2I_OK) return B2I_ERR_UNSUPPORTED;
          continue;
        }
        ph.type = PH_COPY;
        ph.cout = c;
        if (add_lut(ph, e.lut) != B2I_OK) return B2I_ERR_UNSUPPORTED;
        break;
      case B2I_OP_GRAY_REF:
      case B2I_OP_GRAY_CV:
        ph.type = e.kind == B2I_OP_GRAY_REF ? PH_GRAY_REF : PH_GRAY_CV;
        ph.cout = c = 1;
        break;
      case B2I_OP_EXPAND:
        if (!seg.phases.empty() &&
            (seg.phases.back().type == PH_STENCIL || seg.phases.back().type == PH_SOBEL ||
             seg.phases.back().type == PH_SEPV) &&
            !seg.phases.back().expand) {
          seg.phases.back().expand = 1;  // kernel.cu:210 fused into the stencil's store
          seg.phases.back().cout = c = 3;
          continue;
        }
        ph.type = PH_EXPAND;
        ph.cout = c = 3;
        break;
      default: {  // stencil / sobel
        const b2i_op_t& op = *e.op;
        int wv[5] = {0, 0, 0, 0, 0}, wh[5] = {0, 0, 0, 0, 0};
        if (!env_int("B200IMG_NO_SEP", 0) && separable(op, wv, wh)) {
          ph.type = PH_SEPH;  // horizontal pass: bytes -> 16-bit lanes
          ph.cout = c;
          ph.border = op.border;
          ph.ksize = op.ksize;
          for (int q = 0; q < 5; ++q) ph.sep[q] = wh[q];
          seg.phases.push_back(ph);
          PhaseDesc pv;
          memset(&pv, 0, sizeof pv);
          pv.lut = -1;
          pv.fix_border = -1;
          pv.type = PH_SEPV;  // vertical pass: lanes -> bytes, rounded and shifted
          pv.cin = pv.cout = c;
          pv.border = op.border;
          pv.ksize = op.ksize;
          pv.shift = op.shift;
          pv.bias = op.bias;
          for (int q = 0; q < 5; ++q) pv.sep[q] = wv[q];
          seg.phases.push_back(pv);
          continue;
        }
        ph.type = op.kind == B2I_OP_SOBEL ? PH_SOBEL : PH_STENCIL;
        ph.cout = c;
        ph.border = op.border;
        ph.ksize = op.kind == B2I_OP_SOBEL ? 3 : op.ksize;
        ph.shift = op.shift;
        ph.bias = op.bias;
        int w[25];
        if (op.kind == B2I_OP_SOBEL) {
          const int gx[9] = {-1, 0, 1, -2, 0, 2, -1, 0, 1}, gy[9] = {-1, -2, -1, 0, 0, 0, 1, 2, 1};
          uint32_t tmp[5][3];
          pack_rows(c, 3, gx, ph.wpk);
          pack_rows(c, 3, gy, tmp);
          memcpy(ph.wpk2, tmp, sizeof ph.wpk2);
        } else {
          for (int q = 0; q < ph.ksize * ph.ksize; ++q) w[q] = op.weights[q];
          pack_rows(c, ph.ksize, w, ph.wpk);
        }
        break;
      }
    }
    seg.phases.push_back(ph);
  }
  if (seg.phases.empty() || (int)seg.phases.size() > kMaxPhases) return B2I_ERR_UNSUPPORTED;
  seg.cout = c;
  // margins, walking backwards: a phase must produce what the later stencils still consume
  int margin = 0, margin_x = 0;
  for (int k = (int)seg.phases.size() - 1; k >= 0; --k) {
    PhaseDesc& ph = seg.phases[k];
    ph.margin = margin;
    ph.margin_x = margin_x;
    if (ph.type == PH_STENCIL || ph.type == PH_SOBEL) {
      margin += ph.ksize / 2;
      margin_x += ph.ksize / 2;
    }
    if (ph.type == PH_SEPV) margin += ph.ksize / 2;    // the vertical pass widens the rows needed
    if (ph.type == PH_SEPH) margin_x += ph.ksize / 2;  // the horizontal pass the columns
  }
  seg.halo = std::max(margin, margin_x);
  if (seg.halo > kMaxSegHalo) return B2I_ERR_UNSUPPORTED;
  // buffers: RAW -> A -> B -> A ... -> STAGE
  seg.chan_a = seg.chan_b = 0;
  int src = BUF_RAW;
  for (size_t k = 0; k < seg.phases.size(); ++k) {
    PhaseDesc& ph = seg.phases[k];
    ph.src = src;
    const bool last = k + 1 == seg.phases.size();
    ph.dst = last ? BUF_STAGE : (src == BUF_A ? BUF_B : BUF_A);
    const int bpp = ph.type == PH_SEPH ? 2 * ph.cout : ph.cout;  // 16-bit lanes take two bytes per byte
    if (ph.dst == BUF_A) seg.chan_a = std::max(seg.chan_a, bpp);
    if (ph.dst == BUF_B) seg.chan_b = std::max(seg.chan_b, bpp);
    src = ph.dst;
    // the border rule a following stencil needs painted onto this phase's result
    if (!last) {
      const PhaseDesc& nx = seg.phases[k + 1];
      if ((nx.type == PH_STENCIL || nx.type == PH_SOBEL || nx.type == PH_SEPH) && nx.border != B2I_BORDER_REF)
        ph.fix_border = nx.border;
    }
  }
  const PhaseDesc& p0 = seg.phases[0];
  seg.raw_fix_border = -1;
  if ((p0.type == PH_STENCIL || p0.type == PH_SOBEL || p0.type == PH_SEPH) &&
      (p0.border == B2I_BORDER_REFLECT101 || p0.border == B2I_BORDER_REPLICATE))
    seg.raw_fix_border = p0.border;  // ZERO is what TMA's out-of-bounds fill already gives
  return B2I_OK;
}

// [lut] [gray] [lut] [expand] [lut]: the shapes the streaming kernel takes.
bool build_pointwise_segment(const std::vector<Elem>& el, int cin, Segment& seg) {
  seg.pointwise = true;
  seg.cin = cin;
  seg.halo = 0;
  seg.pw_gray = seg.pw_expand = seg.pw_has_lut0 = seg.pw_has_lut1 = 0;
  int c = cin;
  for (const Elem& e : el) {
    switch (e.kind) {
      case B2I_OP_LUT:
        if (!seg.pw_gray && !seg.pw_expand) {  // before any channel change: lut0 (or lut1 if no gray follows)
          if (seg.pw_has_lut0) compose_lut(seg.pw_lut0, e.lut);
          else memcpy(seg.pw_lut0, e.lut, 256), seg.pw_has_lut0 = 1;
        } else {  // after gray and/or expand: per byte, commutes with the replication
          if (seg.pw_has_lut1) compose_lut(seg.pw_lut1, e.lut);
          else memcpy(seg.pw_lut1, e.lut, 256), seg.pw_has_lut1 = 1;
        }
        break;
      case B2I_OP_GRAY_REF:
      case B2I_OP_GRAY_CV:
        if (seg.pw_gray || seg.pw_expand || c != 3) return false;
        seg.pw_gray = e.kind == B2I_OP_GRAY_REF ? 1 : 2;
        c = 1;
        break;
      case B2I_OP_EXPAND:
        if (seg.pw_expand || c != 1) return false;
        seg.pw_expand = 1;
        c = 3;
        break;
      default:
        return false;
    }
  }
  seg.cout = c;
  return true;
}

int align_up(int v, int a);
void layout(const Segment& seg, int th, TileParams& P, int raw_stages);

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

int align_up(int v, int a) { return (v + a - 1) / a * a; }

// Shared-memory layout of a tile segment for tile height th. Every buffer starts 128-byte
// aligned behind a 128-byte gap: stencil threads read 8 bytes to the left of a strip.
void layout(const Segment& seg, int th, TileParams& P, int raw_stages = 2) {
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

// A segment the sweep kernels sweep5.cu / sweep_sobel.cu take (given a fitting image).
bool sweep_shape(const Segment& seg) {
  if (seg.pointwise || seg.cin != 3 || seg.cout != 3) return false;
  if (seg.phases.size() == 2)
    return seg.phases[0].type == PH_SEPH && seg.phases[1].type == PH_SEPV && !seg.phases[1].expand;
  return seg.phases.size() == 1 && seg.phases[0].type == PH_SOBEL && !seg.phases[0].expand &&
         seg.phases[0].border != B2I_BORDER_REF;
}

// Is the table clamp((a v + b) >> s) for small integers? Contrast stretches are (kernel.cu:56 with
// factor 3.5 is clamp((7v - 640) >> 1), kern.cpp:74 is clamp(3v - 256)); the sweep kernel then
// computes them instead of looking them up. Exhaustive over all 256 entries, so exact.
bool affine_lut(const uint8_t* lut, int32_t* a_out, int32_t* b_out, int32_t* s_out) {
  int v1 = -1, v2 = -1;  // the unsaturated span gives the slope; only a few multipliers per shift are tried
  for (int v = 0; v < 256; ++v)
    if (lut[v] != 0 && lut[v] != 255) {
      if (v1 < 0) v1 = v;
      v2 = v;
    }
  if (v1 < 0 || v2 == v1) return false;
  const double slope = ((double)lut[v2] - lut[v1]) / (v2 - v1);
  for (int s = 0; s <= 4; ++s)
    for (int a = (int)std::floor(slope * (1 << s)) - 1; a <= (int)std::ceil(slope * (1 << s)) + 1; ++a) {
      if (a == 0 || a < -2048 || a > 2048) continue;
      // b from the unsaturated entries: (a v + b) >> s == lut[v]  <=>  lut[v] << s <= a v + b < (lut[v] + 1) << s
      long long lo = -(1LL << 40), hi = 1LL << 40;
      bool any = false;
      for (int v = 0; v < 256; ++v) {
        if (lut[v] == 0 || lut[v] == 255) continue;
        any = true;
        lo = std::max(lo, ((long long)lut[v] << s) - (long long)a * v);
        hi = std::min(hi, (((long long)lut[v] + 1) << s) - 1 - (long long)a * v);
      }
      if (!any || lo > hi) continue;
      for (long long b = lo; b <= std::min(hi, lo + 4); ++b) {
        bool ok = true;
        for (int v = 0; v < 256 && ok; ++v) {
          long long t = ((long long)a * v + b) >> s;
          t = t < 0 ? 0 : t > 255 ? 255 : t;
          ok = t == lut[v];
        }
        if (ok) {
          *a_out = a;
          *b_out = (int32_t)b;
          *s_out = s;
          return true;
        }
      }
    }
  return false;
}

// Greedy segmentation: extend the segment while the tile kernel can still absorb it. With
// `split`, a segment the sweep kernels take is closed instead of being fused with what follows.
int segment_chain(const std::vector<Elem>& el, int in_channels, bool split, std::vector<Segment>& out) {
  out.clear();
  size_t i = 0;
  int c = in_channels;
  while (i < el.size()) {
    Segment best;
    size_t best_end = i;
    bool have = false;
    for (size_t end = i + 1; end <= el.size(); ++end) {
      std::vector<Elem> sub(el.begin() + i, el.begin() + end);
      bool stencil = false;
      for (const Elem& e : sub) stencil = stencil || e.kind == B2I_OP_STENCIL || e.kind == B2I_OP_SOBEL;
      Segment seg;
      bool ok = false;
      if (!stencil) ok = build_pointwise_segment(sub, c, seg);
      if (!ok) ok = build_tile_segment(sub, c, seg) == B2I_OK;
      if (ok) {
        TileParams P;  // also has to fit in shared memory at the smallest tile height
        if (!seg.pointwise) {
          layout(seg, 8, P);
          ok = P.smem_bytes <= 227 * 1024;
        }
      }
      if (!ok) break;
      if (split && have && sweep_shape(best) && !sweep_shape(seg)) break;
      best = seg;
      best_end = end;
      have = true;
    }
    if (!have) return B2I_ERR_UNSUPPORTED;
    c = best.cout;
    if (!best.pointwise && !best.phases.empty() && best.phases[0].lut >= 0)
    {
      best.lut0_affine = affine_lut(best.luts[best.phases[0].lut], &best.aff_a, &best.aff_b, &best.aff_s);
      if (env_int("B200IMG_DEBUG", 0) & 2)
        fprintf(stderr, "b200img: first LUT of segment %zu %s clamp((%d v + %d) >> %d)\n", out.size(),
                best.lut0_affine ? "is" : "is not", best.aff_a, best.aff_b, best.aff_s);
    }
    out.push_back(best);
    i = best_end;
  }
  return B2I_OK;
}

int device_sms(int dev, int* sms) {
  static int cache[64] = {0};
  if (dev >= 0 && dev < 64 && cache[dev]) {
    *sms = cache[dev];
    return B2I_OK;
  }
  B2I_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
  if (dev >= 0 && dev < 64) cache[dev] = *sms;
  return B2I_OK;
}

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

// Completion flags of ROOT mode, fused into the end of a sweep kernel (device_utils.cuh:
// sweep_finish). Tells b2i_cuda_run that no separate wait kernel is needed.
thread_local bool g_wait_done_fused = false;
template <class Params>
void sweep_finish_fill(Params& P, const b2i_launch_t& L) {
  P.done_flag = L.done_flag;
  P.wait_done_flags = L.wait_done_flags;
  P.wait_done_mask = L.wait_done_flags ? L.wait_done_mask : 0;
  P.signal_epoch = L.signal_epoch;
  P.counters = L.counters;
  P.status = L.status;
  P.timeout_ns = L.timeout_ns ? L.timeout_ns : 2000000000ull;
  g_wait_done_fused = true;
}

int sweep_seg_rows(int n_strips, int rows, int resident, int chunk, int lo, int hi, int halo_rows);

// The warp-sweep kernel (kernels/sweep3.cu) takes segments of the reference's own shape:
// gray (+LUT) -> one 3x3 stencil on the gray plane (+LUT, +expand). Returns 1 if it launched,
// 0 if the segment or the buffers do not qualify (the tile-chain kernel then runs it).
int try_sweep(b2i_plan* plan, const Segment& seg, const b2i_launch_t& L, int y0, int y1,
              const b2i_view_t& src, const b2i_view_t& dst, int sms, bool last) {
  if (seg.phases.size() != 2 || env_int("B200IMG_NO_SWEEP", 0) || (L.flags & B2I_LAUNCH_NO_TMA) ||
      env_int("B200IMG_NO_TMA", 0))
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
  if ((L.W & 7) || L.W < 16 || y1 - y0 < 1) return 0;
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
  if (g.lut >= 0 && seg.lut0_affine && !env_int("B200IMG_NO_AFFINE_LUT", 0)) {
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
  if (!encode_map(&src_map, src.ptr, src.pitch, (int64_t)L.W * 3, src_rows, kSweepInBytes, kSweepChunk)) return 0;
  const uint64_t dptr = dst.ptr + (uint64_t)((int64_t)(y0 - dst.y_begin) * dst.pitch);
  const bool plain = last && (L.flags & B2I_LAUNCH_DST_MCAST);
  if (plain) {  // NVLS multicast image: 16-byte stores from the staging buffer, no tensor map on that address
    if (((dptr | (uint64_t)dst.pitch) & 15) || (((int64_t)L.W * seg.cout) & 15)) return 0;
    memset(&dst_map, 0, sizeof dst_map);
    P.plain_dst_ptr = dptr;
    P.plain_dst_pitch = dst.pitch;
  } else if (!encode_map(&dst_map, dptr, dst.pitch, (int64_t)L.W * seg.cout, y1 - y0, out_row, kSweepChunk)) {
    return 0;
  }
  // segment height: a multiple of the box height, chosen so that the tasks fill the resident
  // warps evenly and the two halo rows per segment stay cheap
  P.n_strips = (L.W + kSweepStrip - 1) / kSweepStrip;
  const int resident = sms * kSweepCtasPerSm * kSweepWarps;
  P.seg_rows = sweep_seg_rows(P.n_strips, y1 - y0, resident, kSweepChunk, 6, 96, 2);
  P.n_tasks = P.n_strips * ((y1 - y0 + P.seg_rows - 1) / P.seg_rows);
  const int grid = std::max(1, std::min((P.n_tasks + kSweepWarps - 1) / kSweepWarps, sms * kSweepCtasPerSm));
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
bool sweep_hooks_ok(const b2i_launch_t& L, int y0, int y1, bool first, bool last, bool* pushes, bool* flags) {
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
void sweep_hooks_fill(Params& P, const b2i_launch_t& L, int y0, int y1, bool first, bool last, bool pushes) {
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
int sweep_seg_rows(int n_strips, int rows, int resident, int chunk, int lo, int hi, int halo_rows) {
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
  if (seg.phases.size() != 2 || env_int("B200IMG_NO_SWEEP", 0) || env_int("B200IMG_NO_SWEEP5", 0) ||
      (L.flags & B2I_LAUNCH_NO_TMA) || env_int("B200IMG_NO_TMA", 0))
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
  P.seg_rows = sweep_seg_rows(P.n_strips, y1 - y0, resident, kS5Chunk, 5, 120, 4);
  P.n_segs = (y1 - y0 + P.seg_rows - 1) / P.seg_rows;
  P.n_tasks = P.n_strips * P.n_segs;
  if (pushes || flags) sweep_hooks_fill(P, L, y0, y1, first, last, pushes);
  const int grid = std::max(1, std::min((P.n_tasks + kS5Warps - 1) / kS5Warps, sms * kS5CtasPerSm));
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
  if (seg.phases.size() != 1 || env_int("B200IMG_NO_SWEEP", 0) || env_int("B200IMG_NO_SWEEP_SOBEL", 0) ||
      (L.flags & B2I_LAUNCH_NO_TMA) || env_int("B200IMG_NO_TMA", 0))
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
  P.seg_rows = sweep_seg_rows(P.n_strips, y1 - y0, resident, kSSChunk, 4, 120, 2);
  P.n_segs = (y1 - y0 + P.seg_rows - 1) / P.seg_rows;
  P.n_tasks = P.n_strips * P.n_segs;
  if (pushes || flags) sweep_hooks_fill(P, L, y0, y1, first, last, pushes);
  const int grid = std::max(1, std::min((P.n_tasks + kSSWarps - 1) / kSSWarps, sms * kSSCtasPerSm));
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
    int sw = try_sweep(plan, seg, L, y0, y1, src, dst, sms, last);
    if (sw == 0) sw = try_sweep5(plan, seg, L, y0, y1, src, dst, sms, first, last);
    if (sw == 0) sw = try_sweep_sobel(plan, seg, L, y0, y1, src, dst, sms, first, last);
    if (sw < 0) return B2I_ERR_CUDA;
    if (sw > 0) return B2I_OK;
  }
  TileParams P;
  memset(&P, 0, sizeof P);
  // tile height: tallest that still lets two CTAs share an SM; smaller for small images so that
  // every SM gets tiles
  int th = L.tile_h > 0 ? L.tile_h : env_int("B200IMG_TILE_H", 0);
  int raw_stages = env_int("B200IMG_RAW_STAGES", 2);
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
  P.debug = env_int("B200IMG_DEBUG", 0);
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
  const bool no_tma = (L.flags & B2I_LAUNCH_NO_TMA) || env_int("B200IMG_NO_TMA", 0);
  CUtensorMap src_map, dst_map, up_map, dn_map;
  memset(&src_map, 0, sizeof src_map);
  memset(&dst_map, 0, sizeof dst_map);
  memset(&up_map, 0, sizeof up_map);
  memset(&dn_map, 0, sizeof dn_map);
  const bool no_tma_load = no_tma || env_int("B200IMG_NO_TMA_LOAD", 0);
  const bool no_tma_store = no_tma || env_int("B200IMG_NO_TMA_STORE", 0) || (last && (L.flags & B2I_LAUNCH_DST_MCAST));
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
             !env_int("B200IMG_NO_VEC", 0);
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

}  // namespace

extern "C" int b2i_plan_create(const b2i_chain_t* chain, int in_channels, b2i_plan_t** out) {
  B2I_CHECK_ARG(chain && out, "null argument");
  const int cout = b2i_chain_out_channels(chain, in_channels);
  if (cout < 0) return cout;
  b2i_plan* plan = new b2i_plan();
  plan->chain = *chain;
  plan->cin = in_channels;
  plan->cout = cout;
  plan->halo = b2i_chain_halo(chain);
  memset(&plan->info, 0, sizeof plan->info);
  // fold runs of LUTs
  std::vector<Elem> el;
  for (int i = 0; i < chain->n_ops; ++i) {
    const b2i_op_t& op = chain->ops[i];
    if (op.kind == B2I_OP_LUT && !el.empty() && el.back().kind == B2I_OP_LUT) {
      compose_lut(el.back().lut, op.lut);
      continue;
    }
    Elem e;
    e.kind = op.kind;
    e.op = &plan->chain.ops[i];
    if (op.kind == B2I_OP_LUT) memcpy(e.lut, op.lut, 256);
    el.push_back(e);
  }
  if (segment_chain(el, in_channels, false, plan->segs) != B2I_OK) {
    delete plan;
    B2I_FAIL(B2I_ERR_UNSUPPORTED, "the chain cannot be scheduled on the GPU");
  }
  if (segment_chain(el, in_channels, true, plan->segs_split) != B2I_OK ||
      plan->segs_split.size() == plan->segs.size())
    plan->segs_split.clear();
  plan->info.cin = in_channels;
  plan->info.cout = cout;
  plan->info.halo = plan->halo;
  // the cut with more segments is the one reported: callers size their tmp images by it
  const std::vector<Segment>& rep = plan->segs_split.empty() ? plan->segs : plan->segs_split;
  plan->info.n_segments = (int)rep.size();
  for (size_t s = 0; s < rep.size() && s < B2I_MAX_OPS; ++s) {
    plan->info.seg_pointwise[s] = rep[s].pointwise;
    plan->info.seg_halo[s] = rep[s].halo;
    plan->info.seg_phases[s] = (int)rep[s].phases.size();
    plan->info.seg_cout[s] = rep[s].cout;
  }
  *out = plan;
  return B2I_OK;
}

extern "C" void b2i_plan_destroy(b2i_plan_t* plan) { delete plan; }
extern "C" int b2i_plan_halo(const b2i_plan_t* plan) { return plan ? plan->halo : 0; }
extern "C" int b2i_plan_info(const b2i_plan_t* plan, b2i_plan_info_t* info) {
  B2I_CHECK_ARG(plan && info, "null argument");
  *info = plan->info;
  return B2I_OK;
}

extern "C" int b2i_cuda_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

extern "C" int b2i_cuda_run(b2i_plan_t* plan, const b2i_launch_t* Lp) {
  B2I_CHECK_ARG(plan && Lp, "null argument");
  const b2i_launch_t& L = *Lp;
  B2I_CHECK_ARG(L.W > 0 && L.H > 0 && 0 <= L.y0 && L.y0 <= L.y1 && L.y1 <= L.H,
                "bad geometry %dx%d rows [%d,%d)", L.W, L.H, L.y0, L.y1);
  if (L.y0 == L.y1) return B2I_OK;
  plan->info.last_launches = 0;
  // the split cut only where the sweep kernels can take it: 8-pixel widths, TMA-encodable views
  // and tmp images; anything else runs the fused cut on the tile kernel
  auto tma_ok = [](const b2i_view_t& v) { return v.ptr && !(v.ptr & 15) && !(v.pitch & 15); };
  const bool use_split = !plan->segs_split.empty() && !(L.flags & (B2I_LAUNCH_NO_TMA | B2I_LAUNCH_NAIVE)) &&
                         !(L.W & 7) && L.W >= 16 && L.H >= 8 && tma_ok(L.src) && tma_ok(L.dst) &&
                         tma_ok(L.tmp[0]) && (plan->segs_split.size() < 3 || tma_ok(L.tmp[1])) &&
                         L.tmp[0].pitch >= (int64_t)L.W * 3 && tensor_map_encoder() &&
                         !env_int("B200IMG_NO_SWEEP", 0) && !env_int("B200IMG_NO_TMA", 0) &&
                         !env_int("B200IMG_NO_SPLIT", 0);
  const std::vector<Segment>& segs = use_split ? plan->segs_split : plan->segs;
  const int n = (int)segs.size();
  // Degenerate images (narrower or shorter than a border fold can reach) and explicit requests
  // go to the reference-style kernels.
  const bool tiny = plan->halo > 0 && (L.W <= 16 || L.H <= 16);
  if ((L.flags & B2I_LAUNCH_NAIVE) || tiny) {
    B2I_CHECK_ARG(L.y0 == 0 && L.y1 == L.H, "the reference-style path runs whole images only");
    B2I_CHECK_ARG(L.tmp[0].ptr && L.tmp[1].ptr, "the reference-style path needs both tmp images");
    B2I_CHECK_ARG(L.src.y_begin == 0 && L.dst.y_begin == 0 && L.src.y_end >= L.H && L.dst.y_end >= L.H,
                  "the reference-style path needs whole-image views");
    // LUT scratch: the tail of tmp[1] beyond its W*H*3 image bytes is not guaranteed, so LUTs
    // live in the first tmp row of a third region the caller passes through `status`
    B2I_CHECK_ARG(L.counters, "the reference-style path needs 256*B2I_MAX_OPS bytes at `counters`");
    return run_naive(&plan->chain, reinterpret_cast<const uint8_t*>(L.src.ptr), L.W, L.H, plan->cin,
                     (size_t)L.src.pitch, reinterpret_cast<uint8_t*>(L.dst.ptr), (size_t)L.dst.pitch,
                     reinterpret_cast<uint8_t*>(L.tmp[0].ptr), reinterpret_cast<uint8_t*>(L.tmp[1].ptr),
                     reinterpret_cast<uint8_t*>(L.counters), (cudaStream_t)L.stream);
  }
  // Multi-segment plans over-compute: segment s produces the rows later segments still read
  // (communication-avoiding), so no synchronisation is needed inside one call.
  std::vector<int> rem(n + 1, 0);
  for (int s = n - 1; s >= 0; --s) rem[s] = rem[s + 1] + segs[s].halo;
  b2i_view_t cur = L.src;
  for (int s = 0; s < n; ++s) {
    const Segment& seg = segs[s];
    const bool last = s == n - 1;
    const int ya = last ? L.y0 : std::max(0, L.y0 - rem[s + 1]);
    const int yb = last ? L.y1 : std::min(L.H, L.y1 + rem[s + 1]);
    b2i_view_t out = L.dst;
    if (!last) {
      out = L.tmp[s & 1];
      B2I_CHECK_ARG(out.ptr, "plan has %d segments: tmp[%d] is needed", n, s & 1);
      B2I_CHECK_ARG(out.pitch >= (int64_t)L.W * seg.cout, "tmp pitch too small");
    }
    if (last) g_wait_done_fused = false;
    if (seg.pointwise) B2I_TRY(launch_pointwise_segment(plan, seg, L, ya, yb, cur, out));
    else B2I_TRY(launch_tile_segment(plan, seg, L, ya, yb, cur, out, s == 0, last));
    cur = out;
    if (last && L.done_flag && !g_wait_done_fused)  // kernels without the fused flags: one-thread kernels follow
      B2I_TRY(b2i_cuda_set_flags(&L.done_flag, 1, L.signal_epoch, L.stream));
    if (last && L.wait_done_mask && L.wait_done_flags && !g_wait_done_fused) {
      uint64_t fl[B2I_MAX_GPUS];
      int nf = 0;
      for (int r = 0; r < B2I_MAX_GPUS; ++r)
        if (L.wait_done_mask >> r & 1) fl[nf++] = L.wait_done_flags + 4ull * r;
      B2I_CHECK_ARG(L.status, "waiting needs the status word");
      B2I_TRY(b2i_cuda_wait_flags(fl, nf, L.signal_epoch, L.status, L.timeout_ns, L.stream));
    }
    if (!last) {  // later segments may read exactly what this one produced
      cur.ptr = out.ptr + (uint64_t)((int64_t)(ya - out.y_begin) * out.pitch);
      cur.y_begin = ya;
      cur.y_end = yb;
    }
  }
  return B2I_OK;
}

extern "C" int b2i_cuda_run_image(b2i_plan_t* plan, uint64_t src, int64_t src_pitch, uint64_t dst,
                                  int64_t dst_pitch, int W, int H, uint64_t tmp0, uint64_t tmp1,
                                  uint64_t lut_scratch, void* stream, int flags) {
  B2I_CHECK_ARG(plan, "null plan");
  b2i_launch_t L;
  memset(&L, 0, sizeof L);
  L.W = W;
  L.H = H;
  L.y0 = 0;
  L.y1 = H;
  L.src = {src, src_pitch, 0, H};
  L.dst = {dst, dst_pitch, 0, H};
  L.tmp[0] = {tmp0, (int64_t)W * 3, 0, H};
  L.tmp[1] = {tmp1, (int64_t)W * 3, 0, H};
  L.counters = lut_scratch;
  L.stream = stream;
  L.flags = flags;
  return b2i_cuda_run(plan, &L);
}
