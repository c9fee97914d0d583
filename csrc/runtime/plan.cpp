// Chain -> plan compiler and the band-launch entry point.
//
// The reference has no such layer: its pipeline is three literal kernel launches
// (kernel.cu:192-195). Here a chain is compiled once into segments; each segment is one launch
// of a sweep kernel (kernels/sweep3.cu, sweep5.cu, sweep_sobel.cu), of the tile-chain kernel
// (kernels/tile_chain.cu) or of the streaming pointwise kernel (kernels/pointwise.cu).
// Compilation folds every run of u8->u8 pointwise ops into one LUT, attaches LUTs and the
// gray->BGR expand to the neighbouring stencil and lowers stencil weights to packed dp4a operands;
// the launchers are in launch_sweep.cpp and launch_tile.cpp.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../core/trace.hpp"
#include "plan_internal.hpp"

using namespace b2i;

namespace {

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
        if (!B2I_ENV("B200IMG_NO_SEP", 0) && separable(op, wv, wh)) {
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
  // exactly one stencil and nothing else: the tensor-core kernel's shape (kernels/stencil_mma.cu)
  seg.mma_ok = 0;
  if (el.size() == 1 && (el[0].kind == B2I_OP_STENCIL || el[0].kind == B2I_OP_SOBEL) && cin == c &&
      el[0].op->border != B2I_BORDER_REF) {
    const b2i_op_t& op = *el[0].op;
    memset(seg.mma_w, 0, sizeof seg.mma_w);
    seg.mma_border = op.border;
    if (op.kind == B2I_OP_SOBEL) {
      const int8_t gx[9] = {-1, 0, 1, -2, 0, 2, -1, 0, 1}, gy[9] = {-1, -2, -1, 0, 0, 0, 1, 2, 1};
      for (int q = 0; q < 9; ++q) {
        seg.mma_w[0][q / 3][q % 3] = gx[q];
        seg.mma_w[1][q / 3][q % 3] = gy[q];
      }
      seg.mma_mode = SM_SOBEL;
      seg.mma_R = 1;
      seg.mma_shift = seg.mma_bias = 0;
      seg.mma_ok = 1;
    } else if ((op.ksize == 3 || op.ksize == 5) && op.shift >= 0 && op.shift <= 16) {
      for (int q = 0; q < op.ksize * op.ksize; ++q) seg.mma_w[0][q / op.ksize][q % op.ksize] = op.weights[q];
      seg.mma_mode = SM_LINEAR;
      seg.mma_R = op.ksize / 2;
      seg.mma_shift = op.shift;
      seg.mma_bias = op.bias;
      seg.mma_ok = 1;
    }
  }
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

// A segment the sweep kernels sweep5.cu / sweep_sobel.cu take (given a fitting image).
bool sweep_shape(const Segment& seg) {
  if (!seg.pointwise && seg.mma_ok) return true;  // any channel count: the tensor-core kernel works on bytes
  if (seg.pointwise || seg.cin != 3 || seg.cout != 3) return false;
  if (seg.phases.size() == 2)
    return seg.phases[0].type == PH_SEPH && seg.phases[1].type == PH_SEPV && !seg.phases[1].expand;
  return seg.phases.size() == 1 && seg.phases[0].type == PH_SOBEL && !seg.phases[0].expand &&
         seg.phases[0].border != B2I_BORDER_REF;
}

// gray (+LUT) -> one 3x3 stencil on the gray plane (+LUT, +expand): what sweep3.cu takes (try_sweep).
bool sweep3_shape(const Segment& seg) {
  return !seg.pointwise && seg.cin == 3 && seg.phases.size() == 2 &&
         (seg.phases[0].type == PH_GRAY_REF || seg.phases[0].type == PH_GRAY_CV) &&
         seg.phases[1].type == PH_STENCIL && seg.phases[1].ksize == 3 && seg.phases[1].cin == 1;
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
      // a pointwise prefix followed by a stencil the tensor-core kernel takes on its own: two
      // launches (streaming kernel, then stencil_mma) beat the fused tile kernel -- unless the
      // fused segment is the reference pipeline's shape, which sweep3.cu runs in one pass
      if (split && have && best.pointwise && !seg.pointwise && !sweep3_shape(seg)) {
        Segment alone;
        std::vector<Elem> one(el.begin() + end - 1, el.begin() + end);
        if (build_tile_segment(one, best.cout, alone) == B2I_OK && alone.mma_ok) break;
      }
      best = seg;
      best_end = end;
      have = true;
    }
    if (!have) return B2I_ERR_UNSUPPORTED;
    c = best.cout;
    if (!best.pointwise && !best.phases.empty() && best.phases[0].lut >= 0)
    {
      best.lut0_affine = affine_lut(best.luts[best.phases[0].lut], &best.aff_a, &best.aff_b, &best.aff_s);
      if (B2I_ENV("B200IMG_DEBUG", 0) & 2)
        fprintf(stderr, "b200img: first LUT of segment %zu %s clamp((%d v + %d) >> %d)\n", out.size(),
                best.lut0_affine ? "is" : "is not", best.aff_a, best.aff_b, best.aff_s);
    }
    out.push_back(best);
    i = best_end;
  }
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
// A plan keeps the figures of its last launch (b2i_plan_info: last_*), so one plan must not be
// launched from two host threads at once; a thread that needs to gets its own copy.
extern "C" int b2i_plan_clone(const b2i_plan_t* plan, b2i_plan_t** out) {
  B2I_CHECK_ARG(plan && out, "null argument");
  *out = new b2i_plan(*plan);
  return B2I_OK;
}
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
  NvtxRange range("b2i_cuda_run");
  plan->info.last_launches = 0;
  // the split cut only where the sweep / tensor-core kernels can take it: TMA-encodable views (16-byte
  // pitches; the kernels check the widths they need themselves)
  // and tmp images; anything else runs the fused cut on the tile kernel
  auto tma_ok = [](const b2i_view_t& v) { return v.ptr && !(v.ptr & 15) && !(v.pitch & 15); };
  const bool use_split = !plan->segs_split.empty() && !(L.flags & (B2I_LAUNCH_NO_TMA | B2I_LAUNCH_NAIVE)) &&
                         L.W >= 16 && L.H >= 8 && tma_ok(L.src) && tma_ok(L.dst) &&
                         tma_ok(L.tmp[0]) && (plan->segs_split.size() < 3 || tma_ok(L.tmp[1])) &&
                         L.tmp[0].pitch >= (int64_t)L.W * 3 && tensor_map_encoder() &&
                         !B2I_ENV("B200IMG_NO_SWEEP", 0) && !B2I_ENV("B200IMG_NO_TMA", 0) &&
                         !B2I_ENV("B200IMG_NO_SPLIT", 0);
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
    // live in a third region that the caller passes through `counters`
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
