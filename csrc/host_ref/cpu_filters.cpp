// Host reference filters: the numerics oracle for every CUDA kernel, and the compute backend
// of the host "ranks" mode (BASELINE.json config 1).
//
// Semantics restate what the reference computes where it is well defined:
//   gray_ref   kernel.cu:39-42   per-term truncation == B*11/100 + G*59/100 + R*3/10
//   gray_cv    kern.cpp:73       cv::cvtColor 15-bit fixed point
//   lut        kernel.cu:49-58 / kern.cpp:74 (tables built in core/chain.cpp)
//   stencil    kernel.cu:64-94 / kern.cpp:75, but always out-of-place (the reference's in-place
//              update races with its neighbours' reads) and over the whole image, not per band
//   expand     kernel.cu:210 / kern.cpp:77
// All stencil sums are exact int32 (the reference's float accumulation is exact for u8 too).
#include <algorithm>
#include <chrono>
#include <cstring>
#include <thread>
#include <vector>

#include "../core/error.hpp"

namespace b2i {

// Rows [ya, yb) of a W x H image with C interleaved channels, tightly packed.
struct RowWindow {
  std::vector<uint8_t> buf;
  int W = 0, C = 0, ya = 0, yb = 0;
  void reset(int w, int c, int a, int b) {
    W = w;
    C = c;
    ya = a;
    yb = b;
    buf.resize((size_t)std::max(0, b - a) * w * c);
  }
  size_t pitch() const { return (size_t)W * C; }
  uint8_t* row(int y) { return buf.data() + (size_t)(y - ya) * pitch(); }
  const uint8_t* row(int y) const { return buf.data() + (size_t)(y - ya) * pitch(); }
};

// Maps an out-of-range coordinate into [0, n) under `border`; -1 means "reads as zero".
static inline int map_coord(int p, int n, int border) {
  if ((unsigned)p < (unsigned)n) return p;
  switch (border) {
    case B2I_BORDER_REFLECT101:
      if (n == 1) return 0;
      while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
      return p;
    case B2I_BORDER_REPLICATE:
      return p < 0 ? 0 : n - 1;
    default:
      return -1;
  }
}

static inline uint8_t clamp_u8(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }

static inline int radius_of(const b2i_op_t& op) {
  if (op.kind == B2I_OP_STENCIL) return op.ksize / 2;
  if (op.kind == B2I_OP_SOBEL) return 1;
  return 0;
}

static const int8_t kSobelX[9] = {-1, 0, 1, -2, 0, 2, -1, 0, 1};
static const int8_t kSobelY[9] = {-1, -2, -1, 0, 0, 0, 1, 2, 1};

// One stencil / Sobel output pixel-channel with full border handling (slow path).
static inline uint8_t stencil_at(const b2i_op_t& op, const RowWindow& in, int H, int x, int y,
                                 int c) {
  const int k = op.kind == B2I_OP_SOBEL ? 3 : op.ksize, r = k / 2, W = in.W, C = in.C;
  int s = 0, sy = 0;
  for (int j = 0; j < k; ++j) {
    const int my = map_coord(y + j - r, H, op.border);
    for (int i = 0; i < k; ++i) {
      const int mx = map_coord(x + i - r, W, op.border);
      const int v = (my < 0 || mx < 0) ? 0 : in.row(my)[(size_t)mx * C + c];
      if (op.kind == B2I_OP_SOBEL) {
        s += kSobelX[j * 3 + i] * v;
        sy += kSobelY[j * 3 + i] * v;
      } else {
        s += op.weights[j * k + i] * v;
      }
    }
  }
  if (op.kind == B2I_OP_SOBEL) return clamp_u8(std::abs(s) + std::abs(sy));
  return clamp_u8((s + op.bias) >> op.shift);
}

// Pixels [x0, x1] of one row where every tap is inside the image: no coordinate mapping.
template <int K, bool SOBEL>
static void interior_run(const b2i_op_t& op, const uint8_t* const (&rows)[5], uint8_t* o, int x0, int x1, int C) {
  constexpr int r = K / 2;
  int w[K * K];
  for (int q = 0; q < K * K; ++q) w[q] = SOBEL ? 0 : op.weights[q];
  const size_t n = (size_t)(x1 - x0 + 1) * C;  // channel values are independent: one flat loop
  const size_t first = (size_t)(x0 - r) * C;
  uint8_t* out = o + (size_t)x0 * C;
  for (size_t e = 0; e < n; ++e) {
    const size_t base = first + e;
    if (SOBEL) {
      int gx = 0, gy = 0;
      for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i) {
          const int v = rows[j][base + (size_t)i * C];
          gx += kSobelX[j * 3 + i] * v;
          gy += kSobelY[j * 3 + i] * v;
        }
      out[e] = clamp_u8(std::abs(gx) + std::abs(gy));
    } else {
      int s = 0;
      for (int j = 0; j < K; ++j)
        for (int i = 0; i < K; ++i) s += w[j * K + i] * rows[j][base + (size_t)i * C];
      out[e] = clamp_u8((s + op.bias) >> op.shift);
    }
  }
}

// Rows [out.ya, out.yb) of `op` applied to the whole H-row image; `in` must hold every source
// row the border rule can touch for those rows.
static void apply_stencil(const b2i_op_t& op, const RowWindow& in, int H, RowWindow& out) {
  const int W = in.W, C = in.C;
  const int k = op.kind == B2I_OP_SOBEL ? 3 : op.ksize, r = k / 2;
  const bool ref = op.border == B2I_BORDER_REF;
  // REF: computed set x in [r+1, W-r-1], y in [r+1, H-r-1] (kernel.cu:83 minus its wrap/OOB part)
  const int cx0 = ref ? r + 1 : 0, cx1 = ref ? W - r - 1 : W - 1;
  const int cy0 = ref ? r + 1 : 0, cy1 = ref ? H - r - 1 : H - 1;
  // interior: every tap in range, no mapping needed
  const int ix0 = std::max(cx0, r), ix1 = std::min(cx1, W - 1 - r);
  for (int y = out.ya; y < out.yb; ++y) {
    uint8_t* o = out.row(y);
    if (y < cy0 || y > cy1 || cx1 < cx0) {
      memcpy(o, in.row(y), in.pitch());
      continue;
    }
    const bool y_interior = y >= r && y <= H - 1 - r;
    if (ref) {
      if (cx0 > 0) memcpy(o, in.row(y), (size_t)std::min(cx0, W) * C);
      if (cx1 + 1 < W) memcpy(o + (size_t)(cx1 + 1) * C, in.row(y) + (size_t)(cx1 + 1) * C,
                              (size_t)(W - cx1 - 1) * C);
    }
    for (int x = cx0; x <= cx1; ++x) {
      if (y_interior && x >= ix0 && x <= ix1) {
        // fast path for the run [x, ix1]
        const uint8_t* rows[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
        for (int j = 0; j < k; ++j) rows[j] = in.row(y + j - r);
        // the same sums as stencil_at, with the kernel size known to the compiler
        if (op.kind == B2I_OP_SOBEL) interior_run<3, true>(op, rows, o, x, ix1, C);
        else if (k == 3) interior_run<3, false>(op, rows, o, x, ix1, C);
        else interior_run<5, false>(op, rows, o, x, ix1, C);
        x = ix1;  // the outer loop increments
        continue;
      }
      for (int c = 0; c < C; ++c) o[(size_t)x * C + c] = stencil_at(op, in, H, x, y, c);
    }
  }
}

// v*11/100, v*59/100, v*3/10 with the reference's per-term integer truncation (kernel.cu:39-42).
struct GrayRefTables {
  uint8_t b[256], g[256], r[256];
  GrayRefTables() {
    for (int v = 0; v < 256; ++v) {
      b[v] = (uint8_t)(v * 11 / 100);
      g[v] = (uint8_t)(v * 59 / 100);
      r[v] = (uint8_t)(v * 3 / 10);
    }
  }
};

static void apply_pointwise(const b2i_op_t& op, int order, const RowWindow& in, RowWindow& out) {
  const int W = in.W;
  for (int y = out.ya; y < out.yb; ++y) {
    const uint8_t* s = in.row(y);
    uint8_t* o = out.row(y);
    switch (op.kind) {
      case B2I_OP_GRAY_REF: {
        const int ib = order == B2I_ORDER_BGR ? 0 : 2, ir = 2 - ib;
        // kernel.cu:39-42 term by term; the three truncated products come from 256-entry tables
        static const GrayRefTables t;
        for (int x = 0; x < W; ++x)
          o[x] = (uint8_t)(t.b[s[3 * x + ib]] + t.g[s[3 * x + 1]] + t.r[s[3 * x + ir]]);
        break;
      }
      case B2I_OP_GRAY_CV: {
        const int ib = order == B2I_ORDER_BGR ? 0 : 2, ir = 2 - ib;
        for (int x = 0; x < W; ++x)
          o[x] = (uint8_t)((3735 * s[3 * x + ib] + 19235 * s[3 * x + 1] + 9798 * s[3 * x + ir] + 16384) >> 15);
        break;
      }
      case B2I_OP_LUT: {
        const size_t n = in.pitch();
        for (size_t i = 0; i < n; ++i) o[i] = op.lut[s[i]];
        break;
      }
      case B2I_OP_EXPAND:
        for (int x = 0; x < W; ++x) o[3 * x] = o[3 * x + 1] = o[3 * x + 2] = s[x];
        break;
      default:
        break;
    }
  }
}

// Row ranges each stage has to produce, walking backwards from the requested rows [y0, y1):
// need[k] is the range of stage k's input (need[0] = source rows), need[n] = [y0, y1). A border
// rule other than REF may fold a tap back over the edge (reflect101 reaches row r from row 0),
// so near an edge the range is widened to everything the fold can land on.
static void plan_rows(const b2i_chain_t* chain, int H, int y0, int y1, std::vector<int>& need_a,
                      std::vector<int>& need_b) {
  const int n = chain->n_ops;
  need_a.assign(n + 1, 0);
  need_b.assign(n + 1, 0);
  need_a[n] = y0;
  need_b[n] = y1;
  for (int k = n - 1; k >= 0; --k) {
    const int r = radius_of(chain->ops[k]);
    int a = need_a[k + 1] - r, b = need_b[k + 1] + r;
    if (r && H <= 2 * r) {  // degenerate: folds can bounce, take the whole image
      a = 0;
      b = H;
    }
    need_a[k] = std::max(0, a);
    need_b[k] = std::min(H, b);
    if (r && a < 0) need_b[k] = std::min(H, std::max(need_b[k], r + 1));
    if (r && b > H) need_a[k] = std::max(0, std::min(need_a[k], H - r - 1));
  }
}

// Computes rows [y0, y1) of the chain's result. `src` points at global row `src_ya` and holds
// rows up to `src_yb` (exclusive); those must cover [y0-halo, y1+halo) clipped to the image.
int run_rows_impl(const b2i_chain_t* chain, const uint8_t* src, size_t src_pitch, int src_ya,
                  int src_yb, int W, int H, int C, int y0, int y1, uint8_t* dst,
                  size_t dst_pitch) {
  B2I_CHECK_ARG(chain && src && dst, "null argument");
  B2I_CHECK_ARG(W > 0 && H > 0 && (C == 1 || C == 3), "bad geometry %dx%dx%d", W, H, C);
  B2I_CHECK_ARG(0 <= y0 && y0 <= y1 && y1 <= H, "bad row range [%d,%d) of %d", y0, y1, H);
  const int out_c = b2i_chain_out_channels(chain, C);
  if (out_c < 0) return out_c;
  B2I_CHECK_ARG(src_pitch >= (size_t)W * C && dst_pitch >= (size_t)W * out_c, "pitch too small");
  if (y0 == y1) return B2I_OK;
  const int n = chain->n_ops;
  std::vector<int> need_a, need_b;
  plan_rows(chain, H, y0, y1, need_a, need_b);
  B2I_CHECK_ARG(src_ya <= need_a[0] && need_b[0] <= src_yb,
                "source rows [%d,%d) do not cover the needed rows [%d,%d)", src_ya, src_yb,
                need_a[0], need_b[0]);
  RowWindow cur, nxt;
  cur.reset(W, C, need_a[0], need_b[0]);
  for (int y = need_a[0]; y < need_b[0]; ++y)
    memcpy(cur.row(y), src + (size_t)(y - src_ya) * src_pitch, cur.pitch());
  for (int k = 0; k < n; ++k) {
    const b2i_op_t& op = chain->ops[k];
    int oc = cur.C;
    if (op.kind == B2I_OP_GRAY_REF || op.kind == B2I_OP_GRAY_CV) oc = 1;
    if (op.kind == B2I_OP_EXPAND) oc = 3;
    nxt.reset(W, oc, need_a[k + 1], need_b[k + 1]);
    if (radius_of(op)) apply_stencil(op, cur, H, nxt);
    else apply_pointwise(op, chain->order, cur, nxt);
    std::swap(cur, nxt);
  }
  for (int y = y0; y < y1; ++y) memcpy(dst + (size_t)(y - y0) * dst_pitch, cur.row(y), cur.pitch());
  return B2I_OK;
}

}  // namespace b2i

extern "C" int b2i_host_run_rows(const b2i_chain_t* chain, const uint8_t* src, int W, int H,
                                 int C, size_t src_pitch, int y0, int y1, uint8_t* dst,
                                 size_t dst_pitch) {
  return b2i::run_rows_impl(chain, src, src_pitch, 0, H, W, H, C, y0, y1, dst, dst_pitch);
}

extern "C" int b2i_host_run(const b2i_chain_t* chain, const uint8_t* src, int W, int H, int C,
                            size_t src_pitch, uint8_t* dst, size_t dst_pitch) {
  return b2i::run_rows_impl(chain, src, src_pitch, 0, H, W, H, C, 0, H, dst, dst_pitch);
}

extern "C" int b2i_band_rows(int H, int n, int i, int* y0, int* y1) {
  B2I_CHECK_ARG(H >= 0 && n > 0 && i >= 0 && i < n && y0 && y1, "bad band request %d/%d of %d rows",
                i, n, H);
  *y0 = (int)(((int64_t)i * H) / n);
  *y1 = (int)(((int64_t)(i + 1) * H) / n);
  return B2I_OK;
}

// Host workers standing in for the reference's MPI ranks. The data flow is the reference's:
// root-owned source, one private band buffer per rank filled by the root (MPI_Scatter,
// kernel.cu:137), independent filtering, rows handed back into the root-owned destination
// (MPI_Gather, kernel.cu:223). Differences: bands carry halo rows so the result has no seams,
// and remainder rows are not dropped (kernel.cu:117).
extern "C" int b2i_host_run_ranks(const b2i_chain_t* chain, const uint8_t* src, int W, int H,
                                  int C, size_t src_pitch, uint8_t* dst, size_t dst_pitch,
                                  int ranks, int seams, double* elapsed_ms) {
  B2I_CHECK_ARG(chain && src && dst, "null argument");
  B2I_CHECK_ARG(ranks >= 1 && ranks <= 1024, "ranks must be in [1,1024], got %d", ranks);
  const int out_c = b2i_chain_out_channels(chain, C);
  if (out_c < 0) return out_c;
  const auto t0 = std::chrono::steady_clock::now();
  struct Band {
    int y0, y1, ya, yb;
    std::vector<uint8_t> in, out;
    int rc = B2I_OK;
    std::string err;
  };
  std::vector<Band> bands(ranks);
  // "scatter": the root copies each rank's rows (plus halo) into that rank's private buffer
  for (int i = 0; i < ranks; ++i) {
    Band& b = bands[i];
    b2i_band_rows(H, ranks, i, &b.y0, &b.y1);
    b.ya = b.y0;
    b.yb = b.y1;
    if (!seams && b.y1 > b.y0) {
      std::vector<int> na, nb;
      b2i::plan_rows(chain, H, b.y0, b.y1, na, nb);
      b.ya = na[0];
      b.yb = nb[0];
    }
    b.in.resize((size_t)(b.yb - b.ya) * W * C);
    for (int y = b.ya; y < b.yb; ++y)
      memcpy(b.in.data() + (size_t)(y - b.ya) * W * C, src + (size_t)y * src_pitch, (size_t)W * C);
    b.out.resize((size_t)(b.y1 - b.y0) * W * out_c);
  }
  std::vector<std::thread> workers;
  for (int i = 0; i < ranks; ++i) {
    workers.emplace_back([&, i] {
      Band& b = bands[i];
      if (b.y1 == b.y0) return;
      if (seams)  // the band is a whole image of its own, like kern.cpp:73-77 on part_image
        b.rc = b2i::run_rows_impl(chain, b.in.data(), (size_t)W * C, 0, b.y1 - b.y0, W,
                                  b.y1 - b.y0, C, 0, b.y1 - b.y0, b.out.data(), (size_t)W * out_c);
      else
        b.rc = b2i::run_rows_impl(chain, b.in.data(), (size_t)W * C, b.ya, b.yb, W, H, C, b.y0,
                                  b.y1, b.out.data(), (size_t)W * out_c);
      if (b.rc != B2I_OK) b.err = b2i_last_error();  // thread-local: carry it to the caller
    });
  }
  for (auto& t : workers) t.join();
  for (int i = 0; i < ranks; ++i)
    if (bands[i].rc != B2I_OK) {
      b2i::last_error_slot() = "host rank " + std::to_string(i) + ": " + bands[i].err;
      return bands[i].rc;
    }
  // "gather": rows back into the root-owned image
  for (int i = 0; i < ranks; ++i)
    for (int y = bands[i].y0; y < bands[i].y1; ++y)
      memcpy(dst + (size_t)y * dst_pitch,
             bands[i].out.data() + (size_t)(y - bands[i].y0) * W * out_c, (size_t)W * out_c);
  if (elapsed_ms)
    *elapsed_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0)
                      .count();
  return B2I_OK;
}
