// The tile-chain kernel: one persistent sm_100a kernel that runs a whole filter-chain segment
// per tile in shared memory.
//
//   global --TMA(UTMALDG)--> raw tile --phase 0..n-1 (smem -> smem)--> staging --TMA(UTMASTG)--> global
//
// It replaces the reference's grayscaleKernel / contrastKernel / embossKernel launches and the
// host-side GRAY2BGR (kernel.cu:31-94,192-195,210) by one pass over HBM: 3 B/px in, 3 B/px out
// instead of 8 B/px of device traffic plus a D2H copy and a host expand. The source and the
// destination are "views" that may live on another GPU (peer-mapped or IPC-mapped memory), which
// is how scatter (kernel.cu:137), the halo exchange the reference lacks, and gather
// (kernel.cu:223) are fused into this kernel instead of being separate transfers:
//   * scatter : the source tensor map points at the owner GPU's image, tiles are TMA-loaded
//               across NVLink;
//   * halo    : boundary tiles are additionally stored into the neighbours' ghost rows, then a
//               system-scope flag is released; consumers acquire it before loading ghost rows;
//   * gather  : the destination tensor map points into the root GPU's output image.
#include "b200img.h"
#include "device_utils.cuh"
#include "tile_plan.hpp"

namespace b2i {
using namespace dev;

namespace {

struct Tile {
  int tx0, ty0;  // global pixel coordinates of the tile's first output pixel
  bool edge;     // some plane cell of this tile lies outside the image
  bool first_row, last_row;
};

__device__ __forceinline__ Tile tile_of(const TileParams& P, int t) {
  // Boundary-first order: the first and the last tile row of a band are produced first so that
  // their halo pushes reach the neighbours early.
  int trow = t / P.tiles_x, tcol = t - trow * P.tiles_x;
  if (P.push_rows > 0 && P.tiles_y > 2) {
    if (trow == 1) trow = P.tiles_y - 1;
    else if (trow > 1) trow -= 1;
  }
  Tile T;
  T.tx0 = tcol * kTileW;
  T.ty0 = P.y0 + trow * P.th;
  T.first_row = trow == 0;
  T.last_row = trow == P.tiles_y - 1;
  const int R = P.halo;
  T.edge = T.tx0 - kApron < 0 || T.tx0 + kTileW + kApron > P.W || T.ty0 - R < 0 ||
           T.ty0 + P.th + R + kRowGroup > P.H;
  return T;
}

// ---- tile input -------------------------------------------------------------------------------

// Fallback loader for sources TMA cannot describe (pitch or base not 16-byte aligned): every
// thread copies words, zero-filling what lies outside the view, exactly like TMA's OOB fill.
__device__ void load_tile_generic(const TileParams& P, const Tile& T, uint8_t* raw) {
  const int pitch = P.buf[BUF_RAW].pitch, words = pitch >> 2;
  const int64_t xb0 = (int64_t)(T.tx0 - kApron) * P.cin;
  const int64_t row_bytes = (int64_t)P.W * P.cin;
  const uint8_t* src = reinterpret_cast<const uint8_t*>(P.src_ptr);
  for (int idx = threadIdx.x; idx < words * P.raw_rows; idx += kTileThreads) {
    const int pr = idx / words, w = idx - pr * words;
    const int gy = T.ty0 - P.halo + pr;
    const int64_t xb = xb0 + 4 * w;
    uint32_t v = 0;
    if (gy >= P.src_y_begin && gy < P.src_y_end && xb + 3 >= 0 && xb < row_bytes) {
      const uint8_t* p = src + (int64_t)(gy - P.src_y_begin) * P.src_pitch + xb;
      if (xb >= 0 && xb + 3 < row_bytes && (reinterpret_cast<uintptr_t>(p) & 3) == 0) {
        v = *reinterpret_cast<const uint32_t*>(p);
      } else {
#pragma unroll
        for (int b = 0; b < 4; ++b)
          if (xb + b >= 0 && xb + b < row_bytes) v |= (uint32_t)p[b] << (8 * b);
      }
    }
    *reinterpret_cast<uint32_t*>(raw + (size_t)pr * pitch + 4 * w) = v;
  }
}

// Paints the border rule of the next stencil onto the cells of `plane` that lie outside the
// image (edge tiles only). Sources are in-image cells of the same plane, which this never writes.
__device__ void fix_border(const TileParams& P, const Tile& T, uint8_t* plane, int pitch, int C,
                           int margin, int border) {
  const int R = P.halo;
  const int rows = P.th + 2 * margin, cols = kTileW + 2 * margin;
  for (int idx = threadIdx.x; idx < rows * cols; idx += kTileThreads) {
    const int rr = idx / cols, cc = idx - rr * cols;
    const int pr = R - margin + rr, pc = kApron - margin + cc;
    const int gx = T.tx0 - kApron + pc, gy = T.ty0 - R + pr;
    if ((unsigned)gx < (unsigned)P.W && (unsigned)gy < (unsigned)P.H) continue;
    uint8_t* d = plane + (size_t)pr * pitch + pc * C;
    if (border == B2I_BORDER_ZERO) {
      for (int c = 0; c < C; ++c) d[c] = 0;
      continue;
    }
    int sx = gx, sy = gy;
    if (border == B2I_BORDER_REFLECT101) {
      sx = sx < 0 ? -sx : (sx >= P.W ? 2 * P.W - 2 - sx : sx);
      sy = sy < 0 ? -sy : (sy >= P.H ? 2 * P.H - 2 - sy : sy);
    } else {  // replicate
      sx = min(max(sx, 0), P.W - 1);
      sy = min(max(sy, 0), P.H - 1);
    }
    const int spr = sy - (T.ty0 - R), spc = sx - (T.tx0 - kApron);
    // sources outside this tile's planes belong to cells no valid output depends on
    if (spr < 0 || spr >= P.th + 2 * R || spc < 0 || spc >= kPlaneW) continue;
    const uint8_t* s = plane + (size_t)spr * pitch + spc * C;
    for (int c = 0; c < C; ++c) d[c] = s[c];
  }
}

// ---- pointwise phases (smem -> smem) ------------------------------------------------------------

__device__ __forceinline__ uint32_t lut4(const uint8_t* lut, uint32_t w) {
  const uint32_t a = lut[w & 0xFF], b = lut[(w >> 8) & 0xFF], c = lut[(w >> 16) & 0xFF],
                 d = lut[w >> 24];
  return pack4_lo(a, b, c, d);
}

template <int TYPE>
__device__ void pw_phase(const TileParams& P, const PhaseDesc& ph, uint8_t* sbase, int spitch,
                         int sxorg, int syorg, uint8_t* dbase, int dpitch, int dxorg, int dyorg,
                         const uint8_t* lut) {
  const int R = P.halo, m = ph.margin, mc = ph.margin_x > 0 ? kMarginCols : 0;
  const int ux = (kTileW + 2 * mc) >> 2;  // units of 4 pixels per row
  const int rows = P.th + 2 * m;
  for (int idx = threadIdx.x; idx < ux * rows; idx += kTileThreads) {
    const int rr = idx / ux, u = idx - rr * ux;
    const int pr = R - m + rr, pc = kApron - mc + 4 * u;
    const uint8_t* s = sbase + (size_t)(pr - syorg) * spitch + (pc - sxorg) * ph.cin;
    uint8_t* d = dbase + (size_t)(pr - dyorg) * dpitch + (pc - dxorg) * ph.cout;
    if (TYPE == PH_COPY) {
      for (int c = 0; c < ph.cin; ++c) {
        uint32_t w = reinterpret_cast<const uint32_t*>(s)[c];
        if (lut) w = lut4(lut, w);
        reinterpret_cast<uint32_t*>(d)[c] = w;
      }
    } else {  // PH_EXPAND
      uint32_t w = reinterpret_cast<const uint32_t*>(s)[0];
      if (lut) w = lut4(lut, w);
      reinterpret_cast<uint32_t*>(d)[0] = __byte_perm(w, 0, 0x1000);
      reinterpret_cast<uint32_t*>(d)[1] = __byte_perm(w, 0, 0x2211);
      reinterpret_cast<uint32_t*>(d)[2] = __byte_perm(w, 0, 0x3332);
    }
  }
}

// Gray conversion of a plane, 8 pixels per item: three LDS.64 in, one STS.64 out, the source and
// destination addresses advanced incrementally (the first version of this phase spent more
// instructions on indexing 4-pixel units than on the conversion itself: ncu, profiles/).
template <int TYPE, bool LUT>
__device__ __forceinline__ void gray8(const uint8_t* s, uint8_t* d, const GrayRefCoef& gk, int rgb,
                                      const uint8_t* lut) {
  const uint2 v0 = *reinterpret_cast<const uint2*>(s), v1 = *reinterpret_cast<const uint2*>(s + 8),
              v2 = *reinterpret_cast<const uint2*>(s + 16);
  const uint32_t w0 = v0.x, w1 = v0.y, w2 = v1.x, w3 = v1.y, w4 = v2.x, w5 = v2.y;
  uint32_t g[8];
  if (TYPE == PH_GRAY_REF) {
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
  if (LUT) {
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = lut[g[i] & 0xFF];
  }
  *reinterpret_cast<uint2*>(d) = make_uint2(pack4_lo(g[0], g[1], g[2], g[3]), pack4_lo(g[4], g[5], g[6], g[7]));
}

template <int TYPE, bool LUT>
__device__ void gray_phase(const TileParams& P, const PhaseDesc& ph, const uint8_t* sbase, int spitch,
                           int sxorg, int syorg, uint8_t* dbase, int dpitch, int dxorg, int dyorg,
                           const uint8_t* lut) {
  const int R = P.halo, m = ph.margin, mc = ph.margin_x > 0 ? kMarginCols : 0;
  const int ux = ph.units;  // items of 8 pixels per row: 32 or 34
  const int n = ph.n_items;
  const GrayRefCoef gk = gray_ref_coef(P.rgb_order);
  const int rgb = P.rgb_order;
  int idx = threadIdx.x;
  const int row0 = (idx * ph.units_magic) >> 20;
  int col = idx - row0 * ux;
  const int drow = ph.drow, dcol = ph.dcol;
  const uint8_t* ps = sbase + (R - m - syorg + row0) * spitch + (kApron - mc - sxorg) * 3 + col * 24;
  uint8_t* pd = dbase + (R - m - dyorg + row0) * dpitch + (kApron - mc - dxorg) + col * 8;
  const int step_s = drow * spitch + dcol * 24, step_d = drow * dpitch + dcol * 8;
  const int wrap_s = spitch - ux * 24, wrap_d = dpitch - ux * 8;
  for (; idx < n; idx += kTileThreads) {
    gray8<TYPE, LUT>(ps, pd, gk, rgb, lut);
    col += dcol;
    ps += step_s;
    pd += step_d;
    if (col >= ux) {
      col -= ux;
      ps += wrap_s;
      pd += wrap_d;
    }
  }
}

// ---- stencil phases (smem -> smem) ----------------------------------------------------------------

// Bytes [o, o+4) of the 24-byte row window a[0..5].
template <int O>
__device__ __forceinline__ uint32_t window(const uint32_t (&a)[6]) {
  static_assert(O >= 0 && O <= 20, "dp4a window outside the row registers");
  if constexpr ((O & 3) == 0) return a[O >> 2];
  else return __funnelshift_r(a[O >> 2], a[(O >> 2) + 1], (O & 3) * 8);
}

// dp4a window offsets (relative to the output byte) per (channels, ksize): see plan.cpp.
template <int C, int K, int D>
struct DpOff {
  static constexpr int value = C == 1 ? (K == 3 ? -1 : (D == 0 ? -2 : 2))
                                      : (K == 3 ? (D == 0 ? -3 : 0) : (D == 0 ? -6 : (D == 1 ? 0 : 3)));
};
template <int C, int K>
struct DpCount {
  static constexpr int value = C == 1 ? (K == 3 ? 1 : 2) : (K == 3 ? 2 : 3);
};

template <int C, int K, bool SOBEL, int I, int D>
__device__ __forceinline__ void tap_one(const uint32_t (&a)[6], const uint32_t* w,
                                        const uint32_t* w2, int& acc, int& acc2) {
  constexpr int O = I + 8 + DpOff<C, K, D>::value;
  const uint32_t v = window<O>(a);
  acc = dp4a_us(v, w[D], acc);
  if (SOBEL) acc2 = dp4a_us(v, w2[D], acc2);
}
template <int C, int K, bool SOBEL, int I>
__device__ __forceinline__ void tap_px(const uint32_t (&a)[6], const uint32_t* w,
                                       const uint32_t* w2, int& acc, int& acc2) {
  tap_one<C, K, SOBEL, I, 0>(a, w, w2, acc, acc2);
  if constexpr (DpCount<C, K>::value > 1) tap_one<C, K, SOBEL, I, 1>(a, w, w2, acc, acc2);
  if constexpr (DpCount<C, K>::value > 2) tap_one<C, K, SOBEL, I, 2>(a, w, w2, acc, acc2);
}
template <int C, int K, bool SOBEL>
__device__ __forceinline__ void tap_row(const uint32_t (&a)[6], const uint32_t* w,
                                        const uint32_t* w2, int (&acc)[8], int (&acc2)[8]) {
  tap_px<C, K, SOBEL, 0>(a, w, w2, acc[0], acc2[0]);
  tap_px<C, K, SOBEL, 1>(a, w, w2, acc[1], acc2[1]);
  tap_px<C, K, SOBEL, 2>(a, w, w2, acc[2], acc2[2]);
  tap_px<C, K, SOBEL, 3>(a, w, w2, acc[3], acc2[3]);
  tap_px<C, K, SOBEL, 4>(a, w, w2, acc[4], acc2[4]);
  tap_px<C, K, SOBEL, 5>(a, w, w2, acc[5], acc2[5]);
  tap_px<C, K, SOBEL, 6>(a, w, w2, acc[6], acc2[6]);
  tap_px<C, K, SOBEL, 7>(a, w, w2, acc[7], acc2[7]);
}

// One thread: an 8-byte-wide strip, RG output rows, K + RG - 1 input rows each read once (three
// LDS.64), all taps through dp4a. Weights are pulled into registers once per phase; everything
// that is uniform per phase (REF, LUT, expand, scaling) is decided outside the row loop.
template <int C, int K, bool SOBEL>
__device__ void stencil_phase(const TileParams& P, const PhaseDesc& ph, const Tile& T,
                              const uint8_t* sbase, int spitch, uint8_t* dbase, int dpitch,
                              int dxorg, int dyorg, const uint8_t* lut) {
  constexpr int r = K / 2;
  constexpr int RG = SOBEL ? 2 : kRowGroup;
  constexpr int ND = DpCount<C, K>::value;
  const int R = P.halo, m = ph.margin, mc = ph.margin_x > 0 ? kMarginCols : 0;
  const int strips = ph.units;
  const int rows = P.th + 2 * m;
  const bool ref = ph.border == B2I_BORDER_REF;
  const bool scale = (ph.shift | ph.bias) != 0;
  const bool expand = ph.expand != 0;
  const int shift = ph.shift, bias = ph.bias;
  uint32_t w[K][ND], w2[SOBEL ? 3 : 1][ND];
  uint32_t rowmask = 0;  // stencil rows that have a non-zero weight (emboss5: one tap per row)
#pragma unroll
  for (int k = 0; k < K; ++k) {
    uint32_t any = 0;
#pragma unroll
    for (int d = 0; d < ND; ++d) {
      w[k][d] = ph.wpk[k][d];
      any |= ph.wpk[k][d];
      if (SOBEL) w2[k][d] = ph.wpk2[k][d];
    }
    if (any || K == 3) rowmask |= 1u << k;
  }
  const int n_items = ph.n_items;
  for (int item = threadIdx.x; item < n_items; item += kTileThreads) {
    const int g = (item * ph.units_magic) >> 20, strip = item - g * strips;
    const int bc = (kApron - mc) * C + strip * 8;  // plane byte column of the strip
    const int pr0 = R - m + g * RG;                 // plane row of the first output row
    const int nvalid = rows - g * RG;               // output rows of this group inside the region
    int acc[RG][8], acc2[SOBEL ? RG : 1][8];
    uint32_t centre[RG][2];
#pragma unroll
    for (int o = 0; o < RG; ++o)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        acc[o][i] = 0;
        if (SOBEL) acc2[o][i] = 0;
      }
    const uint8_t* srow = sbase + (pr0 - r) * spitch + bc - 8;
#pragma unroll
    for (int j = 0; j < K + RG - 1; ++j) {
      uint32_t a[6];
      const uint2 v0 = *reinterpret_cast<const uint2*>(srow + j * spitch);
      const uint2 v1 = *reinterpret_cast<const uint2*>(srow + j * spitch + 8);
      const uint2 v2 = *reinterpret_cast<const uint2*>(srow + j * spitch + 16);
      a[0] = v0.x; a[1] = v0.y; a[2] = v1.x; a[3] = v1.y; a[4] = v2.x; a[5] = v2.y;
#pragma unroll
      for (int o = 0; o < RG; ++o) {
        constexpr int dummy = 0;
        (void)dummy;
        const int k = j - o;  // stencil row this input row is for output row o (compile time)
        if (k < 0 || k >= K) continue;
        if (k == r) {
          centre[o][0] = a[2];
          centre[o][1] = a[3];
        }
        if (SOBEL) {
          tap_row<C, K, true>(a, w[k], w2[k], acc[o], acc2[o]);
        } else if (K == 3) {
          tap_row<C, K, false>(a, w[k], w[k], acc[o], acc[o]);
        } else if (rowmask & (1u << k)) {
          tap_row<C, K, false>(a, w[k], w[k], acc[o], acc[o]);
        }
      }
    }
    // REF border: only pixels x in [r+1, W-r-1], y in [r+1, H-r-1] are filtered (kernel.cu:83,
    // restricted to its well-defined part); everything else passes the input through.
    uint32_t m0 = 0xFFFFFFFFu, m1 = 0xFFFFFFFFu;
    if (ref) {
      const int gxb0 = (T.tx0 - kApron) * C + bc;
      const int lo = (r + 1) * C - gxb0, hi = (P.W - r) * C - 1 - gxb0;
      if (lo > 0 || hi < 7) {
        m0 = m1 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if (i >= lo && i <= hi) m0 |= 0xFFu << (8 * i);
          if (i + 4 >= lo && i + 4 <= hi) m1 |= 0xFFu << (8 * i);
        }
      }
    }
    const int gy0 = T.ty0 - R + pr0;
    uint32_t res[RG][2];
#pragma unroll
    for (int o = 0; o < RG; ++o) {
      if (SOBEL) {
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[o][i] = abs(acc[o][i]) + abs(acc2[o][i]);
      } else if (scale) {
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[o][i] = (acc[o][i] + bias) >> shift;
      }
      uint32_t r0 = sat_pack4(acc[o][0], acc[o][1], acc[o][2], acc[o][3]);
      uint32_t r1 = sat_pack4(acc[o][4], acc[o][5], acc[o][6], acc[o][7]);
      if (ref) {
        const int gy = gy0 + o;
        const bool row_ok = gy >= r + 1 && gy <= P.H - r - 1;
        const uint32_t mm0 = row_ok ? m0 : 0u, mm1 = row_ok ? m1 : 0u;
        r0 = (r0 & mm0) | (centre[o][0] & ~mm0);
        r1 = (r1 & mm1) | (centre[o][1] & ~mm1);
      }
      res[o][0] = r0;
      res[o][1] = r1;
    }
    if (lut) {
#pragma unroll
      for (int o = 0; o < RG; ++o) {
        res[o][0] = lut4(lut, res[o][0]);
        res[o][1] = lut4(lut, res[o][1]);
      }
    }
    if (expand) {  // C == 1: 8 gray pixels -> 24 bytes of BGR (kernel.cu:210 fused here)
      uint8_t* d = dbase + (pr0 - dyorg) * dpitch + (bc - dxorg) * 3;
#pragma unroll
      for (int o = 0; o < RG; ++o) {
        if (o < nvalid) {
          const uint32_t r0 = res[o][0], r1 = res[o][1];
          uint2* q = reinterpret_cast<uint2*>(d + o * dpitch);
          q[0] = make_uint2(__byte_perm(r0, 0, 0x1000), __byte_perm(r0, 0, 0x2211));
          q[1] = make_uint2(__byte_perm(r0, 0, 0x3332), __byte_perm(r1, 0, 0x1000));
          q[2] = make_uint2(__byte_perm(r1, 0, 0x2211), __byte_perm(r1, 0, 0x3332));
        }
      }
    } else {
      uint8_t* d = dbase + (pr0 - dyorg) * dpitch + (bc - dxorg * C);
#pragma unroll
      for (int o = 0; o < RG; ++o)
        if (o < nvalid) *reinterpret_cast<uint2*>(d + o * dpitch) = make_uint2(res[o][0], res[o][1]);
    }
  }
}

// ---- separable stencils (smem -> smem) --------------------------------------------------------------
// A k x k kernel that factors into a column vector times a row vector (the 5x5 and 3x3 Gaussians)
// costs 2k taps per byte instead of k*k. The horizontal pass widens bytes to 16-bit lanes
// (word m of the source -> two words: the even bytes and the odd bytes of word m) and sums the row
// taps with packed IMADs; the vertical pass sums whole lane words down the rows, rounds, shifts
// and narrows back to bytes. Exact: lanes never exceed 255 * sum(w) < 65536.

// Bytes [O, O+4) of the 24-byte row window a[0..5].
template <int O>
__device__ __forceinline__ uint32_t win24(const uint32_t (&a)[6]) {
  static_assert(O >= 0 && O <= 20, "window outside the row registers");
  if constexpr ((O & 3) == 0) return a[O >> 2];
  else return __funnelshift_r(a[O >> 2], a[(O >> 2) + 1], (O & 3) * 8);
}
__device__ __forceinline__ uint32_t even_lanes(uint32_t w) { return __byte_perm(w, 0u, 0x4240); }
__device__ __forceinline__ uint32_t odd_lanes(uint32_t w) { return __byte_perm(w, 0u, 0x4341); }

// One source word (4 bytes at window offset 8 + 4*WORD) -> its even-lane and odd-lane sums.
template <int C, int K, int WORD>
__device__ __forceinline__ void seph_word(const uint32_t (&a)[6], const int (&w)[5], uint32_t& e, uint32_t& o) {
  constexpr int B = 8 + 4 * WORD;
  constexpr int r = K / 2;
  e = 0;
  o = 0;
  if constexpr (K == 5) {
    const uint32_t s0 = win24<B - 2 * C>(a), s4 = win24<B + 2 * C>(a);
    e = even_lanes(s0) * w[0] + even_lanes(s4) * w[4];
    o = odd_lanes(s0) * w[0] + odd_lanes(s4) * w[4];
  }
  const uint32_t s1 = win24<B - C>(a), s2 = win24<B>(a), s3 = win24<B + C>(a);
  e += even_lanes(s1) * w[r - 1] + even_lanes(s2) * w[r] + even_lanes(s3) * w[r + 1];
  o += odd_lanes(s1) * w[r - 1] + odd_lanes(s2) * w[r] + odd_lanes(s3) * w[r + 1];
}

template <int C, int K>
__device__ void seph_phase(const TileParams& P, const PhaseDesc& ph, const uint8_t* sbase, int spitch,
                           uint8_t* dbase, int dpitch) {
  const int R = P.halo, m = ph.margin, mc = ph.margin_x > 0 ? kMarginCols : 0;
  const int strips = ph.units;
  int w[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) w[i] = ph.sep[i];
  const int n = ph.n_items;
  int idx = threadIdx.x;
  const int row0 = (idx * ph.units_magic) >> 20;
  int col = idx - row0 * strips;
  const int drow = ph.drow, dcol = ph.dcol;
  const uint8_t* ps = sbase + (R - m + row0) * spitch + (kApron - mc) * C + col * 8 - 8;
  uint8_t* pd = dbase + (R - m + row0) * dpitch + 2 * ((kApron - mc) * C + col * 8);
  const int step_s = drow * spitch + dcol * 8, step_d = drow * dpitch + dcol * 16;
  const int wrap_s = spitch - strips * 8, wrap_d = dpitch - strips * 16;
  for (; idx < n; idx += kTileThreads) {
    uint32_t a[6];
    const uint2 v0 = *reinterpret_cast<const uint2*>(ps), v1 = *reinterpret_cast<const uint2*>(ps + 8),
                v2 = *reinterpret_cast<const uint2*>(ps + 16);
    a[0] = v0.x; a[1] = v0.y; a[2] = v1.x; a[3] = v1.y; a[4] = v2.x; a[5] = v2.y;
    uint4 out;
    seph_word<C, K, 0>(a, w, out.x, out.y);
    seph_word<C, K, 1>(a, w, out.z, out.w);
    *reinterpret_cast<uint4*>(pd) = out;
    col += dcol;
    ps += step_s;
    pd += step_d;
    if (col >= strips) {
      col -= strips;
      ps += wrap_s;
      pd += wrap_d;
    }
  }
}

// Vertical pass: one thread = an 8-byte strip (four lane words), RG output rows, K + RG - 1 input rows.
template <int K>
__device__ void sepv_phase(const TileParams& P, const PhaseDesc& ph, const uint8_t* sbase, int spitch,
                           uint8_t* dbase, int dpitch, int dxorg, int dyorg, int C, const uint8_t* lut) {
  constexpr int r = K / 2;
  constexpr int RG = kRowGroup;
  const int R = P.halo, m = ph.margin, mc = ph.margin_x > 0 ? kMarginCols : 0;
  const int strips = ph.units;
  const int rows = P.th + 2 * m;
  int w[5];
#pragma unroll
  for (int i = 0; i < 5; ++i) w[i] = ph.sep[i];
  const uint32_t bias2 = (uint32_t)ph.bias * 0x00010001u;
  const int shift = ph.shift;
  const bool expand = ph.expand != 0;
  const int n_items = ph.n_items;
  for (int item = threadIdx.x; item < n_items; item += kTileThreads) {
    const int g = (item * ph.units_magic) >> 20, strip = item - g * strips;
    const int bc = (kApron - mc) * C + strip * 8;
    const int pr0 = R - m + g * RG;
    const int nvalid = rows - g * RG;
    uint32_t acc[RG][4];
#pragma unroll
    for (int o = 0; o < RG; ++o)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[o][q] = 0;
    const uint8_t* srow = sbase + (pr0 - r) * spitch + 2 * bc;
#pragma unroll
    for (int j = 0; j < K + RG - 1; ++j) {
      const uint4 h = *reinterpret_cast<const uint4*>(srow + j * spitch);
#pragma unroll
      for (int o = 0; o < RG; ++o) {
        const int k = j - o;
        if (k < 0 || k >= K) continue;
        const uint32_t wk = (uint32_t)w[k];
        acc[o][0] += h.x * wk;
        acc[o][1] += h.y * wk;
        acc[o][2] += h.z * wk;
        acc[o][3] += h.w * wk;
      }
    }
    uint32_t res[RG][2];
#pragma unroll
    for (int o = 0; o < RG; ++o) {
      // lanes -> bytes: ((lane + bias) >> shift) of the even word gives bytes 0 and 2, odd word 1 and 3
      uint32_t e0 = acc[o][0] + bias2, o0 = acc[o][1] + bias2, e1 = acc[o][2] + bias2, o1 = acc[o][3] + bias2;
      if (shift == 8) {
        res[o][0] = __byte_perm(e0, o0, 0x7351);
        res[o][1] = __byte_perm(e1, o1, 0x7351);
      } else {
        e0 = (e0 >> shift) & 0x00FF00FFu; o0 = (o0 >> shift) & 0x00FF00FFu;
        e1 = (e1 >> shift) & 0x00FF00FFu; o1 = (o1 >> shift) & 0x00FF00FFu;
        res[o][0] = e0 | (o0 << 8);
        res[o][1] = e1 | (o1 << 8);
      }
    }
    if (lut) {
#pragma unroll
      for (int o = 0; o < RG; ++o) {
        res[o][0] = lut4(lut, res[o][0]);
        res[o][1] = lut4(lut, res[o][1]);
      }
    }
    if (expand) {
      uint8_t* d = dbase + (pr0 - dyorg) * dpitch + (bc - dxorg) * 3;
#pragma unroll
      for (int o = 0; o < RG; ++o) {
        if (o < nvalid) {
          const uint32_t r0 = res[o][0], r1 = res[o][1];
          uint2* q = reinterpret_cast<uint2*>(d + o * dpitch);
          q[0] = make_uint2(__byte_perm(r0, 0, 0x1000), __byte_perm(r0, 0, 0x2211));
          q[1] = make_uint2(__byte_perm(r0, 0, 0x3332), __byte_perm(r1, 0, 0x1000));
          q[2] = make_uint2(__byte_perm(r1, 0, 0x2211), __byte_perm(r1, 0, 0x3332));
        }
      }
    } else {
      uint8_t* d = dbase + (pr0 - dyorg) * dpitch + (bc - dxorg * C);
#pragma unroll
      for (int o = 0; o < RG; ++o)
        if (o < nvalid) *reinterpret_cast<uint2*>(d + o * dpitch) = make_uint2(res[o][0], res[o][1]);
    }
  }
}

// ---- tile output --------------------------------------------------------------------------------

// Fallback for destinations TMA cannot describe. Rows outside [row_lo, row_hi) and pixels at
// or beyond W are dropped, like TMA's clipping.
__device__ void store_tile_generic(const TileParams& P, const Tile& T, const uint8_t* stage,
                                   uint64_t dst_ptr, int64_t dst_pitch, int dst_y_begin,
                                   int row_lo, int row_hi) {
  const int pitch = P.buf[BUF_STAGE].pitch, words = pitch >> 2;
  const int64_t row_bytes = (int64_t)P.W * P.cout;
  const int64_t xb0 = (int64_t)T.tx0 * P.cout;
  uint8_t* dst = reinterpret_cast<uint8_t*>(dst_ptr);
  for (int idx = threadIdx.x; idx < words * P.th; idx += kTileThreads) {
    const int rr = idx / words, w = idx - rr * words;
    const int gy = T.ty0 + rr;
    const int64_t xb = xb0 + 4 * w;
    if (gy < row_lo || gy >= row_hi || xb >= row_bytes) continue;
    const uint32_t v = *reinterpret_cast<const uint32_t*>(stage + (size_t)rr * pitch + 4 * w);
    uint8_t* p = dst + (int64_t)(gy - dst_y_begin) * dst_pitch + xb;
    if (xb + 3 < row_bytes && (reinterpret_cast<uintptr_t>(p) & 3) == 0) {
      *reinterpret_cast<uint32_t*>(p) = v;
    } else {
      for (int b = 0; b < 4 && xb + b < row_bytes; ++b) p[b] = (uint8_t)(v >> (8 * b));
    }
  }
}

// Does tile T take part in a flag this launch publishes? Only those tiles pay for the full
// store completion, the proxy fence and the system-scope fence; all other tiles just wait until
// their TMA store has read the staging buffer. (The first version fenced after every tile: a
// membar.sys per tile on thread 0 cost ~35 us per pass at N = 2.)
__device__ __forceinline__ bool tile_publishes(const TileParams& P, const Tile& T) {
  return (P.signal_up_flag && T.ty0 < P.y0 + P.push_rows) ||
         (P.signal_dn_flag && T.ty0 + P.th > P.y1 - P.push_rows);
}

// Thread 0, after tile T's stores (own rows and halo pushes) are complete and fenced: count the
// tile in; the last tile of each group releases the corresponding flag and re-arms the counter.
__device__ __forceinline__ void publish_flags(const TileParams& P, const Tile& T, uint32_t* counters,
                                              int n_push_up, int n_push_dn) {
  if (P.signal_up_flag && T.ty0 < P.y0 + P.push_rows)
    if (atomicAdd(&counters[0], 1u) + 1 == (uint32_t)n_push_up) {
      atomicExch(&counters[0], 0u);
      st_release_sys(reinterpret_cast<uint32_t*>(P.signal_up_flag), P.signal_epoch);
    }
  if (P.signal_dn_flag && T.ty0 + P.th > P.y1 - P.push_rows)
    if (atomicAdd(&counters[1], 1u) + 1 == (uint32_t)n_push_dn) {
      atomicExch(&counters[1], 0u);
      st_release_sys(reinterpret_cast<uint32_t*>(P.signal_dn_flag), P.signal_epoch);
    }
}

}  // namespace

// ---- the kernel -----------------------------------------------------------------------------------

__global__ void __launch_bounds__(kTileThreads, 2)
tile_chain_kernel(const __grid_constant__ TileParams P, const __grid_constant__ CUtensorMap src_map,
                  const __grid_constant__ CUtensorMap dst_map,
                  const __grid_constant__ CUtensorMap push_up_map,
                  const __grid_constant__ CUtensorMap push_dn_map) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + P.bar_off);
  uint8_t* luts = smem + P.lut_off;
  const int tid = threadIdx.x;

  if (tid == 0) {
    mbar_init(&full_bar[0], 1);
    mbar_init(&full_bar[1], 1);
    fence_mbar_init();
    if (P.use_tma_load && !(P.debug & 1)) tma_prefetch_desc(&src_map);
    if (P.use_tma_store && !(P.debug & 1)) tma_prefetch_desc(&dst_map);
  }
  for (int i = tid; i < kMaxLuts * 256 / 4; i += kTileThreads)
    reinterpret_cast<uint32_t*>(luts)[i] = reinterpret_cast<const uint32_t*>(&P.luts[0][0])[i];
  __syncthreads();

  uint32_t* counters = reinterpret_cast<uint32_t*>(P.counters);
  uint32_t* status = reinterpret_cast<uint32_t*>(P.status);
  // tiles that take part in a halo push: tile rows starting inside the first push_rows rows
  // (up), tile rows reaching into the last push_rows rows (down). The last of them to finish
  // releases the neighbour's flag, so these counts must be exact.
  const int band_rows = P.y1 - P.y0;
  const int n_push_up = P.push_rows > 0 ? ((P.push_rows + P.th - 1) / P.th) * P.tiles_x : 0;
  const int n_push_dn = P.push_rows > 0 ? (P.tiles_y - (band_rows - P.push_rows) / P.th) * P.tiles_x : 0;

  // Thread 0: make sure the neighbours' halo rows for this tile have landed in our ghost rows
  // (acquire their flag), then start the TMA load of the tile into stage `s`.
  auto issue_load = [&](int t, int s) {
    const Tile T = tile_of(P, t);
    if (P.wait_up_flag && T.ty0 - P.halo < P.y0) {
      if (!wait_flag_ge(reinterpret_cast<const uint32_t*>(P.wait_up_flag), P.wait_epoch, P.timeout_ns))
        atomicExch(status, 1u);
      asm volatile("fence.proxy.async;" ::: "memory");
    }
    if (P.wait_dn_flag && T.ty0 + P.th + P.halo > P.y1) {
      if (!wait_flag_ge(reinterpret_cast<const uint32_t*>(P.wait_dn_flag), P.wait_epoch, P.timeout_ns))
        atomicExch(status, 2u);
      asm volatile("fence.proxy.async;" ::: "memory");
    }
    if (P.use_tma_load) {
      fence_proxy_async_smem();  // earlier generic-proxy writes to this stage (border fix-up)
      mbar_arrive_expect_tx(&full_bar[s], (uint32_t)P.raw_tx_bytes);
      tma_load_2d(smem + P.buf[BUF_RAW].off + s * P.raw_stage_bytes, &src_map,
                  ((T.tx0 - kApron) * P.cin) >> 2, T.ty0 - P.halo - P.src_y_begin, &full_bar[s]);
    }
  };

  int it = 0;
  bool store_pending = false;
  Tile prevT = tile_of(P, 0);
  if (tid == 0 && (int)blockIdx.x < P.n_tiles) issue_load(blockIdx.x, 0);
  if (!P.use_tma_load) __syncthreads();  // the flag acquire of thread 0 orders everyone's loads

  const bool prefetch = P.raw_stages == 2;
  for (int t = blockIdx.x; t < P.n_tiles; t += gridDim.x, ++it) {
    const int s = prefetch ? (it & 1) : 0;
    const Tile T = tile_of(P, t);
    uint8_t* raw = smem + P.buf[BUF_RAW].off + s * P.raw_stage_bytes;
    const int tn = t + gridDim.x;
    if (prefetch) {
      if (tid == 0 && tn < P.n_tiles) issue_load(tn, s ^ 1);  // stage s^1 was released by the
                                                               // barrier ending the previous tile
    } else if (tid == 0 && it > 0) {
      issue_load(t, 0);  // single stage: the tile is loaded when its turn comes
    }
    if (P.use_tma_load) {
      mbar_wait(&full_bar[s], prefetch ? ((it >> 1) & 1) : (it & 1));
    } else {
      load_tile_generic(P, T, raw);
      __syncthreads();
    }
    if (P.raw_fix_border >= 0 && T.edge) {
      fix_border(P, T, raw, P.buf[BUF_RAW].pitch, P.cin, P.halo, P.raw_fix_border);
      __syncthreads();
    }

    for (int k = 0; k < P.n_phases; ++k) {
      const PhaseDesc& ph = P.ph[k];
      const BufDesc& sb = P.buf[ph.src];
      const BufDesc& db = P.buf[ph.dst];
      uint8_t* sbase = ph.src == BUF_RAW ? raw : smem + sb.off;
      uint8_t* dbase = smem + db.off;
      const uint8_t* lut = ph.lut >= 0 ? luts + 256 * ph.lut : nullptr;
      if (ph.dst == BUF_STAGE && store_pending) {
        // the previous tile's TMA store must have drained the staging tile before it is reused;
        // waiting for full completion here also lets thread 0 publish that tile's flags
        if (tid == 0) {
          if (tile_publishes(P, prevT)) {
            tma_store_wait_all();
            asm volatile("fence.proxy.async;" ::: "memory");
            __threadfence_system();
            publish_flags(P, prevT, counters, n_push_up, n_push_dn);
          } else {
            tma_store_wait_read();
          }
        }
        __syncthreads();
        store_pending = false;
      }
      switch (ph.type) {
        case PH_GRAY_REF:
          if (lut) gray_phase<PH_GRAY_REF, true>(P, ph, sbase, sb.pitch, sb.xorg, sb.yorg, dbase, db.pitch, db.xorg, db.yorg, lut);
          else gray_phase<PH_GRAY_REF, false>(P, ph, sbase, sb.pitch, sb.xorg, sb.yorg, dbase, db.pitch, db.xorg, db.yorg, lut);
          break;
        case PH_GRAY_CV:
          if (lut) gray_phase<PH_GRAY_CV, true>(P, ph, sbase, sb.pitch, sb.xorg, sb.yorg, dbase, db.pitch, db.xorg, db.yorg, lut);
          else gray_phase<PH_GRAY_CV, false>(P, ph, sbase, sb.pitch, sb.xorg, sb.yorg, dbase, db.pitch, db.xorg, db.yorg, lut);
          break;
        case PH_COPY:
          pw_phase<PH_COPY>(P, ph, sbase, sb.pitch, sb.xorg, sb.yorg, dbase, db.pitch, db.xorg, db.yorg, lut);
          break;
        case PH_EXPAND:
          pw_phase<PH_EXPAND>(P, ph, sbase, sb.pitch, sb.xorg, sb.yorg, dbase, db.pitch, db.xorg, db.yorg, lut);
          break;
        case PH_SEPH:
          if (ph.cin == 1) {
            if (ph.ksize == 3) seph_phase<1, 3>(P, ph, sbase, sb.pitch, dbase, db.pitch);
            else seph_phase<1, 5>(P, ph, sbase, sb.pitch, dbase, db.pitch);
          } else {
            if (ph.ksize == 3) seph_phase<3, 3>(P, ph, sbase, sb.pitch, dbase, db.pitch);
            else seph_phase<3, 5>(P, ph, sbase, sb.pitch, dbase, db.pitch);
          }
          break;
        case PH_SEPV:
          if (ph.ksize == 3) sepv_phase<3>(P, ph, sbase, sb.pitch, dbase, db.pitch, db.xorg, db.yorg, ph.cin, lut);
          else sepv_phase<5>(P, ph, sbase, sb.pitch, dbase, db.pitch, db.xorg, db.yorg, ph.cin, lut);
          break;
        case PH_SOBEL:
          if (ph.cin == 1) stencil_phase<1, 3, true>(P, ph, T, sbase, sb.pitch, dbase, db.pitch, db.xorg, db.yorg, lut);
          else stencil_phase<3, 3, true>(P, ph, T, sbase, sb.pitch, dbase, db.pitch, db.xorg, db.yorg, lut);
          break;
        default:
          if (ph.cin == 1) {
            if (ph.ksize == 3) stencil_phase<1, 3, false>(P, ph, T, sbase, sb.pitch, dbase, db.pitch, db.xorg, db.yorg, lut);
            else stencil_phase<1, 5, false>(P, ph, T, sbase, sb.pitch, dbase, db.pitch, db.xorg, db.yorg, lut);
          } else {
            if (ph.ksize == 3) stencil_phase<3, 3, false>(P, ph, T, sbase, sb.pitch, dbase, db.pitch, db.xorg, db.yorg, lut);
            else stencil_phase<3, 5, false>(P, ph, T, sbase, sb.pitch, dbase, db.pitch, db.xorg, db.yorg, lut);
          }
          break;
      }
      if (ph.dst == BUF_STAGE && P.use_tma_store) fence_proxy_async_smem();
      __syncthreads();
      if (ph.fix_border >= 0 && T.edge) {
        fix_border(P, T, dbase, db.pitch, ph.cout, max(ph.margin, ph.margin_x), ph.fix_border);
        __syncthreads();
      }
    }

    // ---- tile output: own rows, plus the halo rows pushed into the neighbours' ghost rows ----
    const uint8_t* stage = smem + P.buf[BUF_STAGE].off;
    const bool push_up = P.push_up_ptr && T.ty0 < P.y0 + P.push_rows;
    const bool push_dn = P.push_dn_ptr && T.ty0 + P.th > P.y1 - P.push_rows;
    if (P.use_tma_store) {
      if (tid == 0) {
        const int xw = (T.tx0 * P.cout) >> 2;
        tma_store_2d(&dst_map, xw, T.ty0 - P.dst_y_begin, stage);
        // TMA stores clip at the far edge of the tensor only: a negative coordinate faults
        // (measured, scripts/tma_probe_store.cu). Halo rows are therefore pushed with one-row
        // boxes, one per staged row that belongs to the halo; only boundary tiles do this.
        const int spitch = P.buf[BUF_STAGE].pitch;
        if (push_up) {
          const int ra = max(T.ty0, P.y0), rb = min(T.ty0 + P.th, P.y0 + P.push_rows);
          for (int r = ra; r < rb; ++r)
            tma_store_2d(&push_up_map, xw, r - P.push_up_y_begin, stage + (r - T.ty0) * spitch);
        }
        if (push_dn) {
          const int ra = max(T.ty0, P.y1 - P.push_rows), rb = min(T.ty0 + P.th, P.y1);
          for (int r = ra; r < rb; ++r)
            tma_store_2d(&push_dn_map, xw, r - P.push_dn_y_begin, stage + (r - T.ty0) * spitch);
        }
        tma_store_commit();
      }
    } else {
      store_tile_generic(P, T, stage, P.dst_ptr, P.dst_pitch, P.dst_y_begin,
                         max(P.y0, P.dst_y_begin), min(P.y1, P.dst_y_end));
      if (push_up)
        store_tile_generic(P, T, stage, P.push_up_ptr, P.push_pitch, P.push_up_y_begin, P.y0,
                           P.y0 + P.push_rows);
      if (push_dn)
        store_tile_generic(P, T, stage, P.push_dn_ptr, P.push_pitch, P.push_dn_y_begin,
                           P.y1 - P.push_rows, P.y1);
    }
    if (!P.use_tma_store && tile_publishes(P, T))
      __threadfence_system();  // plain stores: make them visible before thread 0 publishes flags
    store_pending = true;
    prevT = T;
    // every thread is past its reads of `raw` (stage s) here: the next iteration may refill it
    __syncthreads();
  }

  // drain: the last tile's stores, then its flags
  if (tid == 0 && store_pending) {
    if (P.use_tma_store) tma_store_wait_all();
    if (tile_publishes(P, prevT)) {
      asm volatile("fence.proxy.async;" ::: "memory");
      __threadfence_system();
      publish_flags(P, prevT, counters, n_push_up, n_push_dn);
    }
  }
}

// The kernel may use every byte of opt-in shared memory; set once per device (setting it per
// launch size would let a small launch lower the limit under a large one).
static cudaError_t allow_max_smem() {
  static bool done[64] = {false};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 64 && done[dev]) return cudaSuccess;
  int optin = 0;
  e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(tile_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
  if (e != cudaSuccess) return e;
  if (dev < 64) done[dev] = true;
  return cudaSuccess;
}

cudaError_t launch_tile_chain(const TileParams& P, const CUtensorMap& src_map,
                              const CUtensorMap& dst_map, const CUtensorMap& push_up_map,
                              const CUtensorMap& push_dn_map, int grid, cudaStream_t stream) {
  cudaError_t e = allow_max_smem();
  if (e != cudaSuccess) return e;
  tile_chain_kernel<<<grid, kTileThreads, P.smem_bytes, stream>>>(P, src_map, dst_map, push_up_map,
                                                                  push_dn_map);
  return cudaGetLastError();
}

cudaError_t tile_chain_occupancy(int smem_bytes, int* blocks_per_sm) {
  cudaError_t e = allow_max_smem();
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, tile_chain_kernel,
                                                       kTileThreads, smem_bytes);
}

}  // namespace b2i
