// The tensor-core stencil kernel: a k x k integer correlation on 8-bit rows (k = 3 or 5; the
// Gaussians, emboss, sharpen, Laplace, generic filter3 / filter5, and Sobel as two of them) as a
// banded GEMM on the 5th-generation tensor cores (tcgen05.mma kind::i8, accumulators in TMEM).
//
// The packed-integer sweeps (sweep5.cu, sweep_sobel.cu) spend ~11 CUDA-core instructions per byte
// and are bound by instruction issue at 0.55 of the HBM roofline. A stencil is linear, so the
// multiply-adds can go to the tensor pipe instead:
//
//   out[y][o] = sum_dy sum_t w[dy][t] in[y + dy - R][o + t * cs]
//             = sum_dy ( A_dy x B_dy )[y][o],   A_dy = the image rows shifted down by dy (128 x K, u8),
//                                               B_dy = a band matrix (K x N, s8) holding w[dy][.]
//
//   * an input slab of 132 rows x 128 bytes arrives by ONE TMA box in the 128-byte-swizzled
//     K-major layout the tensor core reads (rows = M, bytes along a row = K). The vertical taps
//     cost nothing: A_dy is the same slab with the descriptor's start address moved down dy rows
//     (the swizzle is a function of the absolute shared-memory address, so a row-shifted start
//     needs no base offset in the descriptor: measured, scripts/mma_check.py);
//   * a K-step (32 input bytes) touches the 64 result columns [k0 - 16, k0 + 48): 2R+1 MMAs of
//     128 x 64 x 32 with the same 64 x 32 band matrices for every K-step of the image (built in
//     shared memory from the 25 weights at kernel start). Image borders are separate band
//     matrices with the mirrored / replicated taps folded in (left edge, right edge), and
//     copied rows in the slab (top, bottom);
//   * the accumulators live in TMEM as a ring of four 128-column chunks (two for Sobel, which
//     keeps gx and gy side by side). A CTA sweeps along a 128-row band: the MMA warp runs up to
//     three chunks ahead of the epilogue warps, which read a finished chunk with tcgen05.ld,
//     narrow it (three PRMTs per four bytes when the taps sum to 256; the rounding bias is the value
//     the ring is reset to), stage it swizzled and hand it to a TMA store;
//   * warp 0 feeds a six-slab ring (mbarriers, one elected lane), warp 1 issues the MMAs (one
//     elected lane; tcgen05.commit frees slabs and publishes chunks), warps 2-5 are the epilogue.
//     One CTA per SM, persistent: every CTA takes a contiguous run of slabs.
//
// Measured on B200 (8K RGB; scripts/kbench.py, scripts/mma_timeline.py, profiles/mma_notes.md): without
// the MMAs the kernel moves its 210 MB at 6.5 TB/s (0.77 us per slab and SM); an M = 128 MMA with
// both operands in shared memory takes ~110 cycles whatever its N (16, 32 and 64 time the same: the
// 128 operand rows are the cost), so 12 MMAs per slab (3 x 3) hide behind the memory traffic and 20
// (5 x 5) do not: 3 x 3 stencils 47 us, 5 x 5 stencils 55 us, any weights, any channel count. Staging
// A through TMEM (tcgen05.cp, then the MMA reads A from TMEM) was tried and is slower (92 us).
//
// Exact: u8 x s8 products accumulate in s32. Reference: kernel.cu:64-94 is this correlation with
// one thread per pixel and seven byte loads per output; kern.cpp:75 is cv::filter2D.
#include "b200img.h"
#include "device_utils.cuh"
#include "launch_pdl.hpp"
#include "stencil_mma.hpp"

namespace b2i {
using namespace dev;

namespace {

// ---- tcgen05 wrappers ------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc_512(uint32_t* slot) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(slot)) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free_512(uint32_t taddr) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(taddr) : "memory");
}
// All MMAs this thread has issued so far: their completion arrives on the mbarrier.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// One lane of the (converged) warp. The MMA warp runs its issue loop on all 32 lanes so that every
// operand is a warp-uniform value (uniform registers, no per-lane broadcast loops around UTCIMMA);
// only the tcgen05 instructions themselves are predicated on the elected lane.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Shared-memory matrix descriptor (Blackwell format: version 1 in bits 46-47).
//   layout 2 = 128-byte swizzle, K-major: rows 128 bytes apart, SBO between 8-row groups
//   layout 0 = no swizzle, K-major: 8 x 16-byte core matrices, LBO between the two K halves, SBO between 8-row groups
// (the 64-bit descriptor: start address >> 4 in bits 0-13, LBO >> 4 in bits 16-29, SBO >> 4 in bits 32-45,
// version 1 in bits 46-47, base offset in bits 49-51 -- left 0, see the header comment -- layout in bits 61-63)

// D[tmem] += A[smem] x B[smem], A unsigned bytes, B signed bytes, D s32, M = 128.
// The descriptors travel as (lo, hi) words: the issuing thread only ever adds small
// constants to the lo words (start address field), the hi words are compile-time constants.
__device__ __forceinline__ void umma_i8_w(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                          uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], da, db, %5, {%7, %7, %7, %7}, p;\n\t}"
      ::"r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(1u), "r"(0u)
      : "memory");
}
constexpr uint32_t kADescHi = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO 1024, version 1, 128-byte swizzle
constexpr uint32_t kBDescHi = (256u >> 4) | (1u << 14);                // SBO 256, version 1, no swizzle
__device__ __forceinline__ uint32_t a_desc_lo(uint32_t saddr) { return ((saddr & 0x3FFFFu) >> 4) | (1u << 16); }
__device__ __forceinline__ uint32_t b_desc_lo(uint32_t saddr) { return ((saddr & 0x3FFFFu) >> 4) | ((128u >> 4) << 16); }

// c_format s32 (2 << 4), a_format u8 (0 << 7), b_format s8 (1 << 10), both K-major, M = 128 (8 << 24)
__device__ __forceinline__ uint32_t idesc_i8(int n) {
  return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | (8u << 24);
}

// 32 lanes x 32 columns of s32: thread = TMEM lane (result row), v[j] = column col + j.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// The same value into 32 lanes x 32 columns (resets a finished part of the accumulator ring).
__device__ __forceinline__ void tmem_fill32(uint32_t taddr, const uint32_t (&b)[8]) {
  // x8 stores of eight registers that stay live for the whole kernel: no register shuffling per store
#pragma unroll
  for (int j = 0; j < 4; ++j)
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr + 8 * j),
                 "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7])
                 : "memory");
}

// Debug timeline (P.dbg_ptr != 0): globaltimer stamps of blocks 0 and 1, slot i of block b at dbg_ptr[b * 16 + i].
__device__ __forceinline__ void stamp(const StencilMmaParams& P, int slot) {
  if (P.dbg_ptr && blockIdx.x < 2) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    reinterpret_cast<unsigned long long*>(P.dbg_ptr)[blockIdx.x * 16 + slot] = t;
  }
}

// ---- geometry ----------------------------------------------------------------------------------------
// Byte offset of element (n, k) of a 64 x 32 band matrix in the no-swizzle K-major layout.
__device__ __forceinline__ int b_off(int n, int k) { return (n >> 3) * 256 + (k >> 4) * 128 + (n & 7) * 16 + (k & 15); }

__device__ __forceinline__ int fold_px(int p, int W, int border) {  // mirrored / replicated pixel, or -1
  if (p >= 0 && p < W) return p;
  if (border == B2I_BORDER_REFLECT101) return p < 0 ? -p : 2 * (W - 1) - p;
  if (border == B2I_BORDER_REPLICATE) return p < 0 ? 0 : W - 1;
  return -1;
}
__device__ __forceinline__ int fold_row(int g, int H, int border) { return fold_px(g, H, border); }

struct Run {          // slabs [s0, s1) of band b
  int b, s0, s1;
};

// The CTA's next run at global slab index g (band-major), or false when its share is done.
__device__ __forceinline__ bool next_run(int& g, int g_hi, int n_slabs, Run& r) {
  if (g >= g_hi) return false;
  r.b = g / n_slabs;
  r.s0 = g - r.b * n_slabs;
  r.s1 = min(n_slabs, r.s0 + (g_hi - g));
  g += r.s1 - r.s0;
  return true;
}
// First / last slab a run loads: one K-step of the slab before and of the slab after feed its edge columns.
__device__ __forceinline__ int run_first_slab(const Run& r) { return r.s0 > 0 ? r.s0 - 1 : 0; }
__device__ __forceinline__ int run_last_slab(const Run& r, int n_slabs) { return r.s1 < n_slabs ? r.s1 : r.s1 - 1; }

}  // namespace

template <int NDY, int NACC>
__global__ void __launch_bounds__(kSmThreads, 1)
stencil_mma_kernel(const __grid_constant__ StencilMmaParams P, const __grid_constant__ CUtensorMap src_map,
                   const __grid_constant__ CUtensorMap dst_map) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment for the swizzled buffers
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ __align__(8) uint64_t full_slab[kSmStages], empty_slab[kSmStages], chunk_full[4], chunk_empty[4];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) int8_t wtab[2][kSmMaxDy][112];  // weight of input byte i for result byte o, at [d = i - o + 48]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();   // the next launch in the stream may set itself up behind this one
  if (threadIdx.x == 0) stamp(P, 0);
  constexpr int ndy = NDY, nacc = NACC;      // NDY = 2R + 1; NACC = 2: Sobel's gx and gy
  constexpr int ring = kSmRing / nacc;      // columns per accumulator
  constexpr int ring_chunks = ring / 128;   // 4 or 2
  const int W = P.W3 / P.cs;

  uint8_t* slabs = smem;
  uint8_t* bmat = slabs + kSmStages * kSmSlabBytes;                 // [variant][acc][dy] 2 KB each
  uint8_t* stage_out = bmat + kSmVariants * 2 * 3 * kSmBBytes;      // 48 KB covers 1 x 5 and 2 x 3 matrices per variant
  constexpr int bvar_bytes = nacc * ndy * kSmBBytes;

  // ---- set-up: barriers, TMEM, the band matrices ------------------------------------------------------
  if (threadIdx.x == 0) {
    for (int i = 0; i < kSmStages; ++i) {
      mbar_init(&full_slab[i], 1);
      mbar_init(&empty_slab[i], 1);
    }
    for (int i = 0; i < 4; ++i) {
      mbar_init(&chunk_full[i], 1);
      mbar_init(&chunk_empty[i], 4);
    }
    fence_mbar_init();
    tma_prefetch_desc(&src_map);
    tma_prefetch_desc(&dst_map);
  }
  if (warp == 1) tmem_alloc_512(&tmem_slot);
  for (int i = threadIdx.x; i < 2 * kSmMaxDy * 112; i += kSmThreads) {
    const int a = i / (kSmMaxDy * 112), dy = (i / 112) % kSmMaxDy, d = i % 112 - 48;
    int8_t v = 0;
    if (a < nacc && dy < ndy && d % P.cs == 0 && abs(d / P.cs) <= P.R) v = P.w[a][dy][d / P.cs + P.R];
    wtab[a][dy][i % 112] = v;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) stamp(P, 1);

  // work split: global slab index g = band * n_slabs + slab
  const long long G = (long long)P.n_bands * P.n_slabs;
  const int g_lo = (int)(G * blockIdx.x / gridDim.x), g_hi = (int)(G * (blockIdx.x + 1) / gridDim.x);
  const int kl = (P.W3 - 1) / kSmKStep;  // last K-step with input bytes inside the image

  if (warp == 0) {
    // ===== TMA producer =====================================================================
    if (lane == 0) {
      pdl_wait();   // the launch before this one has completed: its output may be this one's input. The
                    // other warps build the band matrices meanwhile (3 us that hide behind that launch).
      uint32_t n = 0;
      int g = g_lo;
      Run r;
      while (next_run(g, g_hi, P.n_slabs, r)) {
        const int yb = P.y0 + r.b * kSmRows;
        for (int j = run_first_slab(r); j <= run_last_slab(r, P.n_slabs); ++j, ++n) {
          const uint32_t st = n % kSmStages;
          mbar_wait(&empty_slab[st], ((n / kSmStages) & 1) ^ 1);
          mbar_arrive_expect_tx(&full_slab[st], (kSmRows + 2 * P.R) * kSmSlab);
          tma_load_2d(slabs + st * kSmSlabBytes, &src_map, j * kSmSlab, yb - P.R - P.src_y_begin, &full_slab[st]);
        }
      }
      stamp(P, 8);
    }
  } else {
    // ---- band matrices: every variant starts as the interior one; 160 threads, 16-byte chunks ---------
    // Row n of a band matrix is the weight table shifted by n: a chunk (n, k = 16 kc ..) is the 16
    // table bytes from d = 16 kc - n + 16 on, i.e. five aligned words and four funnel shifts.
    {
      const int t = threadIdx.x - 32;
      constexpr int per_var = bvar_bytes / 16;
      for (int c = t; c < per_var; c += kSmThreads - 32) {
        const int m = c >> 7, w16 = c & 127;             // matrix (acc * ndy + dy), 16-byte chunk inside it
        const int a = m / ndy, dy = m - a * ndy;
        const int n = (w16 >> 4) * 8 + (w16 & 7), kc = (w16 >> 3) & 1;
        const int base = kc * 16 - n + 16 + 48;          // 1 .. 80
        const uint32_t* wr = reinterpret_cast<const uint32_t*>(wtab[a][dy]) + (base >> 2);
        const uint32_t sh = (base & 3) * 8;
        const uint32_t w0 = wr[0], w1 = wr[1], w2 = wr[2], w3 = wr[3], w4 = wr[4];
        const uint4 val = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh),
                                     __funnelshift_r(w3, w4, sh));
#pragma unroll
        for (int v = 0; v < kSmVariants; ++v) *reinterpret_cast<uint4*>(bmat + v * bvar_bytes + c * 16) = val;
      }
    }
    asm volatile("bar.sync 1, 160;" ::: "memory");
    // ---- edge variants: fold the taps that leave the image onto the mirrored / replicated bytes -----
    if (P.border != B2I_BORDER_ZERO) {
      for (int it = threadIdx.x - 32; it < 3 * kSmWin; it += kSmThreads - 32) {
        const int v = 1 + it / kSmWin, n = it % kSmWin;
        const int k0 = v == 1 ? 0 : v == 2 ? (kl - 1) * kSmKStep : kl * kSmKStep;
        const int o = k0 - 16 + n;
        if (o < 0 || o >= P.W3) continue;
        const int px = o / P.cs, ch = o - px * P.cs;
        for (int tt = -P.R; tt <= P.R; ++tt) {
          const int p = px + tt;
          if (p >= 0 && p < W) continue;
          const int k = fold_px(p, W, P.border) * P.cs + ch - k0;
          if (k < 0 || k >= kSmKStep) continue;
          for (int m = 0; m < nacc * ndy; ++m)
            bmat[v * bvar_bytes + m * kSmBBytes + b_off(n, k)] += (uint8_t)P.w[m / ndy][m % ndy][tt + P.R];
        }
      }
    }
    fence_proxy_async_smem();   // the band matrices are read by the tensor core (async proxy)

    if (warp == 1) {
      // ===== MMA issuer ==========================================================================
      asm volatile("bar.sync 2, 160;" ::: "memory");   // band matrices complete, TMEM ring reset
      tc_fence_after();
      if (lane == 0) stamp(P, 2);
      const uint32_t b_lo_base = b_desc_lo(smem_u32(bmat));
      const bool leader = elect_one() && !(P.dbg & 4);
      uint32_t zmask = 0;        // bit m: row m of the stencil (accumulator-major) has no non-zero weight
      for (int m = 0; m < nacc * ndy; ++m) {
        bool any = false;
        for (int i = 0; i < ndy; ++i) any = any || P.w[m / ndy][m % ndy][i] != 0;
        if (!any) zmask |= 1u << m;
      }
      const bool commit_leader = elect_one();
      uint32_t n = 0, C = 0;     // slabs consumed, chunks started
      uint32_t acquired = 0;     // chunk positions whose previous contents the epilogue has released
      int g = g_lo;
      Run r;
      while (next_run(g, g_hi, P.n_slabs, r)) {
        const int yb = P.y0 + r.b * kSmRows;
        const int ne = r.s1 - r.s0;
        const int oa = r.s0 * 128, ob = r.s1 * 128;   // result columns of the run
        int e_done = 0;                               // chunks published so far
        for (int j = run_first_slab(r); j <= run_last_slab(r, P.n_slabs); ++j, ++n) {
          const uint32_t st = n % kSmStages;
          uint8_t* slab = slabs + st * kSmSlabBytes;
          mbar_wait(&full_slab[st], (n / kSmStages) & 1);
          if (n == 0 && lane == 0) stamp(P, 3);
          // rows above / below the image: copies of the mirrored / replicated rows (swizzled 16-byte chunks)
          if (P.border != B2I_BORDER_ZERO && (yb - P.R < 0 || yb + kSmRows + P.R > P.H)) {
            const int fix = lane >> 3, c16 = lane & 7;                 // up to 4 rows x 8 chunks of 16 bytes
            int row = -1;                                              // slab row this lane repairs
            if (fix < P.R) {                                           // the R rows above the image
              if (yb - P.R + fix < 0) row = fix;
            } else if (fix < 2 * P.R) {                                // the R rows below it
              const int sr = P.H + (fix - P.R) - (yb - P.R);
              if (sr >= 0 && sr < kSmRows + 2 * P.R) row = sr;
            }
            if (row >= 0) {
              const int src_row = fold_row(yb - P.R + row, P.H, P.border) - (yb - P.R);
              if (src_row >= 0 && src_row < kSmRows + 2 * P.R) {
                const uint4 val = *reinterpret_cast<const uint4*>(slab + src_row * 128 + ((c16 ^ (src_row & 7)) << 4));
                *reinterpret_cast<uint4*>(slab + row * 128 + ((c16 ^ (row & 7)) << 4)) = val;
              }
            }
            fence_proxy_async_smem();
            __syncwarp();
          }
          tc_fence_after();
          {
            const uint32_t a_lo_slab = a_desc_lo(smem_u32(slab));
            // 2R+1 (x2 for Sobel) MMAs of one K-step: result columns [rc, rc + n) of the ring, rc + n <= ring
            // 2R+1 (x2 for Sobel) MMAs of one K-step: result columns [rc, rc + n) of the ring, rc + n <= ring.
            // Rows of the stencil that are all zero (Sobel's gy has one) are skipped.
            auto issue = [&](uint32_t a_lo, uint32_t b_lo, uint32_t rc, int n) {
              const uint32_t id = idesc_i8(n);
#pragma unroll
              for (int m = 0; m < nacc * ndy; ++m)
                if (leader && !(zmask >> m & 1))
                  umma_i8_w(tmem + (m / ndy) * ring + rc, a_lo + (m % ndy) * (128 >> 4), kADescHi,
                            b_lo + m * (kSmBBytes >> 4), kBDescHi, id);
            };
            auto issue_wrap = [&](uint32_t a_lo, uint32_t b_lo, uint32_t rc, int n) {  // split where the ring wraps
              const int n1 = min(n, ring - (int)rc);
              issue(a_lo, b_lo, rc, n1);
              if (n > n1) issue(a_lo, b_lo + (n1 >> 3) * (256 >> 4), 0, n - n1);
            };
            auto acquire = [&](uint32_t c_hi) {  // the epilogue has released the ring positions of chunks <= c_hi
              while (acquired <= c_hi) {
                mbar_wait(&chunk_empty[acquired % ring_chunks], ((acquired / ring_chunks) & 1) ^ 1);
                ++acquired;
              }
              tc_fence_after();
            };
            const int t = j - r.s0;
            if (t >= 1 && j + 2 <= r.s1 && j * kSmSlab + kSmSlab <= (kl - 1) * kSmKStep) {
              // ---- interior slab: four full windows, the interior band matrices ---------------------
              acquire(C + t);
              const uint32_t cpos = (C + t) % ring_chunks;
              const uint32_t rc0 = (cpos * 128 + ring - 16) & (ring - 1);
              // Consecutive MMAs into the same accumulator columns form a dependent chain (each waits for
              // the one before to land in TMEM). The windows of K-steps 0 and 2 (and of 1 and 3) do not
              // overlap, so their MMAs alternate: two independent chains in flight.
              const uint32_t id = idesc_i8(kSmWin);
              auto pair = [&](int ia, int ib) {
                const uint32_t ra = (rc0 + 32 * ia) & (ring - 1), rb = (rc0 + 32 * ib) & (ring - 1);
#pragma unroll
                for (int m = 0; m < nacc * ndy; ++m) {
                  if (leader && !(zmask >> m & 1)) {
                    umma_i8_w(tmem + (m / ndy) * ring + ra, a_lo_slab + 2 * ia + (m % ndy) * (128 >> 4), kADescHi,
                              b_lo_base + m * (kSmBBytes >> 4), kBDescHi, id);
                    umma_i8_w(tmem + (m / ndy) * ring + rb, a_lo_slab + 2 * ib + (m % ndy) * (128 >> 4), kADescHi,
                              b_lo_base + m * (kSmBBytes >> 4), kBDescHi, id);
                  }
                }
              };
              if (cpos == 0 || (P.dbg & 32)) {       // window 0 straddles the end of the ring: in two parts
                issue_wrap(a_lo_slab, b_lo_base, rc0, kSmWin);
                issue(a_lo_slab + 4, b_lo_base, (rc0 + 64) & (ring - 1), kSmWin);
              } else {
                pair(0, 2);
              }
              while (e_done < t) {   // the chunk to the left is complete with K-step 0
                if (commit_leader) umma_commit(&chunk_full[(C + e_done) % ring_chunks]);
                ++e_done;
              }
              acquire(C + t + 1);    // the last window reaches 16 columns into the next chunk
              if (cpos == ring_chunks - 1 || (P.dbg & 32)) {   // window 3 straddles the end of the ring
                issue(a_lo_slab + 2, b_lo_base, (rc0 + 32) & (ring - 1), kSmWin);
                issue_wrap(a_lo_slab + 6, b_lo_base, (rc0 + 96) & (ring - 1), kSmWin);
              } else {
                pair(1, 3);
              }
            } else {
              // ---- slabs at the ends of a run or of the image: clipped windows, edge band matrices --------
#pragma unroll 1
              for (int i = 0; i < 4; ++i) {
                const int k0 = j * kSmSlab + i * kSmKStep;
                if (k0 < oa - kSmKStep || k0 > ob || k0 >= P.W3) continue;
                // result columns this K-step touches, clipped to the run: [wa, wb), multiples of 16
                const int wa = max(k0 - 16, oa), wb = min(k0 + 48, ob);
                if (wb <= wa) continue;
                acquire(C + (uint32_t)((wb - 1 - oa) >> 7));
                const int v = k0 == 0 ? 1 : k0 == (kl - 1) * kSmKStep ? 2 : k0 == kl * kSmKStep ? 3 : 0;
                const int noff = wa - (k0 - 16);
                issue_wrap(a_lo_slab + i * (kSmKStep >> 4), b_lo_base + v * (bvar_bytes >> 4) + (noff >> 3) * (256 >> 4),
                           (C * 128u + (uint32_t)(wa - oa)) & (uint32_t)(ring - 1), wb - wa);
                // chunk e is complete once the K-step at its right edge (+6 bytes of reach) has run
                while (e_done < ne && (k0 >= oa + (e_done + 1) * 128 || k0 + kSmKStep >= P.W3)) {
                  if (commit_leader) umma_commit(&chunk_full[(C + e_done) % ring_chunks]);
                  ++e_done;
                }
              }
            }
            if (commit_leader) umma_commit(&empty_slab[st]);   // the slab may be refilled once these MMAs have read it
          }
          __syncwarp();
        }
        {
          while (e_done < ne) {   // (a run ending at the right edge whose last K-steps were outside the image)
            if (commit_leader) umma_commit(&chunk_full[(C + e_done) % ring_chunks]);
            ++e_done;
          }
        }
        C += ne;
      }
      if (lane == 0) stamp(P, 4);
    } else {
      // ===== epilogue warps 2..5: TMEM -> bytes -> staged chunk -> TMA store ============================
      const int q = warp & 3;                       // TMEM lane quarter this warp may access
      const int row = q * 32 + lane;                // result row within the band
      const uint32_t lane_addr = tmem + ((uint32_t)(q * 32) << 16);
      const uint32_t reset = NACC == 2 ? 0u : (uint32_t)P.bias;
      uint32_t rv[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) rv[i] = reset;
      for (int c = 0; c < kSmRing; c += 32) tmem_fill32(lane_addr + c, rv);
      tmem_st_wait();
      tc_fence_before();
      asm volatile("bar.sync 2, 160;" ::: "memory");
      pdl_wait();   // (the results depend on the producer's loads anyway; this orders the stores explicitly)
      const bool fast8 = NACC == 1 && P.fast8;
      const bool leader = threadIdx.x == 64;
      uint32_t C = 0, n_store = 0;
      int g = g_lo;
      Run r;
      while (next_run(g, g_hi, P.n_slabs, r)) {
        const int ne = r.s1 - r.s0;
        for (int e = 0; e < ne; ++e, ++C) {
          const uint32_t pos = C % ring_chunks;
          uint8_t* ob = stage_out + (n_store & 1) * kSmOutBytes;
          if (leader) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");  // the store that used `ob`
          asm volatile("bar.sync 3, 128;" ::: "memory");
          mbar_wait(&chunk_full[pos], (C / ring_chunks) & 1);
          tc_fence_after();
          if (C == 0 && leader) stamp(P, 5);
          // two 16-byte chunks of the row per 32 columns, at their swizzled places (conflict-free per quarter warp)
          auto stage = [&](int gq, const uint32_t (&o)[8]) {
            uint8_t* rowp = ob + row * 128;
            *reinterpret_cast<uint4*>(rowp + (((2 * gq) ^ (row & 7)) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<uint4*>(rowp + (((2 * gq + 1) ^ (row & 7)) << 4)) = make_uint4(o[4], o[5], o[6], o[7]);
          };
          if (P.dbg & 8) {
          } else if constexpr (NACC == 2) {
#pragma unroll 1
            for (int gq = 0; gq < 4; ++gq) {
              const uint32_t col = pos * 128 + gq * 32;
              uint32_t v[32], u[32], o[8];
              tmem_ld32(lane_addr + col, v);
              tmem_ld32(lane_addr + ring + col, u);
              tmem_ld_wait();
              tmem_fill32(lane_addr + col, rv);
              tmem_fill32(lane_addr + ring + col, rv);
#pragma unroll
              for (int k = 0; k < 8; ++k)
                o[k] = sat_pack4(abs((int)v[4 * k]) + abs((int)u[4 * k]), abs((int)v[4 * k + 1]) + abs((int)u[4 * k + 1]),
                                 abs((int)v[4 * k + 2]) + abs((int)u[4 * k + 2]), abs((int)v[4 * k + 3]) + abs((int)u[4 * k + 3]));
              stage(gq, o);
            }
          } else {
#pragma unroll 1
            for (int gp = 0; gp < 2; ++gp) {   // two groups of 32 columns in flight
              const uint32_t col = pos * 128 + gp * 64;
              uint32_t v[32], u[32], o[8];
              tmem_ld32(lane_addr + col, v);
              tmem_ld32(lane_addr + col + 32, u);
              tmem_ld_wait();
              tmem_fill32(lane_addr + col, rv);
              tmem_fill32(lane_addr + col + 32, rv);
              if (fast8) {
#pragma unroll
                for (int k = 0; k < 8; ++k)
                  o[k] = __byte_perm(__byte_perm(v[4 * k], v[4 * k + 1], 0x0051), __byte_perm(v[4 * k + 2], v[4 * k + 3], 0x0051), 0x5410);
                stage(2 * gp, o);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                  o[k] = __byte_perm(__byte_perm(u[4 * k], u[4 * k + 1], 0x0051), __byte_perm(u[4 * k + 2], u[4 * k + 3], 0x0051), 0x5410);
                stage(2 * gp + 1, o);
              } else {
#pragma unroll
                for (int k = 0; k < 8; ++k)
                  o[k] = sat_pack4((int)v[4 * k] >> P.shift, (int)v[4 * k + 1] >> P.shift, (int)v[4 * k + 2] >> P.shift,
                                   (int)v[4 * k + 3] >> P.shift);
                stage(2 * gp, o);
#pragma unroll
                for (int k = 0; k < 8; ++k)
                  o[k] = sat_pack4((int)u[4 * k] >> P.shift, (int)u[4 * k + 1] >> P.shift, (int)u[4 * k + 2] >> P.shift,
                                   (int)u[4 * k + 3] >> P.shift);
                stage(2 * gp + 1, o);
              }
            }
          }
          tmem_st_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&chunk_empty[pos]);
          fence_proxy_async_smem();
          asm volatile("bar.sync 4, 128;" ::: "memory");
          if (leader) {
            tma_store_2d(&dst_map, (r.s0 + e) * 128, r.b * kSmRows, ob);
            tma_store_commit();
          }
          ++n_store;
        }
      }
      if (leader) {
        stamp(P, 6);
        tma_store_wait_all();
        stamp(P, 7);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) stamp(P, 9);
  if (warp == 1) tmem_free_512(tmem);
}

cudaError_t launch_stencil_mma(const StencilMmaParams& P, const CUtensorMap& src_map, const CUtensorMap& dst_map,
                               int grid, cudaStream_t stream) {
  static bool configured[64][3] = {{false}};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  using Kern = void (*)(StencilMmaParams, CUtensorMap, CUtensorMap);
  static const Kern table[3] = {stencil_mma_kernel<3, 1>, stencil_mma_kernel<5, 1>, stencil_mma_kernel<3, 2>};
  const int variant = P.mode == SM_SOBEL ? 2 : P.R == 2 ? 1 : 0;
  if (P.mode == SM_SOBEL && P.R != 1) return cudaErrorInvalidValue;
  const Kern kern = table[variant];
  const int smem = kSmStages * kSmSlabBytes + kSmVariants * 2 * 3 * kSmBBytes + 2 * kSmOutBytes + 1024;
  if (dev < 64 && !configured[dev][variant]) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    configured[dev][variant] = true;
  }
  return launch_pdl(true, kern, grid, kSmThreads, smem, stream, P, src_map, dst_map);
}

}  // namespace b2i
