// Streaming kernel for chains without a stencil: [lut] -> [gray] -> [lut] -> [expand].
//
// This is what the reference's grayscaleKernel + contrastKernel (kernel.cu:31-58) and the host
// GRAY2BGR (kernel.cu:210) collapse to when no stencil sits between them: each thread moves
// 16 pixels with 16-byte loads and stores (3 x LDG.128 -> 1 or 3 x STG.128) instead of three
// byte loads and one byte store per pixel. The pointers may be peer addresses: reading the band
// from the owner GPU fuses the scatter (kernel.cu:137), writing into the root GPU's image fuses
// the gather (kernel.cu:223).
#include "b200img.h"
#include "device_utils.cuh"
#include "launch_pdl.hpp"
#include "tile_plan.hpp"

namespace b2i {
using namespace dev;

namespace {

constexpr int kPwThreads = 256;

__device__ __forceinline__ uint32_t lut4(const uint8_t* lut, uint32_t w) {
  const uint32_t a = lut[w & 0xFF], b = lut[(w >> 8) & 0xFF], c = lut[(w >> 16) & 0xFF],
                 d = lut[w >> 24];
  return pack4_lo(a, b, c, d);
}

// 4 gray pixels from 12 interleaved bytes.
template <int GRAY>
__device__ __forceinline__ uint32_t gray4(uint32_t w0, uint32_t w1, uint32_t w2,
                                          const GrayRefCoef& gk, int rgb) {
  uint32_t g0, g1, g2, g3;
  if (GRAY == 1) {
    g0 = gray_ref_f23(byte_to_f23<0>(w0), byte_to_f23<1>(w0), byte_to_f23<2>(w0), gk);
    g1 = gray_ref_f23(byte_to_f23<3>(w0), byte_to_f23<0>(w1), byte_to_f23<1>(w1), gk);
    g2 = gray_ref_f23(byte_to_f23<2>(w1), byte_to_f23<3>(w1), byte_to_f23<0>(w2), gk);
    g3 = gray_ref_f23(byte_to_f23<1>(w2), byte_to_f23<2>(w2), byte_to_f23<3>(w2), gk);
  } else {
    g0 = gray_cv_px(w0 & 0xFF, (w0 >> 8) & 0xFF, (w0 >> 16) & 0xFF, rgb);
    g1 = gray_cv_px(w0 >> 24, w1 & 0xFF, (w1 >> 8) & 0xFF, rgb);
    g2 = gray_cv_px((w1 >> 16) & 0xFF, w1 >> 24, w2 & 0xFF, rgb);
    g3 = gray_cv_px((w2 >> 8) & 0xFF, (w2 >> 16) & 0xFF, w2 >> 24, rgb);
  }
  return pack4_lo(g0, g1, g2, g3);
}

// One pixel, any alignment: row tails and images whose pitch is not a multiple of 16.
template <int CIN, int GRAY, bool EXPAND>
__device__ __forceinline__ void scalar_px(const PointwiseParams& P, const uint8_t* s, uint8_t* d,
                                          const uint8_t* lut0, const uint8_t* lut1) {
  uint32_t v[3];
#pragma unroll
  for (int c = 0; c < CIN; ++c) v[c] = P.has_lut0 ? lut0[s[c]] : s[c];
  if (GRAY) {
    const int ib = P.rgb_order ? 2 : 0, ir = 2 - ib;
    uint32_t g = GRAY == 1 ? v[ib] * 11 / 100 + v[1] * 59 / 100 + v[ir] * 3 / 10
                           : (3735 * v[ib] + 19235 * v[1] + 9798 * v[ir] + 16384) >> 15;
    if (P.has_lut1) g = lut1[g];
    if (EXPAND) d[0] = d[1] = d[2] = (uint8_t)g;
    else d[0] = (uint8_t)g;
  } else {
#pragma unroll
    for (int c = 0; c < CIN; ++c) {
      const uint32_t o = P.has_lut1 ? lut1[v[c]] : v[c];
      if (EXPAND) d[0] = d[1] = d[2] = (uint8_t)o;
      else d[c] = (uint8_t)o;
    }
  }
}

// One warp moves 512 pixels per step. Every global access is a fully coalesced 16-byte-per-lane
// instruction (lane l touches bytes [16 l, 16 l + 16) of a 512-byte run); where a lane needs its
// own 16 pixels as 48 contiguous bytes (3-channel in or out) the warp transposes through 1536
// bytes of shared memory (both access patterns are bank-conflict free). The first version let
// each lane load its 48 bytes directly: three instructions with a 48-byte lane stride, each
// fetching every sector of the run again (L1 allocation is off for streaming data), which capped
// `invert` at 0.64 of the HBM roofline.
template <int CIN, int GRAY, bool EXPAND>
__global__ void __launch_bounds__(kPwThreads)
pointwise_kernel(const __grid_constant__ PointwiseParams P) {
  __shared__ __align__(16) uint8_t lut0[256];
  __shared__ __align__(16) uint8_t lut1[256];
  __shared__ __align__(16) uint8_t xpose[kPwThreads / 32][1536];
  pdl_launch_dependents();   // the next launch in the stream may set itself up behind this one
  for (int i = threadIdx.x; i < 64; i += kPwThreads) {
    reinterpret_cast<uint32_t*>(lut0)[i] = reinterpret_cast<const uint32_t*>(P.lut0)[i];
    reinterpret_cast<uint32_t*>(lut1)[i] = reinterpret_cast<const uint32_t*>(P.lut1)[i];
  }
  __syncthreads();
  pdl_wait();   // the launch before this one has completed: its output may be this one's input
  constexpr int COUT = GRAY ? (EXPAND ? 3 : 1) : (EXPAND ? 3 : CIN);
  constexpr int MID = GRAY ? 1 : CIN;          // channels between the two LUTs
  constexpr bool BYTEWISE = !GRAY && !EXPAND;  // C -> C per byte: no pixel structure needed
  const GrayRefCoef gk = gray_ref_coef(P.rgb_order);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t* xp = xpose[warp];
  const int su_x = (P.W + 511) >> 9;  // 512-pixel super-units per row
  const long long n_su = (long long)su_x * P.rows;
  const long long warps_total = (long long)gridDim.x * (kPwThreads / 32);
  const uint8_t* src = reinterpret_cast<const uint8_t*>(P.src_ptr);
  uint8_t* dst = reinterpret_cast<uint8_t*>(P.dst_ptr);
  // Kernels that write 3 bytes per pixel issue the loads of a run one iteration ahead (registers
  // `pre`), so that a warp has its next 1.5 KB in flight while it converts and stores the current
  // one (8K invert 35.2 -> 33.5 us, gray,contrast,expand 36.1 -> 33.8 us). Gray output (a third of
  // the stores) measured slower that way (22.8 -> 24.4 us) and loads at the top of the iteration.
  constexpr bool PREFETCH = COUT == 3;
  uint4 pre[CIN];
  auto full_run = [&](long long u) {
    const int y = (int)(u / su_x), sx = (int)(u - (long long)y * su_x);
    return P.vec_ok && (sx << 9) + 512 <= P.W;
  };
  auto issue = [&](long long u) {
    const int y = (int)(u / su_x), sx = (int)(u - (long long)y * su_x);
    const uint8_t* srow = src + (long long)y * P.src_pitch + (long long)(sx << 9) * CIN;
#pragma unroll
    for (int q = 0; q < CIN; ++q) pre[q] = ldg_stream_v4(srow + q * 512 + lane * 16);
  };
  const long long u0 = (long long)blockIdx.x * (kPwThreads / 32) + warp;
  bool pre_ok = PREFETCH && u0 < n_su && full_run(u0);
  if (pre_ok) issue(u0);
  for (long long u = u0; u < n_su; u += warps_total) {
    const int y = (int)(u / su_x), sx = (int)(u - (long long)y * su_x);
    const int xw = sx << 9;  // first pixel of the warp's run
    const uint8_t* srow = src + (long long)y * P.src_pitch + (long long)xw * CIN;
    uint8_t* drow = dst + (long long)y * P.dst_pitch + (long long)xw * COUT;
    bool cur_ok;
    uint4 cur[CIN];
    if constexpr (PREFETCH) {
      cur_ok = pre_ok;
#pragma unroll
      for (int q = 0; q < CIN; ++q) cur[q] = pre[q];
      pre_ok = u + warps_total < n_su && full_run(u + warps_total);
      if (pre_ok) issue(u + warps_total);
    } else {
      cur_ok = P.vec_ok && xw + 512 <= P.W;
    }
    if (cur_ok) {
      uint32_t in[4 * CIN];
      if (BYTEWISE || CIN == 1) {
        // bytes are independent (or one byte is one pixel): keep the coalesced layout
#pragma unroll
        for (int q = 0; q < CIN; ++q) {
          const uint4 v = PREFETCH ? cur[q] : ldg_stream_v4(srow + q * 512 + lane * 16);
          in[4 * q] = v.x; in[4 * q + 1] = v.y; in[4 * q + 2] = v.z; in[4 * q + 3] = v.w;
        }
      } else {
        // 3-channel pixels: coalesced loads, then each lane picks up its own 48 contiguous bytes
#pragma unroll
        for (int q = 0; q < 3; ++q)
          *reinterpret_cast<uint4*>(xp + q * 512 + lane * 16) = PREFETCH ? cur[q] : ldg_stream_v4(srow + q * 512 + lane * 16);
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const uint4 v = *reinterpret_cast<const uint4*>(xp + lane * 48 + q * 16);
          in[4 * q] = v.x; in[4 * q + 1] = v.y; in[4 * q + 2] = v.z; in[4 * q + 3] = v.w;
        }
        __syncwarp();
      }
      if (P.has_lut0 == 2) {  // the table is 255 - v (invert): one NOT per word instead of four lookups
#pragma unroll
        for (int q = 0; q < 4 * CIN; ++q) in[q] = ~in[q];
      } else if (P.has_lut0) {
#pragma unroll
        for (int q = 0; q < 4 * CIN; ++q) in[q] = lut4(lut0, in[q]);
      }
      uint32_t mid[4 * MID];
      if (GRAY) {
#pragma unroll
        for (int q = 0; q < 4; ++q) mid[q] = gray4<GRAY ? GRAY : 1>(in[3 * q], in[3 * q + 1], in[3 * q + 2], gk, P.rgb_order);
      } else {
#pragma unroll
        for (int q = 0; q < 4 * MID; ++q) mid[q] = in[q];
      }
      if (P.has_lut1 == 2) {
#pragma unroll
        for (int q = 0; q < 4 * MID; ++q) mid[q] = ~mid[q];
      } else if (P.has_lut1) {
#pragma unroll
        for (int q = 0; q < 4 * MID; ++q) mid[q] = lut4(lut1, mid[q]);
      }
      if (EXPAND) {  // MID == 1: the lane's 16 gray bytes -> 48 bytes, transposed back to coalesced stores
        uint32_t e[12];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          e[3 * q] = __byte_perm(mid[q], 0, 0x1000);
          e[3 * q + 1] = __byte_perm(mid[q], 0, 0x2211);
          e[3 * q + 2] = __byte_perm(mid[q], 0, 0x3332);
        }
#pragma unroll
        for (int q = 0; q < 3; ++q)  // 16-byte stores at a 48-byte lane stride: conflict free
          *reinterpret_cast<uint4*>(xp + lane * 48 + q * 16) = make_uint4(e[4 * q], e[4 * q + 1], e[4 * q + 2], e[4 * q + 3]);
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 3; ++q)
          stg_stream_v4(drow + q * 512 + lane * 16, *reinterpret_cast<const uint4*>(xp + q * 512 + lane * 16));
        __syncwarp();
      } else {
#pragma unroll
        for (int q = 0; q < MID; ++q)
          stg_stream_v4(drow + q * 512 + lane * 16, make_uint4(mid[4 * q], mid[4 * q + 1], mid[4 * q + 2], mid[4 * q + 3]));
      }
    } else {
      // ragged end of a row, or buffers that are not 16-byte aligned: one pixel at a time
      const int n = min(512, P.W - xw);
      for (int i = lane; i < n; i += 32)
        scalar_px<CIN, GRAY, EXPAND>(P, srow + i * CIN, drow + i * COUT, lut0, lut1);
    }
  }
}

}  // namespace

cudaError_t launch_pointwise(const PointwiseParams& P, int grid, cudaStream_t stream) {
  const int key = P.cin * 100 + P.gray * 10 + (P.expand ? 1 : 0);
  cudaError_t e = cudaSuccess;
  switch (key) {
    case 310: e = launch_pdl(true, pointwise_kernel<3, 1, false>, grid, kPwThreads, 0, stream, P); break;
    case 311: e = launch_pdl(true, pointwise_kernel<3, 1, true>, grid, kPwThreads, 0, stream, P); break;
    case 320: e = launch_pdl(true, pointwise_kernel<3, 2, false>, grid, kPwThreads, 0, stream, P); break;
    case 321: e = launch_pdl(true, pointwise_kernel<3, 2, true>, grid, kPwThreads, 0, stream, P); break;
    case 300: e = launch_pdl(true, pointwise_kernel<3, 0, false>, grid, kPwThreads, 0, stream, P); break;
    case 100: e = launch_pdl(true, pointwise_kernel<1, 0, false>, grid, kPwThreads, 0, stream, P); break;
    case 101: e = launch_pdl(true, pointwise_kernel<1, 0, true>, grid, kPwThreads, 0, stream, P); break;
    default: return cudaErrorInvalidValue;
  }
  return e;
}

}  // namespace b2i
