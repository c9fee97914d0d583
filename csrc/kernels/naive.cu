// Reference-style kernels: one thread per pixel, one launch per filter, 16x16 blocks, the gray
// image bouncing through global memory between launches — the structure of kernel.cu:31-94 and
// its launch sequence (kernel.cu:181-195), generalised to every op of a chain and made
// out-of-place.
//
// They exist for two reasons: (1) BASELINE.md needs a "reference-style" baseline for the filters
// the reference does not have (invert, Gaussian, Sobel); (2) they are an independent on-GPU
// oracle for the tile-chain kernel, and the fallback for degenerate image sizes. They are not
// the product path.
#include "../core/error.hpp"
#include "b200img.h"
#include <cuda_runtime.h>

namespace b2i {

namespace {

__device__ __forceinline__ int map_coord(int p, int n, int border) {
  if ((unsigned)p < (unsigned)n) return p;
  if (border == B2I_BORDER_REFLECT101) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
  }
  if (border == B2I_BORDER_REPLICATE) return p < 0 ? 0 : n - 1;
  return -1;
}

__device__ __forceinline__ float clampf(float v) {  // kernel.cu:19-24
  return v < 0.f ? 0.f : (v > 255.f ? 255.f : v);
}

__global__ void naive_gray(const uint8_t* src, size_t sp, uint8_t* dst, size_t dp, int W, int H,
                           int cv, int rgb) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= W || y >= H) return;
  const uint8_t* p = src + y * sp + 3 * (size_t)x;
  const int b = p[rgb ? 2 : 0], g = p[1], r = p[rgb ? 0 : 2];
  dst[y * dp + x] = cv ? (uint8_t)((3735 * b + 19235 * g + 9798 * r + 16384) >> 15)
                       : (uint8_t)(b * 11 / 100 + g * 59 / 100 + r * 3 / 10);
}

__global__ void naive_lut(const uint8_t* src, size_t sp, uint8_t* dst, size_t dp, int WB, int H,
                          const uint8_t* lut) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= WB || y >= H) return;
  dst[y * dp + x] = lut[src[y * sp + x]];
}

__global__ void naive_expand(const uint8_t* src, size_t sp, uint8_t* dst, size_t dp, int W, int H) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= W || y >= H) return;
  const uint8_t v = src[y * sp + x];
  uint8_t* d = dst + y * dp + 3 * (size_t)x;
  d[0] = d[1] = d[2] = v;
}

struct NaiveStencil {
  int k, shift, bias, border, sobel;
  int w[25];
};

__global__ void naive_stencil(const uint8_t* src, size_t sp, uint8_t* dst, size_t dp, int W, int H,
                              int C, NaiveStencil st) {
  const int xb = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (xb >= W * C || y >= H) return;
  const int x = xb / C, c = xb - x * C, r = st.k / 2;
  if (st.border == B2I_BORDER_REF &&
      (x < r + 1 || x > W - r - 1 || y < r + 1 || y > H - r - 1)) {
    dst[y * dp + xb] = src[y * sp + xb];
    return;
  }
  float sum = 0.f, sum2 = 0.f;  // float accumulation like kernel.cu:65,87-88 (exact for u8 data)
  for (int j = 0; j < st.k; ++j) {
    const int my = map_coord(y + j - r, H, st.border);
    for (int i = 0; i < st.k; ++i) {
      const int mx = map_coord(x + i - r, W, st.border);
      const float v = (my < 0 || mx < 0) ? 0.f : (float)src[my * sp + (size_t)mx * C + c];
      if (st.sobel) {
        const int gx = (i - 1) * (j == 1 ? 2 : 1), gy = (j - 1) * (i == 1 ? 2 : 1);
        sum += gx * v;
        sum2 += gy * v;
      } else {
        sum += st.w[j * st.k + i] * v;
      }
    }
  }
  if (st.sobel) {
    dst[y * dp + xb] = (uint8_t)clampf(fabsf(sum) + fabsf(sum2));
  } else {
    int s = ((int)sum + st.bias) >> st.shift;
    dst[y * dp + xb] = (uint8_t)(s < 0 ? 0 : s > 255 ? 255 : s);
  }
}

}  // namespace

// dst <- chain(src), whole image, one launch per op. tmp0/tmp1 hold W*H*3 bytes each and
// lut_scratch 256 bytes per LUT op (all device memory, caller-owned: nothing is allocated here).
int run_naive(const b2i_chain_t* chain, const uint8_t* src, int W, int H, int C, size_t src_pitch,
              uint8_t* dst, size_t dst_pitch, uint8_t* tmp0, uint8_t* tmp1, uint8_t* lut_scratch,
              cudaStream_t stream) {
  const int out_c = b2i_chain_out_channels(chain, C);
  if (out_c < 0) return out_c;
  B2I_CHECK_ARG(src && dst && tmp0 && tmp1 && lut_scratch, "null buffer");
  const dim3 block(16, 16);  // BLOCK_SIZE 16, kernel.cu:13,181
  const uint8_t* cur = src;
  size_t cur_pitch = src_pitch;
  int c = C, n_lut = 0;
  for (int i = 0; i < chain->n_ops; ++i) {
    const b2i_op_t& op = chain->ops[i];
    const bool last = i == chain->n_ops - 1;
    int oc = c;
    if (op.kind == B2I_OP_GRAY_REF || op.kind == B2I_OP_GRAY_CV) oc = 1;
    if (op.kind == B2I_OP_EXPAND) oc = 3;
    uint8_t* out = last ? dst : (cur == tmp0 ? tmp1 : tmp0);
    const size_t out_pitch = last ? dst_pitch : (size_t)W * oc;
    const dim3 grid_px((W + 15) / 16, (H + 15) / 16), grid_b((W * c + 15) / 16, (H + 15) / 16);
    switch (op.kind) {
      case B2I_OP_GRAY_REF:
      case B2I_OP_GRAY_CV:
        naive_gray<<<grid_px, block, 0, stream>>>(cur, cur_pitch, out, out_pitch, W, H,
                                                 op.kind == B2I_OP_GRAY_CV,
                                                 chain->order == B2I_ORDER_RGB);
        break;
      case B2I_OP_LUT: {
        uint8_t* l = lut_scratch + 256 * (n_lut++);
        cudaError_t e = cudaMemcpyAsync(l, op.lut, 256, cudaMemcpyHostToDevice, stream);
        if (e != cudaSuccess) B2I_FAIL(B2I_ERR_CUDA, "LUT upload: %s", cudaGetErrorString(e));
        naive_lut<<<grid_b, block, 0, stream>>>(cur, cur_pitch, out, out_pitch, W * c, H, l);
        break;
      }
      case B2I_OP_EXPAND:
        naive_expand<<<grid_px, block, 0, stream>>>(cur, cur_pitch, out, out_pitch, W, H);
        break;
      default: {
        NaiveStencil st;
        st.k = op.kind == B2I_OP_SOBEL ? 3 : op.ksize;
        st.shift = op.shift;
        st.bias = op.bias;
        st.border = op.border;
        st.sobel = op.kind == B2I_OP_SOBEL;
        for (int q = 0; q < 25; ++q) st.w[q] = op.weights[q];
        naive_stencil<<<grid_b, block, 0, stream>>>(cur, cur_pitch, out, out_pitch, W, H, c, st);
        break;
      }
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) B2I_FAIL(B2I_ERR_CUDA, "naive launch %d: %s", i, cudaGetErrorString(e));
    cur = out;
    cur_pitch = out_pitch;
    c = oc;
  }
  return B2I_OK;
}

}  // namespace b2i
