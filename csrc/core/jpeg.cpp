// JPEG input and output on the root, off the hot path.
//
// The reference's only file format is JPEG: rank 0 decodes one with cv::imread and encodes the
// result with cv::imwrite (kernel.cu:110,236; kern.cpp:33,92). OpenCV's C++ library is not on this
// image; nvJPEG is, so the codec is that library's (a library codec is fine here: file I/O is
// outside every timed region, exactly as in the reference, kernel.cu:190). It is loaded with
// dlopen on first use: libb200img.so has no link-time dependency on it, and a machine without
// nvJPEG still runs everything else (JPEG calls then fail with B2I_ERR_UNSUPPORTED). Decoding
// and encoding run on the current CUDA device.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nvjpeg.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "error.hpp"

namespace {

struct NvJpegApi {
  void* lib = nullptr;
  nvjpegStatus_t (*CreateSimple)(nvjpegHandle_t*) = nullptr;
  nvjpegStatus_t (*Destroy)(nvjpegHandle_t) = nullptr;
  nvjpegStatus_t (*JpegStateCreate)(nvjpegHandle_t, nvjpegJpegState_t*) = nullptr;
  nvjpegStatus_t (*JpegStateDestroy)(nvjpegJpegState_t) = nullptr;
  nvjpegStatus_t (*GetImageInfo)(nvjpegHandle_t, const unsigned char*, size_t, int*, nvjpegChromaSubsampling_t*, int*,
                                 int*) = nullptr;
  nvjpegStatus_t (*Decode)(nvjpegHandle_t, nvjpegJpegState_t, const unsigned char*, size_t, nvjpegOutputFormat_t,
                           nvjpegImage_t*, cudaStream_t) = nullptr;
  nvjpegStatus_t (*EncoderStateCreate)(nvjpegHandle_t, nvjpegEncoderState_t*, cudaStream_t) = nullptr;
  nvjpegStatus_t (*EncoderStateDestroy)(nvjpegEncoderState_t) = nullptr;
  nvjpegStatus_t (*EncoderParamsCreate)(nvjpegHandle_t, nvjpegEncoderParams_t*, cudaStream_t) = nullptr;
  nvjpegStatus_t (*EncoderParamsDestroy)(nvjpegEncoderParams_t) = nullptr;
  nvjpegStatus_t (*EncoderParamsSetQuality)(nvjpegEncoderParams_t, int, cudaStream_t) = nullptr;
  nvjpegStatus_t (*EncoderParamsSetSamplingFactors)(nvjpegEncoderParams_t, nvjpegChromaSubsampling_t,
                                                    cudaStream_t) = nullptr;
  nvjpegStatus_t (*EncodeImage)(nvjpegHandle_t, nvjpegEncoderState_t, const nvjpegEncoderParams_t, const nvjpegImage_t*,
                                nvjpegInputFormat_t, int, int, cudaStream_t) = nullptr;
  nvjpegStatus_t (*EncodeRetrieveBitstream)(nvjpegHandle_t, nvjpegEncoderState_t, unsigned char*, size_t*,
                                            cudaStream_t) = nullptr;
  bool ok = false;
};

template <class F>
bool sym(void* lib, const char* name, F& f) {
  f = reinterpret_cast<F>(dlsym(lib, name));
  return f != nullptr;
}

const NvJpegApi& api() {
  static NvJpegApi a;
  static std::once_flag once;
  std::call_once(once, [] {
    for (const char* n : {"libnvjpeg.so.12", "libnvjpeg.so", "/usr/local/cuda/lib64/libnvjpeg.so.12"}) {
      a.lib = dlopen(n, RTLD_NOW | RTLD_LOCAL);
      if (a.lib) break;
    }
    if (!a.lib) return;
    a.ok = sym(a.lib, "nvjpegCreateSimple", a.CreateSimple) && sym(a.lib, "nvjpegDestroy", a.Destroy) &&
           sym(a.lib, "nvjpegJpegStateCreate", a.JpegStateCreate) && sym(a.lib, "nvjpegJpegStateDestroy", a.JpegStateDestroy) &&
           sym(a.lib, "nvjpegGetImageInfo", a.GetImageInfo) && sym(a.lib, "nvjpegDecode", a.Decode) &&
           sym(a.lib, "nvjpegEncoderStateCreate", a.EncoderStateCreate) &&
           sym(a.lib, "nvjpegEncoderStateDestroy", a.EncoderStateDestroy) &&
           sym(a.lib, "nvjpegEncoderParamsCreate", a.EncoderParamsCreate) &&
           sym(a.lib, "nvjpegEncoderParamsDestroy", a.EncoderParamsDestroy) &&
           sym(a.lib, "nvjpegEncoderParamsSetQuality", a.EncoderParamsSetQuality) &&
           sym(a.lib, "nvjpegEncoderParamsSetSamplingFactors", a.EncoderParamsSetSamplingFactors) &&
           sym(a.lib, "nvjpegEncodeImage", a.EncodeImage) &&
           sym(a.lib, "nvjpegEncodeRetrieveBitstream", a.EncodeRetrieveBitstream);
  });
  return a;
}

struct DevBuf {  // RAII: nothing leaks on the error paths
  void* p = nullptr;
  ~DevBuf() {
    if (p) cudaFree(p);
  }
};

bool ends_with_ci(const char* s, const char* suffix) {
  const size_t n = strlen(s), m = strlen(suffix);
  if (n < m) return false;
  for (size_t i = 0; i < m; ++i)
    if (tolower((unsigned char)s[n - m + i]) != suffix[i]) return false;
  return true;
}

}  // namespace

#define B2I_NVJ(expr)                                                                             \
  do {                                                                                            \
    nvjpegStatus_t s_ = (expr);                                                                   \
    if (s_ != NVJPEG_STATUS_SUCCESS) B2I_FAIL(B2I_ERR_IO, "%s: nvJPEG status %d", #expr, (int)s_); \
  } while (0)
#define B2I_CUDA_IO(expr)                                                                        \
  do {                                                                                           \
    cudaError_t e_ = (expr);                                                                     \
    if (e_ != cudaSuccess) B2I_FAIL(B2I_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_));      \
  } while (0)

extern "C" int b2i_jpeg_available(void) { return api().ok ? 1 : 0; }

// Decodes a JPEG file into a malloc'ed interleaved image (free with b2i_free): 3 channels in
// `order`, or 1 channel for a grayscale file -- what cv::imread(IMREAD_UNCHANGED) returns
// (kernel.cu:110).
extern "C" int b2i_read_jpeg(const char* path, int order, uint8_t** data, int* W, int* H, int* C) {
  B2I_CHECK_ARG(path && data && W && H && C, "null argument");
  const NvJpegApi& nj = api();
  if (!nj.ok) B2I_FAIL(B2I_ERR_UNSUPPORTED, "nvJPEG (libnvjpeg.so.12) is not available on this machine");
  FILE* f = fopen(path, "rb");
  if (!f) B2I_FAIL(B2I_ERR_IO, "cannot open '%s' for reading", path);
  std::vector<unsigned char> bytes;
  unsigned char chunk[1 << 16];
  size_t n;
  while ((n = fread(chunk, 1, sizeof chunk, f)) > 0) bytes.insert(bytes.end(), chunk, chunk + n);
  fclose(f);
  if (bytes.size() < 4 || bytes[0] != 0xFF || bytes[1] != 0xD8) B2I_FAIL(B2I_ERR_IO, "'%s' is not a JPEG file", path);
  nvjpegHandle_t h = nullptr;
  nvjpegJpegState_t st = nullptr;
  B2I_NVJ(nj.CreateSimple(&h));
  struct Guard {
    const NvJpegApi& nj;
    nvjpegHandle_t& h;
    nvjpegJpegState_t& st;
    ~Guard() {
      if (st) nj.JpegStateDestroy(st);
      if (h) nj.Destroy(h);
    }
  } guard{nj, h, st};
  B2I_NVJ(nj.JpegStateCreate(h, &st));
  int ncomp = 0, ws[NVJPEG_MAX_COMPONENT] = {0}, hs[NVJPEG_MAX_COMPONENT] = {0};
  nvjpegChromaSubsampling_t css;
  B2I_NVJ(nj.GetImageInfo(h, bytes.data(), bytes.size(), &ncomp, &css, ws, hs));
  const int w = ws[0], hh = hs[0], c = ncomp == 1 ? 1 : 3;
  if (w <= 0 || hh <= 0) B2I_FAIL(B2I_ERR_IO, "'%s': bad JPEG geometry", path);
  const size_t pitch = (size_t)w * c, total = pitch * hh;
  DevBuf dev;
  B2I_CUDA_IO(cudaMalloc(&dev.p, total));
  nvjpegImage_t out;
  memset(&out, 0, sizeof out);
  out.channel[0] = static_cast<unsigned char*>(dev.p);
  out.pitch[0] = pitch;
  const nvjpegOutputFormat_t fmt = c == 1 ? NVJPEG_OUTPUT_Y : (order == B2I_ORDER_RGB ? NVJPEG_OUTPUT_RGBI : NVJPEG_OUTPUT_BGRI);
  B2I_NVJ(nj.Decode(h, st, bytes.data(), bytes.size(), fmt, &out, nullptr));
  B2I_CUDA_IO(cudaStreamSynchronize(nullptr));
  uint8_t* host = static_cast<uint8_t*>(malloc(total));
  if (!host) B2I_FAIL(B2I_ERR_NOMEM, "out of memory (%zu bytes)", total);
  cudaError_t e = cudaMemcpy(host, dev.p, total, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    free(host);
    B2I_FAIL(B2I_ERR_CUDA, "cudaMemcpy: %s", cudaGetErrorString(e));
  }
  *data = host;
  *W = w;
  *H = hh;
  *C = c;
  return B2I_OK;
}

// Encodes an interleaved image as a baseline JPEG (quality 1..100; 4:4:4 so that the three equal
// channels of the reference pipeline's result stay equal). One-channel images are replicated to
// three channels first, the way the reference writes its gray result (kernel.cu:210,236).
extern "C" int b2i_write_jpeg(const char* path, int order, const uint8_t* data, int W, int H, int C, size_t pitch,
                              int quality) {
  B2I_CHECK_ARG(path && data, "null argument");
  B2I_CHECK_ARG(W > 0 && H > 0 && (C == 1 || C == 3) && pitch >= (size_t)W * C, "bad geometry");
  const NvJpegApi& nj = api();
  if (!nj.ok) B2I_FAIL(B2I_ERR_UNSUPPORTED, "nvJPEG (libnvjpeg.so.12) is not available on this machine");
  if (quality <= 0 || quality > 100) quality = 95;
  const size_t row = (size_t)W * 3, total = row * H;
  std::vector<uint8_t> packed(total);
  for (int y = 0; y < H; ++y) {
    const uint8_t* s = data + (size_t)y * pitch;
    uint8_t* d = packed.data() + (size_t)y * row;
    if (C == 3) memcpy(d, s, row);
    else
      for (int x = 0; x < W; ++x) d[3 * x] = d[3 * x + 1] = d[3 * x + 2] = s[x];
  }
  DevBuf dev;
  B2I_CUDA_IO(cudaMalloc(&dev.p, total));
  B2I_CUDA_IO(cudaMemcpy(dev.p, packed.data(), total, cudaMemcpyHostToDevice));
  nvjpegHandle_t h = nullptr;
  nvjpegEncoderState_t es = nullptr;
  nvjpegEncoderParams_t ep = nullptr;
  B2I_NVJ(nj.CreateSimple(&h));
  struct Guard {
    const NvJpegApi& nj;
    nvjpegHandle_t& h;
    nvjpegEncoderState_t& es;
    nvjpegEncoderParams_t& ep;
    ~Guard() {
      if (ep) nj.EncoderParamsDestroy(ep);
      if (es) nj.EncoderStateDestroy(es);
      if (h) nj.Destroy(h);
    }
  } guard{nj, h, es, ep};
  B2I_NVJ(nj.EncoderStateCreate(h, &es, nullptr));
  B2I_NVJ(nj.EncoderParamsCreate(h, &ep, nullptr));
  B2I_NVJ(nj.EncoderParamsSetQuality(ep, quality, nullptr));
  B2I_NVJ(nj.EncoderParamsSetSamplingFactors(ep, NVJPEG_CSS_444, nullptr));
  nvjpegImage_t img;
  memset(&img, 0, sizeof img);
  img.channel[0] = static_cast<unsigned char*>(dev.p);
  img.pitch[0] = row;
  B2I_NVJ(nj.EncodeImage(h, es, ep, &img, order == B2I_ORDER_RGB ? NVJPEG_INPUT_RGBI : NVJPEG_INPUT_BGRI, W, H, nullptr));
  size_t len = 0;
  B2I_NVJ(nj.EncodeRetrieveBitstream(h, es, nullptr, &len, nullptr));
  std::vector<unsigned char> out(len);
  B2I_NVJ(nj.EncodeRetrieveBitstream(h, es, out.data(), &len, nullptr));
  B2I_CUDA_IO(cudaStreamSynchronize(nullptr));
  FILE* f = fopen(path, "wb");
  if (!f) B2I_FAIL(B2I_ERR_IO, "cannot open '%s' for writing", path);
  const bool ok = fwrite(out.data(), 1, len, f) == len;
  if (fclose(f) != 0 || !ok) B2I_FAIL(B2I_ERR_IO, "short write to '%s'", path);
  return B2I_OK;
}

// Any supported image file: JPEG (by content) or binary PPM / PGM.
extern "C" int b2i_read_image(const char* path, int order, uint8_t** data, int* W, int* H, int* C) {
  B2I_CHECK_ARG(path, "null argument");
  FILE* f = fopen(path, "rb");
  if (!f) B2I_FAIL(B2I_ERR_IO, "cannot open '%s' for reading", path);
  const int m0 = fgetc(f), m1 = fgetc(f);
  fclose(f);
  if (m0 == 0xFF && m1 == 0xD8) return b2i_read_jpeg(path, order, data, W, H, C);
  return b2i_read_pnm(path, order, data, W, H, C);
}

// By extension: .jpg / .jpeg -> JPEG (quality as given, 0 = 95), anything else -> PPM / PGM.
extern "C" int b2i_write_image(const char* path, int order, const uint8_t* data, int W, int H, int C, size_t pitch,
                               int quality) {
  B2I_CHECK_ARG(path, "null argument");
  if (ends_with_ci(path, ".jpg") || ends_with_ci(path, ".jpeg"))
    return b2i_write_jpeg(path, order, data, W, H, C, pitch, quality);
  return b2i_write_pnm(path, order, data, W, H, C, pitch);
}
