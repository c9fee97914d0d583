// File I/O and synthetic inputs.
//
// The reference decodes / encodes JPEG through OpenCV on rank 0 (kernel.cu:110,236;
// kern.cpp:33,92). JPEG is lossy and OpenCV's C++ API is not on this image, so the framework
// reads and writes the same bytes losslessly: binary PPM (P6) / PGM (P5) and headerless raw
// in the reference's scatter layout (interleaved u8, W*C bytes per row: kernel.cu:135).
#include <cctype>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "error.hpp"

namespace {

// Skips blanks and '#' comments, then reads one decimal integer of a PNM header.
int pnm_int(FILE* f, int* out) {
  int ch = fgetc(f);
  for (;;) {
    while (ch != EOF && isspace(ch)) ch = fgetc(f);
    if (ch == '#') {
      while (ch != EOF && ch != '\n') ch = fgetc(f);
      continue;
    }
    break;
  }
  if (ch == EOF || !isdigit(ch)) return -1;
  long v = 0;
  while (ch != EOF && isdigit(ch)) {
    v = v * 10 + (ch - '0');
    if (v > (1L << 30)) return -1;
    ch = fgetc(f);
  }
  // exactly one whitespace byte follows the last header field; the caller relies on that
  *out = (int)v;
  return 0;
}

void swap_rb_rows(uint8_t* data, int W, int H, size_t pitch) {
  for (int y = 0; y < H; ++y) {
    uint8_t* row = data + (size_t)y * pitch;
    for (int x = 0; x < W; ++x) {
      uint8_t t = row[3 * x];
      row[3 * x] = row[3 * x + 2];
      row[3 * x + 2] = t;
    }
  }
}

}  // namespace

extern "C" void b2i_free(void* p) { free(p); }

extern "C" int b2i_read_pnm(const char* path, int order, uint8_t** data, int* W, int* H,
                            int* C) {
  B2I_CHECK_ARG(path && data && W && H && C, "null argument");
  FILE* f = fopen(path, "rb");
  if (!f) B2I_FAIL(B2I_ERR_IO, "cannot open '%s' for reading", path);
  int m0 = fgetc(f), m1 = fgetc(f);
  int channels = (m0 == 'P' && m1 == '6') ? 3 : (m0 == 'P' && m1 == '5') ? 1 : 0;
  int w = 0, h = 0, maxv = 0;
  if (!channels || pnm_int(f, &w) || pnm_int(f, &h) || pnm_int(f, &maxv)) {
    fclose(f);
    B2I_FAIL(B2I_ERR_IO, "'%s' is not a binary PPM (P6) or PGM (P5) file", path);
  }
  if (maxv != 255 || w <= 0 || h <= 0) {
    fclose(f);
    B2I_FAIL(B2I_ERR_UNSUPPORTED, "'%s': only 8-bit images are supported (maxval %d, %dx%d)",
             path, maxv, w, h);
  }
  size_t n = (size_t)w * h * channels;
  uint8_t* buf = (uint8_t*)malloc(n ? n : 1);
  if (!buf) {
    fclose(f);
    B2I_FAIL(B2I_ERR_NOMEM, "out of memory reading '%s' (%zu bytes)", path, n);
  }
  size_t got = fread(buf, 1, n, f);
  fclose(f);
  if (got != n) {
    free(buf);
    B2I_FAIL(B2I_ERR_IO, "'%s' is truncated: %zu of %zu pixel bytes", path, got, n);
  }
  if (channels == 3 && order == B2I_ORDER_BGR) swap_rb_rows(buf, w, h, (size_t)w * 3);
  *data = buf;
  *W = w;
  *H = h;
  *C = channels;
  return B2I_OK;
}

extern "C" int b2i_write_pnm(const char* path, int order, const uint8_t* data, int W, int H,
                             int C, size_t pitch) {
  B2I_CHECK_ARG(path && data, "null argument");
  B2I_CHECK_ARG(C == 1 || C == 3, "PNM output needs 1 or 3 channels, got %d", C);
  B2I_CHECK_ARG(W > 0 && H > 0 && pitch >= (size_t)W * C, "bad geometry %dx%dx%d pitch %zu", W,
                H, C, pitch);
  FILE* f = fopen(path, "wb");
  if (!f) B2I_FAIL(B2I_ERR_IO, "cannot open '%s' for writing", path);
  fprintf(f, "P%c\n%d %d\n255\n", C == 3 ? '6' : '5', W, H);
  const size_t rowb = (size_t)W * C;
  uint8_t* tmp = nullptr;
  if (C == 3 && order == B2I_ORDER_BGR) tmp = (uint8_t*)malloc(rowb);
  bool ok = true;
  for (int y = 0; y < H && ok; ++y) {
    const uint8_t* row = data + (size_t)y * pitch;
    if (tmp) {
      for (int x = 0; x < W; ++x) {
        tmp[3 * x] = row[3 * x + 2];
        tmp[3 * x + 1] = row[3 * x + 1];
        tmp[3 * x + 2] = row[3 * x];
      }
      row = tmp;
    }
    ok = fwrite(row, 1, rowb, f) == rowb;
  }
  free(tmp);
  ok = (fclose(f) == 0) && ok;
  if (!ok) B2I_FAIL(B2I_ERR_IO, "short write to '%s'", path);
  return B2I_OK;
}

extern "C" int b2i_read_raw(const char* path, int W, int H, int C, uint8_t** data) {
  B2I_CHECK_ARG(path && data, "null argument");
  B2I_CHECK_ARG(W > 0 && H > 0 && (C == 1 || C == 3), "raw input needs --width/--height (got "
                "%dx%dx%d)", W, H, C);
  FILE* f = fopen(path, "rb");
  if (!f) B2I_FAIL(B2I_ERR_IO, "cannot open '%s' for reading", path);
  size_t n = (size_t)W * H * C;
  uint8_t* buf = (uint8_t*)malloc(n);
  if (!buf) {
    fclose(f);
    B2I_FAIL(B2I_ERR_NOMEM, "out of memory reading '%s' (%zu bytes)", path, n);
  }
  size_t got = fread(buf, 1, n, f);
  fclose(f);
  if (got != n) {
    free(buf);
    B2I_FAIL(B2I_ERR_IO, "'%s' holds %zu bytes, %dx%dx%d needs %zu", path, got, W, H, C, n);
  }
  *data = buf;
  return B2I_OK;
}

extern "C" int b2i_write_raw(const char* path, const uint8_t* data, int W, int H, int C,
                             size_t pitch) {
  B2I_CHECK_ARG(path && data, "null argument");
  B2I_CHECK_ARG(W > 0 && H > 0 && C > 0 && pitch >= (size_t)W * C, "bad geometry");
  FILE* f = fopen(path, "wb");
  if (!f) B2I_FAIL(B2I_ERR_IO, "cannot open '%s' for writing", path);
  bool ok = true;
  for (int y = 0; y < H && ok; ++y)
    ok = fwrite(data + (size_t)y * pitch, 1, (size_t)W * C, f) == (size_t)W * C;
  ok = (fclose(f) == 0) && ok;
  if (!ok) B2I_FAIL(B2I_ERR_IO, "short write to '%s'", path);
  return B2I_OK;
}

// Synthetic test card. There is no network and the reference's input (C:\BA_Tests\PKA_img\3.jpg,
// kernel.cu:110) is not available, so benchmarks and tests use this: smooth gradients (so
// contrast does not saturate everything), hard edges and rings (so emboss / Sobel have
// structure) and hashed noise (so every byte value and every LUT entry is exercised).
extern "C" int b2i_synth_image(uint8_t* dst, int W, int H, int C, size_t pitch, uint32_t seed) {
  B2I_CHECK_ARG(dst && W > 0 && H > 0 && (C == 1 || C == 3) && pitch >= (size_t)W * C,
                "bad geometry %dx%dx%d pitch %zu", W, H, C, pitch);
  for (int y = 0; y < H; ++y) {
    uint8_t* row = dst + (size_t)y * pitch;
    for (int x = 0; x < W; ++x) {
      uint32_t h = (uint32_t)x * 0x9E3779B1u ^ (uint32_t)y * 0x85EBCA77u ^ seed * 0xC2B2AE3Du;
      h ^= h >> 15;
      h *= 0x2C1B3C6Du;
      h ^= h >> 12;
      h *= 0x297A2D39u;
      h ^= h >> 15;
      const int noise = (int)(h & 63) - 32;
      const int grad = (int)(((int64_t)x * 255) / (W > 1 ? W - 1 : 1));
      const int vgrad = (int)(((int64_t)y * 255) / (H > 1 ? H - 1 : 1));
      const int checker = (((x >> 5) ^ (y >> 5)) & 1) ? 40 : -40;
      const long dx = x - W / 2, dy = y - H / 2;
      const int ring = (((dx * dx + dy * dy) >> 12) & 1) ? 25 : -25;
      for (int c = 0; c < C; ++c) {
        int v;
        switch (c) {
          case 0: v = grad + checker + noise; break;
          case 1: v = vgrad + ring + (noise >> 1); break;
          default: v = (grad + vgrad) / 2 - checker + (int)((h >> 8) & 31) - 16; break;
        }
        if ((h >> 20) % 97 == 0) v = (int)((h >> 24) & 255);  // salt: full value range
        row[(size_t)x * C + c] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
      }
    }
  }
  return B2I_OK;
}
