// Harness code, not product code: the slice of the OpenCV C++ API that the unmodified reference
// uses (kernel.cu:109-139,210,233-236; kern.cpp:32-92), because the image has Python
// cv2 but no C++ OpenCV headers. Numerics follow what OpenCV 4.13 does and are checked against
// Python cv2 by tests/test_reference_shim.py and tests/test_host_numerics.py:
//   cvtColor(BGR2GRAY)  (3735 B + 19235 G + 9798 R + 16384) >> 15
//   a*(m-b)+c           folded like cv::MatExpr into one saturate_cast<uchar>(cvRound(alpha*v+beta))
//   filter2D(-1, CV_8S) correlation, anchor centre, BORDER_REFLECT_101, saturate
// imread ignores the reference's hard-coded Windows path and loads the PPM named by
// $B2I_REF_INPUT (as BGR); imshow/waitKey are no-ops (no display on this box); imwrite writes a PPM.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#if defined(__SSSE3__) || defined(__AVX2__)
#include <tmmintrin.h>
#endif

typedef unsigned char uchar;
typedef signed char schar;

#define CV_8U 0
#define CV_8S 1
#define CV_CN_SHIFT 3
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn)-1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)

namespace cv {

enum { IMREAD_UNCHANGED = -1, IMREAD_COLOR = 1 };
enum { COLOR_BGR2GRAY = 6, COLOR_GRAY2BGR = 8 };

class Mat;
struct MatExpr {  // alpha * src + beta, evaluated lazily with one saturation (like cv::MatExpr)
  const Mat* src;
  double alpha, beta;
};

class Mat {
 public:
  uchar* data = nullptr;
  int rows = 0, cols = 0;

  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  void create(int r, int c, int type) {
    rows = r;
    cols = c;
    type_ = type;
    const size_t n = (size_t)(r > 0 ? r : 0) * (c > 0 ? c : 0) * channels();
    buf_ = n ? std::shared_ptr<uchar>((uchar*)malloc(n), free) : std::shared_ptr<uchar>();
    data = buf_.get();
  }
  int type() const { return type_; }
  int depth() const { return type_ & 7; }
  int channels() const { return (type_ >> CV_CN_SHIFT) + 1; }
  size_t total() const { return (size_t)rows * cols; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  template <typename T>
  T& at(int r, int c) { return reinterpret_cast<T*>(data)[(size_t)r * cols * channels() + c]; }
  Mat& operator=(const MatExpr& e);

 private:
  int type_ = 0;
  std::shared_ptr<uchar> buf_;
};

inline MatExpr operator-(const Mat& m, int b) { return {&m, 1.0, (double)-b}; }
inline MatExpr operator+(const Mat& m, int b) { return {&m, 1.0, (double)b}; }
inline MatExpr operator*(int a, const Mat& m) { return {&m, (double)a, 0.0}; }
inline MatExpr operator*(int a, const MatExpr& e) { return {e.src, e.alpha * a, e.beta * a}; }
inline MatExpr operator*(double a, const MatExpr& e) { return {e.src, e.alpha * a, e.beta * a}; }
inline MatExpr operator+(const MatExpr& e, int b) { return {e.src, e.alpha, e.beta + b}; }
inline MatExpr operator-(const MatExpr& e, int b) { return {e.src, e.alpha, e.beta - b}; }

inline Mat& Mat::operator=(const MatExpr& e) {
  const Mat src = *e.src;  // may alias *this (tempImage = 3*(tempImage-128)+128, kern.cpp:74)
  Mat out(src.rows, src.cols, src.type());
  const size_t n = src.total() * src.channels();
  const float a = (float)e.alpha, b = (float)e.beta;
  for (size_t i = 0; i < n; ++i) {
    const long v = lrintf(src.data[i] * a + b);
    out.data[i] = (uchar)(v < 0 ? 0 : v > 255 ? 255 : v);
  }
  *this = out;
  return *this;
}

inline Mat imread(const std::string&, int = IMREAD_COLOR) {
  const char* path = getenv("B2I_REF_INPUT");
  Mat m;
  if (!path) return m;
  FILE* f = fopen(path, "rb");
  if (!f) return m;
  int w = 0, h = 0, maxv = 0;
  char magic[3] = {0, 0, 0};
  if (fscanf(f, "%2s %d %d %d", magic, &w, &h, &maxv) != 4 || strcmp(magic, "P6") || maxv != 255) {
    fclose(f);
    return m;
  }
  fgetc(f);
  m.create(h, w, CV_8UC3);
  const size_t n = (size_t)w * h * 3;
  if (fread(m.data, 1, n, f) != n) {
    fclose(f);
    return Mat();
  }
  fclose(f);
  for (size_t i = 0; i < n; i += 3) {  // PPM is RGB, cv::imread delivers BGR
    const uchar t = m.data[i];
    m.data[i] = m.data[i + 2];
    m.data[i + 2] = t;
  }
  return m;
}

inline bool imwrite(const std::string& name, const Mat& m) {
  const char* dir = getenv("B2I_REF_OUTDIR");
  const std::string path = (dir ? std::string(dir) + "/" : std::string()) + name + ".ppm";
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) return false;
  fprintf(f, "P6\n%d %d\n255\n", m.cols, m.rows);
  std::vector<uchar> row((size_t)m.cols * 3);
  for (int y = 0; y < m.rows; ++y) {
    const uchar* s = m.data + (size_t)y * m.cols * 3;
    for (int x = 0; x < m.cols; ++x) {
      row[3 * x] = s[3 * x + 2];
      row[3 * x + 1] = s[3 * x + 1];
      row[3 * x + 2] = s[3 * x];
    }
    fwrite(row.data(), 1, row.size(), f);
  }
  fclose(f);
  return true;
}

inline void imshow(const std::string&, const Mat&) {}  // no display on this box
inline int waitKey(int = 0) { return -1; }
inline void destroyAllWindows() {}

inline void cvtColor(const Mat& src, Mat& dst, int code) {
  if (code == COLOR_BGR2GRAY) {
    Mat out(src.rows, src.cols, CV_8UC1);
    const size_t n = src.total();
    for (size_t i = 0; i < n; ++i) {
      const uchar* p = src.data + 3 * i;
      out.data[i] = (uchar)((3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15);
    }
    dst = out;
  } else {  // COLOR_GRAY2BGR
    Mat out(src.rows, src.cols, CV_8UC3);
    const size_t n = src.total();
    size_t i = 0;
#if defined(__SSSE3__) || defined(__AVX2__)
    // 16 pixels per step with byte shuffles, as OpenCV's own SIMD path does: a scalar loop here
    // would charge the reference for the shim's slowness (kernel.cu:210 is inside its timed window)
    const __m128i m0 = _mm_setr_epi8(0, 0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 3, 4, 4, 4, 5);
    const __m128i m1 = _mm_setr_epi8(5, 5, 6, 6, 6, 7, 7, 7, 8, 8, 8, 9, 9, 9, 10, 10);
    const __m128i m2 = _mm_setr_epi8(10, 11, 11, 11, 12, 12, 12, 13, 13, 13, 14, 14, 14, 15, 15, 15);
    for (; i + 16 <= n; i += 16) {
      const __m128i g = _mm_loadu_si128((const __m128i*)(src.data + i));
      _mm_storeu_si128((__m128i*)(out.data + 3 * i), _mm_shuffle_epi8(g, m0));
      _mm_storeu_si128((__m128i*)(out.data + 3 * i + 16), _mm_shuffle_epi8(g, m1));
      _mm_storeu_si128((__m128i*)(out.data + 3 * i + 32), _mm_shuffle_epi8(g, m2));
    }
#endif
    for (; i < n; ++i) out.data[3 * i] = out.data[3 * i + 1] = out.data[3 * i + 2] = src.data[i];
    dst = out;
  }
}

inline void filter2D(const Mat& src, Mat& dst, int /*ddepth*/, const Mat& kernel) {
  const int k = kernel.rows, r = k / 2, W = src.cols, H = src.rows;
  const schar* w = reinterpret_cast<const schar*>(kernel.data);
  Mat out(H, W, CV_8UC1);
  auto fold = [](int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
    return p;
  };
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) {
      int s = 0;
      for (int j = 0; j < k; ++j) {
        const uchar* row = src.data + (size_t)fold(y + j - r, H) * W;
        for (int i = 0; i < k; ++i) s += w[j * k + i] * row[fold(x + i - r, W)];
      }
      out.data[(size_t)y * W + x] = (uchar)(s < 0 ? 0 : s > 255 ? 255 : s);
    }
  dst = out;
}

}  // namespace cv
