// Filter-chain construction: token parser, presets and LUT builders.
//
// The reference has no configuration at all — contrast 3.5 / 3, the emboss weights and the
// 3x3-vs-5x5 switch are literals (kernel.cu:50,71-82,195; kern.cpp:62-74). Here the same
// pipelines are data: a b2i_chain_t that the host reference, the CUDA planner and the CLI share.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "error.hpp"

namespace b2i {
std::string& last_error_slot() {
  thread_local std::string s;
  return s;
}
}  // namespace b2i

extern "C" const char* b2i_last_error(void) { return b2i::last_error_slot().c_str(); }
extern "C" const char* b2i_version(void) { return B2I_VERSION; }

// kernel.cu:56 — `clamp(contrast * (v - 128) + 128)` stored to uchar. nvcc contracts the
// multiply-add to one FFMA, hence fmaf; the store truncates.
extern "C" void b2i_lut_contrast_ref(float factor, uint8_t* lut) {
  for (int v = 0; v < 256; ++v) {
    float f = fmaf(factor, (float)(v - 128), 128.0f);
    if (f < 0.f) f = 0.f;
    if (f > 255.f) f = 255.f;
    lut[v] = (uint8_t)f;
  }
}

// kern.cpp:74 — `a*(m-128)+128` is folded by cv::MatExpr into one convertTo(alpha=a,
// beta=128-128a) with a single saturate_cast<uchar>(cvRound(.)) at the end.
extern "C" void b2i_lut_contrast_cv(double factor, uint8_t* lut) {
  const float alpha = (float)factor, beta = (float)(128.0 - 128.0 * factor);
  for (int v = 0; v < 256; ++v) {
    long r = lrintf((float)v * alpha + beta);  // round-half-even like cvRound
    lut[v] = (uint8_t)(r < 0 ? 0 : r > 255 ? 255 : r);
  }
}

namespace {

void set_name(b2i_op_t& op, const std::string& s) {
  memset(op.name, 0, sizeof op.name);
  strncpy(op.name, s.c_str(), sizeof op.name - 1);
}

b2i_op_t make_op(int kind, const std::string& name) {
  b2i_op_t op;
  memset(&op, 0, sizeof op);
  op.kind = kind;
  set_name(op, name);
  return op;
}

b2i_op_t make_stencil(const std::string& name, int k, const int* w, int shift, int bias,
                      int border) {
  b2i_op_t op = make_op(B2I_OP_STENCIL, name);
  op.ksize = k;
  op.shift = shift;
  op.bias = bias;
  op.border = border;
  for (int i = 0; i < k * k; ++i) op.weights[i] = (int8_t)w[i];
  return op;
}

int parse_border(const std::string& s, int* out) {
  if (s == "ref") *out = B2I_BORDER_REF;
  else if (s == "reflect101" || s == "reflect") *out = B2I_BORDER_REFLECT101;
  else if (s == "replicate") *out = B2I_BORDER_REPLICATE;
  else if (s == "zero") *out = B2I_BORDER_ZERO;
  else B2I_FAIL(B2I_ERR_ARG, "unknown border '%s' (ref|reflect101|replicate|zero)", s.c_str());
  return B2I_OK;
}

const int kEmboss3[9] = {-2, -1, 0, -1, 1, 1, 0, 1, 2};  // kernel.cu:71-75, kern.cpp:62-71
const int kEmboss5[25] = {4, 0, 0, 0, 0, 0, 4, 0, 0, 0, 0, 0, 1,
                          0, 0, 0, 0, 0, -4, 0, 0, 0, 0, 0, -4};  // kernel.cu:76-82
const int kSharpen[9] = {0, -1, 0, -1, 5, -1, 0, -1, 0};
const int kLaplace[9] = {0, 1, 0, 1, -4, 1, 0, 1, 0};

int parse_token(const std::string& tok_in, int default_border, b2i_chain_t* c) {
  std::string tok = tok_in, border_s, arg;
  size_t at = tok.find('@');
  if (at != std::string::npos) {
    border_s = tok.substr(at + 1);
    tok = tok.substr(0, at);
  }
  size_t colon = tok.find(':');
  if (colon != std::string::npos) {
    arg = tok.substr(colon + 1);
    tok = tok.substr(0, colon);
  }
  B2I_CHECK_ARG(c->n_ops < B2I_MAX_OPS, "chain longer than %d ops", B2I_MAX_OPS);
  b2i_op_t& op = c->ops[c->n_ops];
  auto border_for = [&](int own_default, int* out) -> int {
    if (!border_s.empty()) return parse_border(border_s, out);
    *out = default_border >= 0 ? default_border : own_default;
    return B2I_OK;
  };
  auto need_arg = [&]() -> int {
    B2I_CHECK_ARG(!arg.empty(), "filter '%s' needs an argument, e.g. %s:3.5", tok.c_str(),
                  tok.c_str());
    return B2I_OK;
  };

  if (tok == "gray" || tok == "gray_ref" || tok == "grayscale") {
    op = make_op(B2I_OP_GRAY_REF, "gray_ref");
  } else if (tok == "gray_cv") {
    op = make_op(B2I_OP_GRAY_CV, "gray_cv");
  } else if (tok == "expand" || tok == "bgr" || tok == "gray2bgr") {
    op = make_op(B2I_OP_EXPAND, "expand");
  } else if (tok == "invert") {
    op = make_op(B2I_OP_LUT, "invert");
    for (int v = 0; v < 256; ++v) op.lut[v] = (uint8_t)(255 - v);
  } else if (tok == "contrast" || tok == "contrast_ref") {
    float f = arg.empty() ? 3.5f : strtof(arg.c_str(), nullptr);  // kernel.cu:50
    op = make_op(B2I_OP_LUT, "contrast");
    b2i_lut_contrast_ref(f, op.lut);
  } else if (tok == "contrast_cv") {
    double f = arg.empty() ? 3.0 : strtod(arg.c_str(), nullptr);  // kern.cpp:74
    op = make_op(B2I_OP_LUT, "contrast_cv");
    b2i_lut_contrast_cv(f, op.lut);
  } else if (tok == "brightness") {
    B2I_TRY(need_arg());
    int d = atoi(arg.c_str());
    op = make_op(B2I_OP_LUT, "brightness");
    for (int v = 0; v < 256; ++v) {
      int r = v + d;
      op.lut[v] = (uint8_t)(r < 0 ? 0 : r > 255 ? 255 : r);
    }
  } else if (tok == "threshold") {
    B2I_TRY(need_arg());
    int t = atoi(arg.c_str());
    op = make_op(B2I_OP_LUT, "threshold");
    for (int v = 0; v < 256; ++v) op.lut[v] = v > t ? 255 : 0;
  } else if (tok == "gamma") {
    B2I_TRY(need_arg());
    double g = strtod(arg.c_str(), nullptr);
    B2I_CHECK_ARG(g > 0, "gamma must be positive");
    op = make_op(B2I_OP_LUT, "gamma");
    for (int v = 0; v < 256; ++v) op.lut[v] = (uint8_t)lrint(255.0 * pow(v / 255.0, 1.0 / g));
  } else if (tok == "emboss3" || tok == "emboss") {
    int b;
    B2I_TRY(border_for(B2I_BORDER_REF, &b));
    op = make_stencil("emboss3", 3, kEmboss3, 0, 0, b);
  } else if (tok == "emboss5") {
    int b;
    B2I_TRY(border_for(B2I_BORDER_REF, &b));
    op = make_stencil("emboss5", 5, kEmboss5, 0, 0, b);
  } else if (tok == "sharpen") {
    int b;
    B2I_TRY(border_for(B2I_BORDER_REFLECT101, &b));
    op = make_stencil("sharpen", 3, kSharpen, 0, 0, b);
  } else if (tok == "laplace" || tok == "laplacian") {
    int b;
    B2I_TRY(border_for(B2I_BORDER_REFLECT101, &b));
    op = make_stencil("laplace", 3, kLaplace, 0, 0, b);
  } else if (tok == "gauss5" || tok == "blur" || tok == "gauss" || tok == "gaussian") {
    // cv::GaussianBlur(5x5, sigma 0) on u8: [1 4 6 4 1]^2 / 256 with round-half-up.
    int b, w[25];
    const int k1[5] = {1, 4, 6, 4, 1};
    for (int i = 0; i < 25; ++i) w[i] = k1[i / 5] * k1[i % 5];
    B2I_TRY(border_for(B2I_BORDER_REFLECT101, &b));
    op = make_stencil("gauss5", 5, w, 8, 128, b);
  } else if (tok == "gauss3") {
    int b, w[9];
    const int k1[3] = {1, 2, 1};
    for (int i = 0; i < 9; ++i) w[i] = k1[i / 3] * k1[i % 3];
    B2I_TRY(border_for(B2I_BORDER_REFLECT101, &b));
    op = make_stencil("gauss3", 3, w, 4, 8, b);
  } else if (tok == "sobel" || tok == "edge" || tok == "edges") {
    int b;
    B2I_TRY(border_for(B2I_BORDER_REFLECT101, &b));
    op = make_op(B2I_OP_SOBEL, "sobel");
    op.border = b;
    op.ksize = 3;
  } else if (tok == "filter3" || tok == "filter5") {
    const int k = tok == "filter3" ? 3 : 5;
    B2I_TRY(need_arg());
    std::vector<int> vals;
    size_t pos = 0;
    while (pos <= arg.size()) {
      size_t slash = arg.find('/', pos);
      if (slash == std::string::npos) slash = arg.size();
      vals.push_back(atoi(arg.substr(pos, slash - pos).c_str()));
      pos = slash + 1;
    }
    B2I_CHECK_ARG((int)vals.size() == k * k || (int)vals.size() == k * k + 1,
                  "%s needs %d weights (optionally one more value: the right shift)",
                  tok.c_str(), k * k);
    for (int i = 0; i < k * k; ++i)
      B2I_CHECK_ARG(vals[i] >= -128 && vals[i] <= 127, "weight %d out of int8 range", vals[i]);
    int shift = (int)vals.size() == k * k + 1 ? vals[k * k] : 0;
    B2I_CHECK_ARG(shift >= 0 && shift <= 16, "shift %d out of range", shift);
    int b;
    B2I_TRY(border_for(B2I_BORDER_REFLECT101, &b));
    op = make_stencil(tok, k, vals.data(), shift, shift ? 1 << (shift - 1) : 0, b);
  } else {
    B2I_FAIL(B2I_ERR_ARG, "unknown filter '%s'", tok_in.c_str());
  }
  c->n_ops++;
  return B2I_OK;
}

}  // namespace

extern "C" int b2i_chain_parse(const char* spec, int default_border, int order,
                               b2i_chain_t* out) {
  B2I_CHECK_ARG(spec && out, "null argument");
  B2I_CHECK_ARG(order == B2I_ORDER_BGR || order == B2I_ORDER_RGB, "bad channel order %d", order);
  memset(out, 0, sizeof *out);
  out->order = order;
  std::string s(spec);
  size_t pos = 0;
  while (pos <= s.size()) {
    size_t comma = s.find(',', pos);
    if (comma == std::string::npos) comma = s.size();
    std::string tok = s.substr(pos, comma - pos);
    // trim blanks
    while (!tok.empty() && (tok.front() == ' ' || tok.front() == '\t')) tok.erase(tok.begin());
    while (!tok.empty() && (tok.back() == ' ' || tok.back() == '\t')) tok.pop_back();
    if (!tok.empty()) B2I_TRY(parse_token(tok, default_border, out));
    pos = comma + 1;
  }
  B2I_CHECK_ARG(out->n_ops > 0, "empty filter chain '%s'", spec);
  return B2I_OK;
}

extern "C" int b2i_chain_preset(const char* name, int order, b2i_chain_t* out) {
  B2I_CHECK_ARG(name && out, "null argument");
  std::string n(name);
  if (n == "kernel_cu" || n == "kernel.cu" || n == "reference")
    return b2i_chain_parse("gray_ref,contrast:3.5,emboss3@ref,expand", -1, order, out);
  if (n == "kernel_cu_5x5")  // the smallEmboss=false branch the reference never launches
    return b2i_chain_parse("gray_ref,contrast:3.5,emboss5@ref,expand", -1, order, out);
  if (n == "kern_cpp" || n == "kern.cpp" || n == "opencv")
    return b2i_chain_parse("gray_cv,contrast_cv:3,emboss3@reflect101,expand", -1, order, out);
  B2I_FAIL(B2I_ERR_ARG, "unknown preset '%s' (kernel_cu|kernel_cu_5x5|kern_cpp)", name);
}

extern "C" int b2i_chain_out_channels(const b2i_chain_t* chain, int in_channels) {
  B2I_CHECK_ARG(chain, "null chain");
  B2I_CHECK_ARG(in_channels == 1 || in_channels == 3, "images have 1 or 3 channels, got %d",
                in_channels);
  int c = in_channels;
  for (int i = 0; i < chain->n_ops; ++i) {
    const b2i_op_t& op = chain->ops[i];
    switch (op.kind) {
      case B2I_OP_GRAY_REF:
      case B2I_OP_GRAY_CV:
        B2I_CHECK_ARG(c == 3, "op %d (%s) needs 3-channel input, has %d", i, op.name, c);
        c = 1;
        break;
      case B2I_OP_EXPAND:
        B2I_CHECK_ARG(c == 1, "op %d (expand) needs 1-channel input, has %d", i, c);
        c = 3;
        break;
      case B2I_OP_LUT:
      case B2I_OP_SOBEL:
        break;
      case B2I_OP_STENCIL:
        B2I_CHECK_ARG(op.ksize == 3 || op.ksize == 5, "op %d: stencil size %d not 3 or 5", i,
                      op.ksize);
        break;
      default:
        B2I_FAIL(B2I_ERR_ARG, "op %d has unknown kind %d", i, op.kind);
    }
  }
  return c;
}

extern "C" int b2i_chain_halo(const b2i_chain_t* chain) {
  if (!chain) return 0;
  int r = 0;
  for (int i = 0; i < chain->n_ops; ++i) {
    const b2i_op_t& op = chain->ops[i];
    if (op.kind == B2I_OP_STENCIL) r += op.ksize / 2;
    if (op.kind == B2I_OP_SOBEL) r += 1;
  }
  return r;
}

extern "C" int b2i_chain_describe(const b2i_chain_t* chain, char* buf, size_t buflen) {
  B2I_CHECK_ARG(chain && buf && buflen > 0, "null argument");
  static const char* bn[] = {"ref", "reflect101", "replicate", "zero"};
  std::string s;
  for (int i = 0; i < chain->n_ops; ++i) {
    const b2i_op_t& op = chain->ops[i];
    if (i) s += ",";
    s += op.name[0] ? op.name : "?";
    if (op.kind == B2I_OP_STENCIL || op.kind == B2I_OP_SOBEL) {
      s += "@";
      s += bn[op.border & 3];
    }
  }
  snprintf(buf, buflen, "%s", s.c_str());
  return B2I_OK;
}
