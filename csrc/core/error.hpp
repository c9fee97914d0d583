// Error plumbing shared by every translation unit of libb200img.
// The reference checks cudaError_t by hand and returns 1 without telling its peers
// (kernel.cu:148-151,197-200); here every failure carries file:line text and a status code.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <string>

#include "b200img.h"

namespace b2i {

std::string& last_error_slot();

inline int fail(int code, const char* file, int line, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  char full[1200];
  snprintf(full, sizeof full, "%s (%s:%d)", buf, file, line);
  last_error_slot() = full;
  return code;
}

}  // namespace b2i

#define B2I_FAIL(code, ...) return ::b2i::fail((code), __FILE__, __LINE__, __VA_ARGS__)
#define B2I_CHECK_ARG(cond, ...)                                   \
  do {                                                             \
    if (!(cond)) B2I_FAIL(B2I_ERR_ARG, __VA_ARGS__);               \
  } while (0)
#define B2I_TRY(expr)                 \
  do {                                \
    int b2i_rc_ = (expr);             \
    if (b2i_rc_ != B2I_OK) return b2i_rc_; \
  } while (0)
