// Error plumbing shared by every translation unit of libb200gru.
// The reference aborts the process on any CUDA error (CUDA_CALL, ref:11-19);
// a library must not, so errors are recorded per thread and returned as codes.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <string>

#include "../include/b200gru.h"

namespace b200gru {

std::string &last_error_slot();

inline int fail(int code, const char *fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error_slot() = buf;
  return code;
}

}  // namespace b200gru

#define B200_REQUIRE(cond, ...)                                   \
  do {                                                            \
    if (!(cond)) return ::b200gru::fail(B200GRU_ERR_ARG, __VA_ARGS__); \
  } while (0)

#ifdef __CUDACC__
#define B200_CUDA(expr)                                                        \
  do {                                                                         \
    cudaError_t err__ = (expr);                                                \
    if (err__ != cudaSuccess)                                                  \
      return ::b200gru::fail(B200GRU_ERR_CUDA, "%s:%d: %s failed: %s (%d)",    \
                             __FILE__, __LINE__, #expr,                        \
                             cudaGetErrorString(err__), (int)err__);           \
  } while (0)
#endif
