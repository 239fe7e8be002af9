// Error plumbing for the C ABI: every entry point returns 0 on success or a non-zero code, and the message is kept in
// a thread-local buffer readable through dpb_last_error().
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cuda_runtime.h>

namespace dpb {

inline char* error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}

inline int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

#define DPB_CUDA(expr)                                                                               \
  do {                                                                                               \
    cudaError_t _e = (expr);                                                                         \
    if (_e != cudaSuccess)                                                                           \
      return ::dpb::set_error((int)_e, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)

}  // namespace dpb
