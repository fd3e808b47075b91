// Error handling, logging and small helpers shared by the host side of the library.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

namespace lgb {

inline std::string StrFormat(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  return std::string(buf);
}

#define LGB_CUDA(expr)                                                                                   \
  do {                                                                                                   \
    cudaError_t _e = (expr);                                                                             \
    if (_e != cudaSuccess)                                                                               \
      throw std::runtime_error(lgb::StrFormat("CUDA error '%s' at %s:%d (%s)", cudaGetErrorString(_e),   \
                                              __FILE__, __LINE__, #expr));                               \
  } while (0)

#define LGB_CHECK(cond, ...)                                                  \
  do {                                                                        \
    if (!(cond)) throw std::runtime_error(lgb::StrFormat(__VA_ARGS__));       \
  } while (0)

extern int g_verbosity;
inline void LogInfo(const char* fmt, ...) {
  if (g_verbosity < 1) return;
  va_list ap;
  va_start(ap, fmt);
  fprintf(stderr, "[LightGBM-B200] [Info] ");
  vfprintf(stderr, fmt, ap);
  fprintf(stderr, "\n");
  va_end(ap);
}
inline void LogWarning(const char* fmt, ...) {
  if (g_verbosity < 0) return;
  va_list ap;
  va_start(ap, fmt);
  fprintf(stderr, "[LightGBM-B200] [Warning] ");
  vfprintf(stderr, fmt, ap);
  fprintf(stderr, "\n");
  va_end(ap);
}

template <typename T>
inline T CeilDiv(T a, T b) { return (a + b - 1) / b; }
template <typename T>
inline T RoundUp(T a, T b) { return CeilDiv(a, b) * b; }

// Simple owning device buffer.
template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { Free(); }
  void Alloc(size_t count) {
    Free();
    n = count;
    if (count) LGB_CUDA(cudaMalloc(&p, count * sizeof(T)));
  }
  void Reserve(size_t count) { if (n < count) Alloc(count); }      // grow-only: keeps the buffer between calls
  void AllocZero(size_t count) {
    Alloc(count);
    if (count) LGB_CUDA(cudaMemset(p, 0, count * sizeof(T)));
  }
  void Free() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  size_t bytes() const { return n * sizeof(T); }
};

}  // namespace lgb
