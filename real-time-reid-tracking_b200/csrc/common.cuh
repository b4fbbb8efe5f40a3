// Shared helpers for libreidb200 (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/reidb200.h"

namespace reidb200 {

void set_error(const char* fmt, ...);

inline int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("%s: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
  return REIDB200_ECUDA;
}

#define RB_CUDA(call)                                                        \
  do {                                                                       \
    cudaError_t _e = (call);                                                 \
    if (_e != cudaSuccess) return reidb200::cuda_fail(_e, #call, __FILE__, __LINE__); \
  } while (0)

#define RB_LAUNCH_CHECK(name)                                                \
  do {                                                                       \
    cudaError_t _e = cudaGetLastError();                                     \
    if (_e != cudaSuccess) return reidb200::cuda_fail(_e, name, __FILE__, __LINE__); \
  } while (0)

#define RB_REQUIRE(cond, ...)                                                \
  do {                                                                       \
    if (!(cond)) {                                                           \
      reidb200::set_error(__VA_ARGS__);                                      \
      return REIDB200_EINVAL;                                                \
    }                                                                        \
  } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

int num_sms();

// RAII span: when profiling is enabled (reidb200_profile_enable) records a CUDA event pair on `st` around a launch.
class ProfScope {
 public:
  ProfScope(const char* name, cudaStream_t st);
  ~ProfScope();
 private:
  const char* name_;
  cudaStream_t st_;
  bool on_;
  cudaEvent_t a_, b_;
};

// Bump allocator over a caller-provided workspace.
struct Carver {
  char* base;
  size_t off;
  size_t cap;
  Carver(void* p, size_t bytes) : base(static_cast<char*>(p)), off(0), cap(bytes) {}
  template <typename T>
  T* take(size_t count) {
    off = align_up(off, 256);
    T* r = reinterpret_cast<T*>(base + off);
    off += count * sizeof(T);
    return r;
  }
  bool ok() const { return off <= cap && (base != nullptr || off == 0); }
};

struct CarveSize {
  size_t off = 0;
  template <typename T>
  void take(size_t count) {
    off = align_up(off, 256);
    off += count * sizeof(T);
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Order-preserving map float -> uint32 (ascending).
__device__ __forceinline__ uint32_t float_to_ordered(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

}  // namespace reidb200
