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

// cudaFuncSetAttribute (dynamic shared memory opt-in) is a PER-DEVICE setting: one flag per (call site, device), so a
// process that touches a second GPU (faiss_utils.index_init_gpu(ngpus > 1)) opts in there too.
struct DeviceOnce {
  bool done[64] = {};
  bool pending() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
    if (done[dev]) return false;
    done[dev] = true;
    return true;
  }
};

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

// Bitonic sort of 32*S keys held by one warp, ascending, element e = lane*S + s (S consecutive elements per lane, S a
// power of two), with an optional 32-bit payload.  Compare-exchanges at distance < S stay inside a lane (registers);
// larger distances are one shuffle per element.  Everything unrolls: k[] and v[] stay in registers.
template <int S, bool HAS_VAL>
__device__ __forceinline__ void warp_bitonic_sort(uint32_t (&k)[S], uint32_t (&v)[S], int lane) {
  // every compare-exchange is written as selects (no branches: the compiler otherwise emits a divergent-branch
  // sequence per exchange, which tripled the instruction count of the sort)
#pragma unroll
  for (int kk = 2; kk <= 32 * S; kk <<= 1) {
#pragma unroll
    for (int j = kk >> 1; j > 0; j >>= 1) {
      if (j >= S) {
        const int lj = j / S;
        const bool lower = (lane & lj) == 0;
#pragma unroll
        for (int s = 0; s < S; ++s) {
          const bool up = ((lane * S + s) & kk) == 0;
          const bool keep_min = lower == up;
          const uint32_t mine = k[s];
          const uint32_t ok = __shfl_xor_sync(0xffffffffu, mine, lj);
          if (HAS_VAL) {
            const uint32_t ov = __shfl_xor_sync(0xffffffffu, v[s], lj);
            const bool take = keep_min ? (ok < mine) : (ok > mine);
            k[s] = take ? ok : mine;
            v[s] = take ? ov : v[s];
          } else {
            k[s] = keep_min ? min(mine, ok) : max(mine, ok);
          }
        }
      } else {
#pragma unroll
        for (int s = 0; s < S; ++s) {
          if ((s & j) == 0) {
            const bool up = ((lane * S + s) & kk) == 0;
            const uint32_t a = k[s], b = k[s | j];
            if (HAS_VAL) {
              const bool sw = (a > b) == up;
              const uint32_t va = v[s], vb = v[s | j];
              k[s] = sw ? b : a;
              k[s | j] = sw ? a : b;
              v[s] = sw ? vb : va;
              v[s | j] = sw ? va : vb;
            } else {
              const uint32_t lo = min(a, b), hi = max(a, b);
              k[s] = up ? lo : hi;
              k[s | j] = up ? hi : lo;
            }
          }
        }
      }
    }
  }
}

// exclusive prefix of x over the lanes of a warp; total returned through `total`
__device__ __forceinline__ int warp_exclusive_scan(int x, int lane, int& total) {
  int incl = x;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += n;
  }
  total = __shfl_sync(0xffffffffu, incl, 31);
  return incl - x;
}

// Exclusive scan of in[0..n) by ONE CTA of 1024 threads through shared memory: coalesced loads into buf[0..n), every
// thread scans a contiguous chunk (odd chunk length: conflict-free banks), one block-wide scan of the chunk totals.
// Afterwards buf[r] = sum of in[0..r); returns the total.  wsum = 32 ints of shared memory.  Ends with a barrier.
__device__ __forceinline__ int cta_exclusive_scan_smem(const int32_t* __restrict__ in, int n, int* buf, int* wsum) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int r = tid; r < n; r += 1024) buf[r] = in[r];
  __syncthreads();
  const int per = ((n + 1023) / 1024) | 1;
  const int b = min(n, tid * per), e = min(n, b + per);
  int tot = 0;
  for (int r = b; r < e; ++r) tot += buf[r];
  int incl = tot;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  int run = incl - tot, total = 0;
  for (int w = 0; w < 32; ++w) { if (w < warp) run += wsum[w]; total += wsum[w]; }
  for (int r = b; r < e; ++r) { const int v = buf[r]; buf[r] = run; run += v; }
  __syncthreads();
  return total;
}

}  // namespace reidb200
