// Shared helpers for libfdlp_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/fdlp_b200.h"

namespace fdlp {

extern thread_local char g_err[512];

inline int set_err(int code, const char* fmt, const char* a = "", long b = 0, long c = 0) {
  snprintf(g_err, sizeof(g_err), fmt, a, b, c);
  return code;
}

#define FDLP_CUDA(call)                                                                      \
  do {                                                                                       \
    cudaError_t e__ = (call);                                                                \
    if (e__ != cudaSuccess) {                                                                \
      snprintf(fdlp::g_err, sizeof(fdlp::g_err), "%s failed: %s (%s:%d)", #call,             \
               cudaGetErrorString(e__), __FILE__, __LINE__);                                 \
      return FDLP_E_CUDA;                                                                    \
    }                                                                                        \
  } while (0)

template <typename T>
struct cplx {
  T x, y;
};

template <typename T>
__device__ __forceinline__ cplx<T> cadd(cplx<T> a, cplx<T> b) { return {a.x + b.x, a.y + b.y}; }
template <typename T>
__device__ __forceinline__ cplx<T> csub(cplx<T> a, cplx<T> b) { return {a.x - b.x, a.y - b.y}; }
template <typename T>
__device__ __forceinline__ cplx<T> cmul(cplx<T> a, cplx<T> b) {
  return {a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x};
}
// multiply by -j : (x + jy)(-j) = y - jx
template <typename T>
__device__ __forceinline__ cplx<T> mulmj(cplx<T> a) { return {a.y, -a.x}; }
template <typename T>
__device__ __forceinline__ cplx<T> cscale(cplx<T> a, T s) { return {a.x * s, a.y * s}; }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// np.pad(..., 'reflect') index (no edge repeat, multi-bounce): maps any integer q onto [0, T), T >= 2.
__device__ __forceinline__ int64_t reflect_index(int64_t q, int64_t T) {
  const int64_t P = 2 * (T - 1);
  q %= P;
  if (q < 0) q += P;
  return q < T ? q : P - q;
}

}  // namespace fdlp
