// Developer probe: throughput of the packed FP32 instructions of sm_100 (FFMA2 / FADD2 / FMUL2: fma.rn.f32x2 ...)
// against scalar FFMA.   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/ffma2_probe tools/ffma2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) probe(float* out, int iters) {
  float2 a[8], b = make_float2(1.0001f, 0.9999f), c = make_float2(1e-7f, -1e-7f);
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = make_float2(threadIdx.x * 1e-3f + i, blockIdx.x * 1e-3f - i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) { a[i].x = fmaf(a[i].x, b.x, c.x); a[i].y = fmaf(a[i].y, b.y, c.y); }   // 2 scalar FFMA
      else a[i] = __ffma2_rn(a[i], b, c);                                                   // 1 FFMA2
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  float* out;
  const int blocks = sms * 8, iters = 20000;
  cudaMalloc(&out, (size_t)blocks * 256 * 4);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int mode = 0; mode < 2; ++mode) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (mode == 0) probe<0><<<blocks, 256>>>(out, iters); else probe<1><<<blocks, 256>>>(out, iters);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      const double flops = 2.0 * 2 * 8 * (double)iters * blocks * 256;
      if (rep) printf("%s: %.3f ms, %.1f TFLOP/s fp32\n", mode ? "FFMA2 (fma.rn.f32x2)" : "2 x FFMA", ms, flops / ms * 1e-9);
    }
  }
  return 0;
}
