// DMMA throughput vs resident warps per SM sub-partition (how many warps does the FP64 tensor path need?)
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int NACC>
__global__ void __launch_bounds__(128) k(double* out, double a, double b, int iters) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma884(c[i][0], c[i][1], a + i, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC> void run(int kwarps) {
  double* out; cudaMalloc(&out, 8 * 148 * 16 * 128);
  int iters = 8192; int blocks = 148 * kwarps;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0); k<NACC><<<blocks, 128>>>(out, 1.0000001, 1e-9, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
  }
  double tf = (double)blocks * 4 * iters * NACC * 512.0 / (best * 1e-3) / 1e12;
  printf("nacc=%d warps/SMSP=%d  %.2f TFLOP/s\n", NACC, kwarps, tf);
  cudaFree(out);
}
int main() {
  for (int kw = 1; kw <= 6; ++kw) run<20>(kw);
  for (int kw = 1; kw <= 4; ++kw) run<10>(kw);
  for (int kw = 1; kw <= 4; ++kw) run<4>(kw);
  return 0;
}
