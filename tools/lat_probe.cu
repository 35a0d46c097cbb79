// Dependent-issue latency probe for B200 (sm_100a): one warp per SM sub-partition, serial chains.
// Prints cycles per dependent DFMA, DADD, 64-bit shuffle (+DADD), FSEL pair, and the rcp seed.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lat_probe lat_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int N = 2048;

template <int MODE>
__global__ void probe(double* out, long long* cyc, double a, double b) {
  double x = a + threadIdx.x, y = b;
  long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) {
    if (MODE == 0) x = fma(x, y, a);
    else if (MODE == 1) x = x + y;
    else if (MODE == 2) x += __shfl_xor_sync(0xffffffffu, x, 1);
    else if (MODE == 3) x = __shfl_sync(0xffffffffu, x, (threadIdx.x + 31) & 31);
    else if (MODE == 4) { asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(x)); }
    else if (MODE == 5) x = x * y;
    else if (MODE == 6) { float f = (float)x; f = fmaf(f, 1.0001f, 0.5f); x = f; }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static double run(int warps) {
  double* out; long long* cyc;
  cudaMalloc(&out, sizeof(double) * 1024); cudaMalloc(&cyc, sizeof(long long));
  probe<MODE><<<1, 32 * warps>>>(out, cyc, 1.0000001, 0.9999999);
  probe<MODE><<<1, 32 * warps>>>(out, cyc, 1.0000001, 0.9999999);
  long long h; cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(out); cudaFree(cyc);
  return (double)h / N;
}

int main() {
  printf("{");
  for (int w : {1, 4, 8, 12, 16}) {
    printf("\"w%d\": {\"dfma\": %.1f, \"dadd\": %.1f, \"shflxor64_dadd\": %.1f, \"shfl64\": %.1f, \"rcp64h\": %.1f, \"dmul\": %.1f, \"cvt_ffma_cvt\": %.1f}%s",
           w, run<0>(w), run<1>(w), run<2>(w), run<3>(w), run<4>(w), run<5>(w), run<6>(w), w == 16 ? "" : ", ");
  }
  printf("}\n");
  return 0;
}
