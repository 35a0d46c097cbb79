// FP64 pipe peak microbenchmark for B200 (sm_100a).
// MEASURED_PEAKS.json carries only HBM and bf16 numbers; the FDLP path is bound by the FP64
// pipe, so the denominators for its roofline are measured here:
//   dfma      : register-resident DFMA chains (vector FP64 pipe)
//   dmma884   : mma.sync.m8n8k4.f64 (FP64 tensor path), register-resident
//   dmma1688  : mma.sync.m16n8k8.f64
//   dmma16816 : mma.sync.m16n8k16.f64
// Prints one JSON line.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int ITERS = 4096;

__global__ void __launch_bounds__(256) k_dfma(double* out, double a, double b) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void dmma16816(double* c, const double* a, const double* b) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                 "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma884(double* out, double a, double b) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma1688(double* out, double a, double b) {
  double c[NACC][4];
  double av[4] = {a, a + 1, a + 2, a + 3}, bv[2] = {b, b + 1};
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma1688(c[i], av, bv);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_dmma16816(double* out, double a, double b) {
  double c[NACC][4];
  double av[8], bv[4];
#pragma unroll
  for (int i = 0; i < 8; ++i) av[i] = a + i;
#pragma unroll
  for (int i = 0; i < 4; ++i) bv[i] = b + i;
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma16816(c[i], av, bv);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FP32 FMA peak for context
__global__ void __launch_bounds__(256) k_ffma(float* out, float a, float b) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double time_ms(F launch, int reps) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; ++i) launch();
  CK(cudaDeviceSynchronize());
  double best = 1e30;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(e0));
    launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  return best;
}

// sustained: back-to-back launches for ~secs seconds, returns average ms per launch
template <typename F>
static double sustained_ms(F launch, double secs, double est_ms) {
  int n = (int)(secs * 1000.0 / est_ms) + 1;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  CK(cudaEventRecord(e0));
  for (int i = 0; i < n; ++i) launch();
  CK(cudaEventRecord(e1));
  CK(cudaEventSynchronize(e1));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  return ms / n;
}

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  int sms = prop.multiProcessorCount;
  int blocks = sms * 8, threads = 256;
  double* out; CK(cudaMalloc(&out, sizeof(double) * blocks * threads));
  const double total_threads = (double)blocks * threads;
  const double warps = total_threads / 32;

  double t_dfma = time_ms([&] { k_dfma<<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 10);
  double f_dfma = total_threads * ITERS * 16 * 2 / (t_dfma * 1e-3) / 1e12;
  double t_dfma_s = sustained_ms([&] { k_dfma<<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 3.0, t_dfma);
  double f_dfma_s = total_threads * ITERS * 16 * 2 / (t_dfma_s * 1e-3) / 1e12;

  double t_884 = time_ms([&] { k_dmma884<8><<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 10);
  double f_884 = warps * ITERS * 8 * (8 * 8 * 4 * 2.0) / (t_884 * 1e-3) / 1e12;
  double t_884_s = sustained_ms([&] { k_dmma884<8><<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 3.0, t_884);
  double f_884_s = warps * ITERS * 8 * (8 * 8 * 4 * 2.0) / (t_884_s * 1e-3) / 1e12;
  double t_884b = time_ms([&] { k_dmma884<20><<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 10);
  double f_884b = warps * ITERS * 20 * (8 * 8 * 4 * 2.0) / (t_884b * 1e-3) / 1e12;

  double t_1688 = time_ms([&] { k_dmma1688<8><<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 10);
  double f_1688 = warps * ITERS * 8 * (16 * 8 * 8 * 2.0) / (t_1688 * 1e-3) / 1e12;
  double t_16816 = time_ms([&] { k_dmma16816<8><<<blocks, threads>>>(out, 1.0000001, 1e-9); }, 10);
  double f_16816 = warps * ITERS * 8 * (16 * 8 * 16 * 2.0) / (t_16816 * 1e-3) / 1e12;

  float* outf = (float*)out;
  double t_ffma = time_ms([&] { k_ffma<<<blocks, threads>>>(outf, 1.0000001f, 1e-9f); }, 10);
  double f_ffma = total_threads * ITERS * 16 * 2 / (t_ffma * 1e-3) / 1e12;

  printf("{\"gpu\": \"%s\", \"sms\": %d, \"dfma_tflops\": %.2f, \"dfma_tflops_sustained\": %.2f, "
         "\"dmma884_tflops\": %.2f, \"dmma884_tflops_sustained\": %.2f, \"dmma884_20acc_tflops\": %.2f, "
         "\"dmma1688_tflops\": %.2f, \"dmma16816_tflops\": %.2f, \"ffma_tflops\": %.2f}\n",
         prop.name, sms, f_dfma, f_dfma_s, f_884, f_884_s, f_884b, f_1688, f_16816, f_ffma);
  return 0;
}
