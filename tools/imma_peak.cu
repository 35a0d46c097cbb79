// Legacy warp-level MMA peak microbenchmark for B200 (sm_100a): register-resident mma.sync loops.
//   imma16832 : mma.sync.m16n8k32.s32.s8.s8.s32   (int8 -> int32, exact)
//   hmma16816 : mma.sync.m16n8k16.f32.bf16.bf16.f32
// Used to decide whether an exact int8-slice (Ozaki) autocorrelation on the legacy tensor path is worth it.
// Prints one JSON line.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o imma_peak imma_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int ITERS = 4096;

__device__ __forceinline__ void imma16832(int* c, const uint32_t* a, const uint32_t* b) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void hmma16816(float* c, const uint32_t* a, const uint32_t* b) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int NACC>
__global__ void __launch_bounds__(256) k_imma(int* out, uint32_t a0, uint32_t b0) {
  int c[NACC][4];
  uint32_t a[4] = {a0, a0 + 1, a0 + 2, a0 + 3}, b[2] = {b0, b0 + 1};
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) imma16832(c[i], a, b);
  }
  int s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NACC>
__global__ void __launch_bounds__(256) k_hmma(float* out, uint32_t a0, uint32_t b0) {
  float c[NACC][4];
  uint32_t a[4] = {a0, a0 + 1, a0 + 2, a0 + 3}, b[2] = {b0, b0 + 1};
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) hmma16816(c[i], a, b);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
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

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  int sms = prop.multiProcessorCount;
  int* out; CK(cudaMalloc(&out, sizeof(int) * sms * 8 * 256));
  printf("{\"gpu\": \"%s\", \"sms\": %d", prop.name, sms);
  for (int wpb : {4, 8}) {
    for (int bps : {1, 2, 4}) {
      int blocks = sms * bps, threads = wpb * 32;
      if (blocks * threads > sms * 8 * 256) continue;
      const double warps = (double)blocks * wpb;
      double t = time_ms([&] { k_imma<16><<<blocks, threads>>>(out, 0x01020304u, 0x04030201u); }, 5);
      double tops = warps * ITERS * 16 * (16 * 8 * 32 * 2.0) / (t * 1e-3) / 1e12;
      printf(", \"imma16832_w%d_b%d_tops\": %.1f", wpb, bps, tops);
      double th = time_ms([&] { k_hmma<16><<<blocks, threads>>>((float*)out, 0x3f803f80u, 0x3f803f80u); }, 5);
      double tf = warps * ITERS * 16 * (16 * 8 * 16 * 2.0) / (th * 1e-3) / 1e12;
      printf(", \"hmma16816_w%d_b%d_tflops\": %.1f", wpb, bps, tf);
    }
  }
  printf("}\n");
  return 0;
}
