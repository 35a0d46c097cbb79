// Km, second generation (nfft = 2048: the recipes' mel baseline, e2e/*/run_melspec.sh): registers + tensor cores.
//
// Reference: computeMelSpectrum.py:147-152 -- Hamming frames, abs(fft(frames, nfft))[:nfft/2+1] @ fbank.T, log10.
//
// ncu of the first warp-per-frame kernel (profiles/ncu_full_r01_v12_mel_summary.json): shared-memory pipe 79 % busy,
// ~910 wavefronts per frame, 560 of them in the filterbank dot products (8-byte-strided magnitude reads at unrelated
// offsets: 3.5 wavefronts per load), the rest in three trips of the spectrum through shared memory.  This kernel keeps
// the 4-step FFT of mel.cuh and removes everything else from shared memory:
//  * after the second FFT step lane k1 holds Z[k1 + 32 k2] in registers.  The real-transform partner Z[n - k] of
//    every bin sits in lane (32 - k1) & 31 at register 31 - k2 (lane 0: its own register (32 - k2) & 31): ONE shuffle
//    per bin instead of a store / two loads; the unpack factor exp(-2 pi i k / nfft) is W^k1 (a lane constant) times
//    exp(-2 pi i k2 / 64) (a compile-time constant); the step-1 twiddles W_n^(k1 lane) come from four lane constants by
//    short recurrences.  The 1025 magnitudes of a frame end up in registers, bin k1 + 32 k2 in lane k1.
//  * the filterbank is a banded GEMM on the tensor cores: a CTA of 8 warps transforms 16 frames per round (2 per
//    warp), parks the magnitudes as A[16 frames][bins] (fp32, row pitch = 4 mod 32 banks: conflict-free fragment
//    loads), and multiplies by the filterbank tile by tile -- 8 bands per tile, only the k-steps the tile's bands
//    touch (163 of 1290 at the recipe's 80 mel bands) -- with mma.sync.m16n8k8 TF32 in the 3xTF32 scheme
//    (a_lo b_hi + a_hi b_lo + a_hi b_hi, fp32 accumulate: 2^-21 relative, the error of an fp32 dot product).
//    tcgen05 is not used: the whole filterbank is 30 MMAs per frame, 3 % of the kernel's instructions.
// Shared-memory wavefronts per frame: 128 (transpose) + 33 (magnitudes) + ~45 (A fragments) = ~206, against ~910.
#pragma once
#include "mel.cuh"

namespace fdlp {

constexpr int kMmaWarps = 8;
constexpr int kMmaFrames = 16;            // frames per round = rows of the MMA
constexpr int kMmaPitch = 1060;           // floats per A row: 1025 bins + zero pad to the last k-step, = 4 mod 32
constexpr int kMmaMaxTiles = 16;          // J <= 128
constexpr int kMmaHalfTile = 16 * 33;     // float2 per warp = a 32 x 33 float tile: real and imaginary parts transposed in turn

struct MelMmaArgs {
  const void* wav;
  const int64_t* utt_wav_off;
  const int64_t* utt_frame_off;
  const int32_t* frame_utt;  // [n_frames]
  int64_t n_frames;
  int flen, shift, nfft;
  const float* ham;          // [nfft] window, zero beyond min(flen, nfft)
  const float2* lane_tw;     // [5][32]: W_n^(lane), W_n^(8 lane), W_n^(16 lane), W_n^(24 lane) (n = 1024), W_2048^(lane)
  const float4* bfrag;       // [k-steps of all tiles][32 lanes]: {b_hi[k = t], b_hi[k = t+4], b_lo[t], b_lo[t+4]}, band n0 + lane/4
  float* feats;              // [n_frames, J]
  int J, power;
  // filterbank tiles of 8 bands, grouped by warp (longest first): tiles warp_tile0[w] .. warp_tile0[w+1]
  int32_t warp_tile0[kMmaWarps + 1];
  int32_t tile_n0[kMmaMaxTiles];   // first band
  int32_t tile_k0[kMmaMaxTiles];   // first bin (multiple of 8)
  int32_t tile_nk[kMmaMaxTiles];   // k-steps of 8 bins
  int32_t tile_b0[kMmaMaxTiles];   // first k-step in bfrag
};

__host__ __device__ constexpr size_t mel_mma_smem_bytes() {
  return (size_t)kMmaFrames * kMmaPitch * 4 + (size_t)kMmaWarps * kMmaHalfTile * 8;
}

__device__ __forceinline__ float2 cmulf(const float2 a, const float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t b0, const uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// exp(-2 pi i k2 / 64), k2 < 32, as (cos, -sin): compile-time after unrolling
__device__ __forceinline__ float2 w64(int k2) {
  // W_64^k2 = W_32^(k2/2) for even k2; odd k2 from the 64-point table below
  constexpr float c64[32] = {1.0000000000f, 0.9951847267f, 0.9807852804f, 0.9569403357f, 0.9238795325f, 0.8819212643f,
                             0.8314696123f, 0.7730104534f, 0.7071067812f, 0.6343932842f, 0.5555702330f, 0.4713967368f,
                             0.3826834324f, 0.2902846773f, 0.1950903220f, 0.0980171403f, 0.0000000000f, -0.0980171403f,
                             -0.1950903220f, -0.2902846773f, -0.3826834324f, -0.4713967368f, -0.5555702330f, -0.6343932842f,
                             -0.7071067812f, -0.7730104534f, -0.8314696123f, -0.8819212643f, -0.9238795325f, -0.9569403357f,
                             -0.9807852804f, -0.9951847267f};
  // sin(2 pi k2 / 64) = cos(2 pi (k2 - 16) / 64)
  const float c = c64[k2];
  const float s = k2 >= 16 ? c64[k2 - 16] : -c64[k2 + 16];
  return make_float2(c, -s);
}

// One frame of one warp: samples -> 1025 magnitudes; lane k1 produces bins k1 + 32 k2 (lane 0 also bin 1024) and stores
// them as it goes into the frame's row of A (arow = row + lane: consecutive lanes, consecutive floats).
template <int M1, typename S>
__device__ __forceinline__ void mel_mma_frame(const MelMmaArgs& a, long long fr, int lane, float2* tile, float* arow) {
  constexpr int R2 = 32, n = 1024, TS = 33;
  constexpr int NV = M1 > 0 ? M1 : R2;
  const int ext = (a.flen % 2 == 0) ? a.flen / 2 - 1 : (a.flen - 1) / 2;
  const int used = a.flen < a.nfft ? a.flen : a.nfft;  // fft(x, nfft) truncates longer frames
  const int m1_hi = (used + 63) / 64;                   // rows m1 with a non-zero sample
  const int u = a.frame_utt[fr];
  const int kf = (int)(fr - a.utt_frame_off[u]);
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  const S* sig = reinterpret_cast<const S*>(a.wav) + woff;
  const int64_t q0 = (int64_t)kf * a.shift - ext;
  const int rows = M1 > 0 ? M1 : m1_hi;
  const bool interior = q0 >= 0 && q0 + 64 * rows <= T;
  const bool aligned = (reinterpret_cast<uintptr_t>(sig + q0) & 3) == 0;
  float2 v[NV];
#pragma unroll
  for (int m1 = 0; m1 < NV; ++m1) {
    float2 z = make_float2(0.f, 0.f);
    if (M1 > 0 || m1 < m1_hi) {
      const int i0 = 2 * (32 * m1 + lane);
      float2 x;
      if (interior) x = load_pair<S>(sig + q0 + i0, aligned);
      else x = make_float2((float)sig[reflect_index(q0 + i0, T)], (float)sig[reflect_index(q0 + i0 + 1, T)]);
      // window and lane twiddles are re-read per frame (L1 hits): kept in registers for the kernel's lifetime they
      // cost 20 registers through the phases that hold 32 complex values, i.e. spills
      const float2 h = __ldg(reinterpret_cast<const float2*>(a.ham + i0));
      z = make_float2(x.x * h.x, x.y * h.y);
    }
    v[m1] = z;
  }
  // step 1: R2-point DFT over m1 (lane = m2), then the twiddle W_n^(k1 m2) by recurrences of at most 7 products
  float2 y[R2];
  if constexpr (M1 > 0) dft_pruned<R2, M1>(v, y);
  else {
#pragma unroll
    for (int k1 = 0; k1 < R2; ++k1) y[k1] = v[k1];
    fft_reg<R2>(y);
  }
  const float2 w1 = __ldg(a.lane_tw + lane);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    float2 t = q == 0 ? make_float2(1.f, 0.f) : __ldg(a.lane_tw + 32 * q + lane);
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      if (q > 0 || b > 0) y[8 * q + b] = cmulf(y[8 * q + b], t);
      if (b < 7) t = cmulf(t, w1);
    }
  }
  // step 2 input: transpose through the warp's tile, real parts then imaginary parts (a 32 x 33 float tile: half the
  // shared memory of a float2 tile, and at most 32 + 32 of the 128 y / c registers are live at any point)
  float2 c[32];
  float* tf = reinterpret_cast<float*>(tile);
#pragma unroll
  for (int k1 = 0; k1 < 32; ++k1) tf[k1 * TS + lane] = y[k1].x;
  __syncwarp();
#pragma unroll
  for (int m2 = 0; m2 < 32; ++m2) c[m2].x = tf[lane * TS + m2];
  __syncwarp();
#pragma unroll
  for (int k1 = 0; k1 < 32; ++k1) tf[k1 * TS + lane] = y[k1].y;
  __syncwarp();
#pragma unroll
  for (int m2 = 0; m2 < 32; ++m2) c[m2].y = tf[lane * TS + m2];
  __syncwarp();
  fft_reg<32>(c);  // c[k2] = Z[k1 + 32 k2], k1 = lane
  // real-transform unpack, one shuffle per bin: X[k] = E - i A D, E = (Z[k] + conj Z[n-k]) / 2, D = (Z[k] - conj Z[n-k]) / 2
  const int src = (32 - lane) & 31;
  const float2 wk1 = __ldg(a.lane_tw + 128 + lane);
#pragma unroll
  for (int k2 = 0; k2 < 32; ++k2) {
    float2 zc;
    zc.x = __shfl_sync(0xffffffffu, c[31 - k2].x, src);
    zc.y = __shfl_sync(0xffffffffu, c[31 - k2].y, src);
    if (lane == 0) zc = c[(32 - k2) & 31];
    const float2 zk = c[k2];
    const float2 A = cmulf(wk1, w64(k2));
    // (the factors 1/2 of E and D are folded into the filterbank weights: A holds 2 |X|)
    const float ex = zk.x + zc.x, ey = zk.y - zc.y;
    const float dx = zk.x - zc.x, dy = zk.y + zc.y;
    const float adx = A.x * dx - A.y * dy, ady = A.x * dy + A.y * dx;
    const float px = ex + ady, py = ey - adx;
    arow[32 * k2] = sqrt_approx(px * px + py * py);
  }
  if (lane == 0) arow[1024] = 2.f * fabsf(c[0].x - c[0].y);  // bin n: X[n] = Re Z[0] - Im Z[0]
  (void)n;
}

template <int M1, typename S>
__global__ void __launch_bounds__(kMmaWarps * 32, 2) mel_frames_mma_kernel(const MelMmaArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* sA = reinterpret_cast<float*>(smem_raw);                                   // [16][kMmaPitch]
  float2* tiles = reinterpret_cast<float2*>(sA + kMmaFrames * kMmaPitch);           // [warps][16 * 33]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float2* tile = tiles + (size_t)warp * kMmaHalfTile;
  // pad columns (bins 1025 .. pitch) are read by the last k-steps with zero weights: they must hold finite values
  for (int i = threadIdx.x; i < kMmaFrames * (kMmaPitch - 1025); i += blockDim.x)
    sA[(i / (kMmaPitch - 1025)) * kMmaPitch + 1025 + i % (kMmaPitch - 1025)] = 0.f;
  const int g = lane >> 2, t = lane & 3;
  const long long n_rounds = (a.n_frames + kMmaFrames - 1) / kMmaFrames;
  __syncthreads();
  for (long long rd = blockIdx.x; rd < n_rounds; rd += gridDim.x) {
    const long long f0 = rd * kMmaFrames;
    // ---- phase 1: two frames per warp -> magnitudes into A ----
#pragma unroll 1
    for (int i = 0; i < 2; ++i) {
      const int row = 2 * warp + i;
      const long long fr = f0 + row;
      if (fr < a.n_frames)
        mel_mma_frame<M1, S>(a, fr, lane, tile, sA + row * kMmaPitch + lane);
    }
    __syncthreads();
    // ---- phase 2: banded GEMM A[16 x bins] x fbank^T, this warp's tiles of 8 bands; 3xTF32 ----
    for (int ti = a.warp_tile0[warp]; ti < a.warp_tile0[warp + 1]; ++ti) {
      const int n0 = a.tile_n0[ti], k0 = a.tile_k0[ti], nk = a.tile_nk[ti];
      const float4* bp = a.bfrag + (size_t)a.tile_b0[ti] * 32 + lane;
      const float* ap = sA + g * kMmaPitch + k0 + t;
      float d[4] = {0.f, 0.f, 0.f, 0.f};
      float4 bn = __ldg(bp);
      for (int s = 0; s < nk; ++s) {
        const float4 b = bn;
        if (s + 1 < nk) bn = __ldg(bp + (size_t)(s + 1) * 32);
        const float a0 = ap[8 * s], a1 = ap[8 * kMmaPitch + 8 * s], a2 = ap[8 * s + 4], a3 = ap[8 * kMmaPitch + 8 * s + 4];
        uint32_t ah[4], al[4];
        ah[0] = to_tf32(a0); ah[1] = to_tf32(a1); ah[2] = to_tf32(a2); ah[3] = to_tf32(a3);
        al[0] = to_tf32(a0 - __uint_as_float(ah[0])); al[1] = to_tf32(a1 - __uint_as_float(ah[1]));
        al[2] = to_tf32(a2 - __uint_as_float(ah[2])); al[3] = to_tf32(a3 - __uint_as_float(ah[3]));
        mma_tf32(d, al, __float_as_uint(b.x), __float_as_uint(b.y));
        mma_tf32(d, ah, __float_as_uint(b.z), __float_as_uint(b.w));
        mma_tf32(d, ah, __float_as_uint(b.x), __float_as_uint(b.y));
      }
      // D[row g / g+8][band n0 + 2t, + 1]
#pragma unroll
      for (int hrow = 0; hrow < 2; ++hrow) {
        const long long fr = f0 + g + 8 * hrow;
        if (fr < a.n_frames) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int j = n0 + 2 * t + e;
            if (j < a.J) {
              const float s_ = d[2 * hrow + e];
              a.feats[fr * a.J + j] = a.power ? s_ * s_ : log10f(s_);
            }
          }
        }
      }
    }
    __syncthreads();  // A is rewritten by the next round
  }
}

}  // namespace fdlp
