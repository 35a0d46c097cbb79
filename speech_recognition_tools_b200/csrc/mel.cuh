// Km: mel-spectrogram baseline path.
//
// Reference: compute_mel_spectrum  src/featgen/computeMelSpectrum.py:147-157
//     frames = getFrames(signal, 16000, frate, fduration, np.hamming)            (features.py:118-154)
//     out    = log10( abs(fft(frames, nfft))[:, :nfft/2+1] @ fbank.T )           (magnitude, not power)
//            | ( ... ) ** 2 for --spectrum_type power
//
// One CTA per 100 Hz frame: the 20 ms Hamming frame (zero-padded / truncated to nfft) is packed as an
// nfft/2-point complex sequence in shared memory, transformed with the same in-place radix-4/2 DIF passes
// as the DCT kernel (FP32: the stated tolerance is 1e-4 in log10), unpacked to |X[k]|, and contracted with
// the filterbank one warp per band over the band's non-zero support (mel triangles hold 1932 non-zeros of
// 80 x 1025 at the recipe setting, so a dense GEMM would be 97 % zeros).
#pragma once
#include "common.cuh"
#include "dct.cuh"

namespace fdlp {

struct MelArgs {
  const void* wav;
  const int64_t* utt_wav_off;
  const int64_t* utt_frame_off;
  const int32_t* frame_utt;  // [n_frames] owning utterance (built on device by mel_frame_map_kernel)
  int flen, shift, nfft, n;  // n = nfft/2
  int npass;
  int radix[kMaxRadixPasses];
  const double* ham;       // [flen]
  const double2* tw;       // [n]
  const int32_t* perm;     // [n]
  const double2* unpackA;  // [n+1] exp(-2 pi i k / nfft)
  const float* fbank;      // [J, n+1]
  const int32_t* band_lo;  // [J]
  const int32_t* band_hi;  // [J]
  float* feats;            // [n_frames, J]
  double* cmvn;            // [2, J+1] or null
  int J, power;
};

__global__ void mel_frame_map_kernel(const int64_t* utt_frame_off, int n_utt, int32_t* frame_utt) {
  const int u = blockIdx.x;
  if (u >= n_utt) return;
  const int64_t f0 = utt_frame_off[u], f1 = utt_frame_off[u + 1];
  for (int64_t f = f0 + threadIdx.x; f < f1; f += blockDim.x) frame_utt[f] = u;
}

template <typename S>
__global__ void __launch_bounds__(128) mel_frames_kernel(const MelArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* sv = reinterpret_cast<float*>(smem_raw);            // nfft reals == n complex
  cplx<float>* z = reinterpret_cast<cplx<float>*>(smem_raw);
  float* mag = sv + a.nfft;                                   // [n+1]

  const int64_t fr = blockIdx.x;
  const int u = a.frame_utt[fr];
  const int k = (int)(fr - a.utt_frame_off[u]);
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  const S* sig = reinterpret_cast<const S*>(a.wav) + woff;
  const int ext = (a.flen % 2 == 0) ? a.flen / 2 - 1 : (a.flen - 1) / 2;
  const int64_t q0 = (int64_t)k * a.shift - ext;
  const int used = a.flen < a.nfft ? a.flen : a.nfft;  // fft(x, nfft) truncates longer frames

  for (int i = threadIdx.x; i < a.nfft; i += blockDim.x) {
    float v = 0.f;
    if (i < used) {
      const int64_t q = reflect_index(q0 + i, T);
      v = (float)((double)sig[q] * __ldg(a.ham + i));
    }
    sv[i] = v;  // z[m] = x[2m] + j x[2m+1]
  }
  __syncthreads();

  int L = a.n;
  for (int p = 0; p < a.npass; ++p) {
    const int R = a.radix[p];
    if (R == 4) fft_pass<float, 4>(z, a.n, L, a.tw);
    else if (R == 2) fft_pass<float, 2>(z, a.n, L, a.tw);
    else if (R == 3) fft_pass<float, 3>(z, a.n, L, a.tw);
    else fft_pass<float, 5>(z, a.n, L, a.tw);
    L /= R;
  }

  const int n = a.n;
  for (int kq = threadIdx.x; kq <= n; kq += blockDim.x) {
    const int k1 = (kq == n) ? 0 : kq, k2 = (kq == 0) ? 0 : n - kq;
    const cplx<float> zk = z[a.perm[k1]];
    cplx<float> zc = z[a.perm[k2]];
    zc.y = -zc.y;
    const cplx<float> E = {0.5f * (zk.x + zc.x), 0.5f * (zk.y + zc.y)};
    const cplx<float> D = {0.5f * (zk.x - zc.x), 0.5f * (zk.y - zc.y)};
    const double2 Ad = __ldg(a.unpackA + kq);
    const cplx<float> AD = cmul(cplx<float>{(float)Ad.x, (float)Ad.y}, D);
    const float vx = E.x + AD.y, vy = E.y - AD.x;
    mag[kq] = sqrtf(vx * vx + vy * vy);
  }
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int j = warp; j < a.J; j += nwarp) {
    const int lo = a.band_lo[j], hi = a.band_hi[j];
    const float* w = a.fbank + (size_t)j * (n + 1);
    float s = 0.f;
    for (int b = lo + lane; b < hi; b += 32) s = fmaf(mag[b], __ldg(w + b), s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) {
      const float o = a.power ? s * s : log10f(s);
      a.feats[fr * a.J + j] = o;
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// Km, warp-per-frame variant (nfft/2 = 32 * R2 complex points, R2 in {16, 32}): the CTA-per-frame kernel above
// spends its time in barriers between radix passes.  Here one warp owns a frame and the nfft/2-point complex
// FFT is a 4-step transform held in registers: lane m2 holds z[32 m1 + m2] (m1 < R2), does an R2-point FFT over
// m1 in registers, multiplies by W_n^(m2 k1), the warp transposes through its private shared-memory tile, and
// lane k1 does a 32-point FFT over m2: X[k1 + R2 k2].  Butterfly twiddles are compile-time constants.
__device__ constexpr float kW32c[16] = {1.000000000f, 0.980785280f, 0.923879533f, 0.831469612f, 0.707106781f, 0.555570233f, 0.382683432f, 0.195090322f, 0.000000000f, -0.195090322f, -0.382683432f, -0.555570233f, -0.707106781f, -0.831469612f, -0.923879533f, -0.980785280f};
__device__ constexpr float kW32s[16] = {0.000000000f, 0.195090322f, 0.382683432f, 0.555570233f, 0.707106781f, 0.831469612f, 0.923879533f, 0.980785280f, 1.000000000f, 0.980785280f, 0.923879533f, 0.831469612f, 0.707106781f, 0.555570233f, 0.382683432f, 0.195090322f};

__host__ __device__ constexpr int bitrev_c(int i, int bits) {
  int r = 0;
  for (int b = 0; b < bits; ++b) r |= ((i >> b) & 1) << (bits - 1 - b);
  return r;
}
__host__ __device__ constexpr int ilog2_c(int n) { return n <= 1 ? 0 : 1 + ilog2_c(n >> 1); }

// w^i for w = exp(-2 pi j / 32) as (c, s) with w^i = c - j s; i is a compile-time constant after unrolling
__device__ __forceinline__ float w32c(int i) { i &= 31; return i < 16 ? kW32c[i] : -kW32c[i - 16]; }
__device__ __forceinline__ float w32s(int i) { i &= 31; return i < 16 ? kW32s[i] : -kW32s[i - 16]; }

// acc += z * (c - j s); the comparisons fold at compile time, so the trivial roots cost two adds
__device__ __forceinline__ void cmac_w(float2& acc, const float2 z, const float c, const float s) {
  if (s == 0.f && c == 1.f) { acc.x += z.x; acc.y += z.y; }
  else if (s == 0.f && c == -1.f) { acc.x -= z.x; acc.y -= z.y; }
  else if (c == 0.f && s == 1.f) { acc.x += z.y; acc.y -= z.x; }
  else if (c == 0.f && s == -1.f) { acc.x -= z.y; acc.y += z.x; }
  else {
    acc.x = fmaf(z.y, s, fmaf(z.x, c, acc.x));
    acc.y = fmaf(-z.x, s, fmaf(z.y, c, acc.y));
  }
}

// forward DFT of N (16 or 32) points, natural order in and out, fully unrolled (all indices compile-time)
template <int N, int LEN>
__device__ __forceinline__ void fft_reg_stage(float2 (&a)[N]) {
#pragma unroll
  for (int i = 0; i < N; i += LEN) {
#pragma unroll
    for (int k = 0; k < LEN / 2; ++k) {
      const float c = kW32c[(32 / LEN) * k], s = kW32s[(32 / LEN) * k];  // w = c - j s
      const float2 v = a[i + k + LEN / 2];
      const float2 u = a[i + k];
      float2 p = u, m = u;
      cmac_w(p, v, c, s);
      cmac_w(m, v, -c, -s);
      a[i + k] = p;
      a[i + k + LEN / 2] = m;
    }
  }
  if constexpr (LEN < N) fft_reg_stage<N, LEN * 2>(a);
}

template <int N>
__device__ __forceinline__ void fft_reg(float2 (&a)[N]) {
  constexpr int BITS = ilog2_c(N);
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const int j = bitrev_c(i, BITS);
    if (i < j) { const float2 t = a[i]; a[i] = a[j]; a[j] = t; }
  }
  fft_reg_stage<N, 2>(a);
}

// N-point DFT of a sequence whose rows m >= M are zero (short frames in a long FFT): direct sums split into the
// even and odd rows, y[k] = E + O, y[k + N/2] = E - O
template <int N, int M>
__device__ __forceinline__ void dft_pruned(const float2 (&v)[M], float2 (&y)[N]) {
#pragma unroll
  for (int k = 0; k < N / 2; ++k) {
    float2 E = v[0], O = make_float2(0.f, 0.f);
#pragma unroll
    for (int m = 1; m < M; ++m) {
      const int idx = (32 / N) * ((m * k) % N);
      if (m & 1) {
        if (m == 1) {
          const float c = w32c(idx), s = w32s(idx);
          if (s == 0.f) O = v[1];
          else O = make_float2(fmaf(v[1].y, s, v[1].x * c), fmaf(-v[1].x, s, v[1].y * c));
        } else cmac_w(O, v[m], w32c(idx), w32s(idx));
      } else cmac_w(E, v[m], w32c(idx), w32s(idx));
    }
    y[k] = make_float2(E.x + O.x, E.y + O.y);
    y[k + N / 2] = make_float2(E.x - O.x, E.y - O.y);
  }
}

struct MelWarpArgs {
  const void* wav;
  const int64_t* utt_wav_off;
  const int64_t* utt_frame_off;
  const int32_t* frame_utt;  // [n_frames]
  int64_t n_frames;
  int flen, shift, nfft, n;  // n = nfft/2 = 32 * R2
  const float* ham;          // [nfft] window, zero beyond min(flen, nfft)
  const float2* tw4;         // [R2][32]  W_n^(k1 * lane)
  const float2* unpackA;     // [n+1] exp(-2 pi i k / nfft)
  const float* wts;          // filter weights: band j, bin b at wts[band[j].z + b - band[j].x]
  const int4* band;          // [J] {first bin, bin count, offset into wts, 0}
  int wts_smem;              // > 0: number of weights to stage in shared memory; 0: read wts from global
  float* feats;              // [n_frames, J]
  double* cmvn;              // [2, J+1] or null
  int J, power;
};

constexpr int kMelWarps = 8;
constexpr int kMelMaxBandsPerLane = 4;  // J <= 128

__host__ __device__ constexpr size_t mel_warp_smem_bytes(int R2, int J, int wts_smem) {
  // per warp tile [R2][33] float2 | twiddles [R2][32] float2 | unpack [n/2+1] float2 (+pad) | band int4[J] |
  // band sums [warps][J] float | staged weights
  return (size_t)kMelWarps * R2 * 33 * 8 + (size_t)R2 * 32 * 8 + (size_t)(16 * R2 + 2) * 8 + (size_t)J * 16 +
         (size_t)kMelWarps * J * 4 + (size_t)wts_smem * 4;
}

template <typename S>
__device__ __forceinline__ float2 load_pair(const S* p, bool aligned);
template <>
__device__ __forceinline__ float2 load_pair<int16_t>(const int16_t* p, bool aligned) {
  if (aligned) {
    const uint32_t u = __ldg(reinterpret_cast<const uint32_t*>(p));
    return make_float2((float)(int16_t)(u & 0xffffu), (float)(int16_t)(u >> 16));
  }
  return make_float2((float)__ldg(p), (float)__ldg(p + 1));
}
template <>
__device__ __forceinline__ float2 load_pair<float>(const float* p, bool) {
  return make_float2(__ldg(p), __ldg(p + 1));
}

__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// |X[k]| and |X[n-k]| of the real transform from Z[k], Z[n-k] of the packed complex transform
__device__ __forceinline__ void unpack_mag(const float2 zk, const float2 zc, const float2 A, float& pm, float& mm) {
  const float ex = 0.5f * (zk.x + zc.x), ey = 0.5f * (zk.y - zc.y);
  const float dx = 0.5f * (zk.x - zc.x), dy = 0.5f * (zk.y + zc.y);
  const float adx = A.x * dx - A.y * dy, ady = A.x * dy + A.y * dx;
  const float px = ex + ady, py = ey - adx, mx = ex - ady, my = ey + adx;
  pm = sqrt_approx(px * px + py * py);
  mm = sqrt_approx(mx * mx + my * my);
}

#ifndef FDLP_MEL_LPB
#define FDLP_MEL_LPB 2
#endif
// lanes per band in the filterbank dot products.  Measured on the recipe (80 bands, 1025 bins, 852 k frames):
// 1 lane 4.17 ms, 2 lanes 3.72, 4 lanes 3.98, 8 lanes 4.59, 16 lanes 6.19, 32 lanes 8.62: few lanes per band mean
// few rounds and few shuffles, and two lanes still halve the longest band's serial chain
constexpr int kMelLpb = FDLP_MEL_LPB;
constexpr int kMelLpbLog = kMelLpb == 1 ? 0 : kMelLpb == 2 ? 1 : kMelLpb == 4 ? 2 : kMelLpb == 8 ? 3 : kMelLpb == 16 ? 4 : 5;

// (A dense in-place magnitude array -- bins <= n/2 at float m, bins > n/2 at float n + m, so that a warp's reads spread
// over all 32 banks instead of the 16 even ones -- was measured slower: 3.83 vs 3.72 ms; the index select costs more
// than the conflicts it removes.  Four accumulators instead of two: 4.05 ms.)
__device__ __forceinline__ float band_dot(const float* magf, const float* w, int lo, int cnt, int q) {
  float s0 = 0.f, s1 = 0.f;
  int t = q;
  for (; t + kMelLpb < cnt; t += 2 * kMelLpb) {
    s0 = fmaf(magf[2 * (lo + t)], w[t], s0);
    s1 = fmaf(magf[2 * (lo + t + kMelLpb)], w[t + kMelLpb], s1);
  }
  if (t < cnt) s0 = fmaf(magf[2 * (lo + t)], w[t], s0);
  return s0 + s1;
}

// M1 > 0: only the first M1 packed rows (64 M1 samples) of a frame are non-zero and the first FFT step is the
// pruned DFT; M1 == 0: any frame length, full R2-point FFT.
template <int R2, int M1, typename S>
__global__ void __launch_bounds__(kMelWarps * 32, 2) mel_frames_warp_kernel(const MelWarpArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int n = 32 * R2;
  constexpr int TS = 33;  // transpose row stride (float2)
  constexpr int NV = M1 > 0 ? M1 : R2;
  float2* tiles = reinterpret_cast<float2*>(smem_raw);
  float2* s_tw = tiles + kMelWarps * R2 * TS;
  float2* s_A = s_tw + R2 * 32;
  int4* s_band = reinterpret_cast<int4*>(s_A + 16 * R2 + 2);
  float* s_sum = reinterpret_cast<float*>(s_band + a.J) + warp * a.J;
  float* s_wts = reinterpret_cast<float*>(s_band + a.J) + kMelWarps * a.J;
  for (int i = threadIdx.x; i < R2 * 32; i += blockDim.x) s_tw[i] = a.tw4[i];
  for (int i = threadIdx.x; i <= n / 2; i += blockDim.x) s_A[i] = a.unpackA[i];
  for (int i = threadIdx.x; i < a.J; i += blockDim.x) s_band[i] = a.band[i];
  for (int i = threadIdx.x; i < a.wts_smem; i += blockDim.x) s_wts[i] = a.wts[i];
  __syncthreads();
  // per warp: transpose tile [R2][33] float2, reused for Z[n] (natural order) once every lane holds its row in
  // registers; the magnitudes then overwrite the real parts in place (|X[k]| at zl[k].x, k <= n)
  float2* tile = tiles + (size_t)warp * R2 * TS;
  float2* zl = tile;
  const float* magf = reinterpret_cast<const float*>(tile);
  const int ext = (a.flen % 2 == 0) ? a.flen / 2 - 1 : (a.flen - 1) / 2;
  const int used = a.flen < a.nfft ? a.flen : a.nfft;  // fft(x, nfft) truncates longer frames
  const int m1_hi = (used + 63) / 64;                   // rows m1 with a non-zero sample (m = 32 m1 + lane < used/2)
  float2 hw[NV];
  if constexpr (M1 > 0) {
#pragma unroll
    for (int m1 = 0; m1 < M1; ++m1) hw[m1] = make_float2(__ldg(a.ham + 2 * (32 * m1 + lane)), __ldg(a.ham + 2 * (32 * m1 + lane) + 1));
  }
  const int g = lane >> kMelLpbLog, q = lane & (kMelLpb - 1);

  const long long wstride = (long long)gridDim.x * kMelWarps;
  for (long long fr = (long long)blockIdx.x * kMelWarps + warp; fr < a.n_frames; fr += wstride) {
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
        float2 h;
        if constexpr (M1 > 0) h = hw[m1];
        else h = make_float2(__ldg(a.ham + i0), __ldg(a.ham + i0 + 1));
        z = make_float2(x.x * h.x, x.y * h.y);
      }
      v[m1] = z;
    }
    float2 y[R2];
    if constexpr (M1 > 0) dft_pruned<R2, M1>(v, y);
    else {
#pragma unroll
      for (int k1 = 0; k1 < R2; ++k1) y[k1] = v[k1];
      fft_reg<R2>(y);
    }
#pragma unroll
    for (int k1 = 0; k1 < R2; ++k1) {
      const float2 w = s_tw[k1 * 32 + lane];
      tile[k1 * TS + lane] = make_float2(y[k1].x * w.x - y[k1].y * w.y, y[k1].x * w.y + y[k1].y * w.x);
    }
    __syncwarp();
    if (lane < R2) {
      float2 c[32];
#pragma unroll
      for (int m2 = 0; m2 < 32; ++m2) c[m2] = tile[lane * TS + m2];
      fft_reg<32>(c);
      __syncwarp((R2 == 32) ? 0xffffffffu : 0x0000ffffu);  // every row is in registers: the tile may be overwritten
#pragma unroll
      for (int k2 = 0; k2 < 32; ++k2) zl[lane + R2 * k2] = c[k2];
    }
    __syncwarp();
    // unpack the packed real transform in place, bins k and n-k together:
    //   X[k] = E - j A D,  X[n-k] = conj(E + j A D),  E = (Z[k] + conj Z[n-k]) / 2,  D = (Z[k] - conj Z[n-k]) / 2
    for (int t0 = 0; t0 < n / 64; t0 += 4) {
      float2 zk[4], zc[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = lane + 32 * (t0 + i);
        zk[i] = zl[k];
        zc[i] = zl[(n - k) & (n - 1)];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k = lane + 32 * (t0 + i);
        float pm, mm;
        unpack_mag(zk[i], zc[i], s_A[k], pm, mm);
        zl[k].x = pm;
        zl[n - k].x = mm;
      }
    }
    if (lane == 0) {
      float pm, mm;
      unpack_mag(zl[n / 2], zl[n / 2], s_A[n / 2], pm, mm);
      zl[n / 2].x = pm;
    }
    __syncwarp();
    // filter bank: kMelLpb lanes per band, 32 / kMelLpb bands per round; the band sums go through shared memory so that
    // the log, the store and the CMVN sums run one band per lane
    for (int j0 = 0; j0 < a.J; j0 += 32 / kMelLpb) {
      const int j = j0 + g;
      float s = 0.f;
      if (j < a.J) {
        const int4 bi = s_band[j];
        if (a.wts_smem > 0) s = band_dot(magf, s_wts + bi.z, bi.x, bi.y, q);
        else s = band_dot(magf, a.wts + bi.z, bi.x, bi.y, q);
      }
#pragma unroll
      for (int o = 1; o < kMelLpb; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (q == 0 && j < a.J) s_sum[j] = s;
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < kMelMaxBandsPerLane; ++i) {
      const int j = lane + 32 * i;
      if (j < a.J) {
        const float s = s_sum[j];
        const float o = a.power ? s * s : log10f(s);
        a.feats[fr * a.J + j] = o;
      }
    }
    __syncwarp();
  }
}

}  // namespace fdlp
