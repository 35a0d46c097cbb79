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
      if (a.cmvn) {
        atomicAdd(a.cmvn + j, (double)o);
        atomicAdd(a.cmvn + (a.J + 1) + j, (double)o * (double)o);
      }
    }
  }
  if (a.cmvn && threadIdx.x == 0) atomicAdd(a.cmvn + a.J, 1.0);
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

// forward DFT of N (16 or 32) points, natural order in and out, fully unrolled (all indices compile-time)
template <int N, int LEN>
__device__ __forceinline__ void fft_reg_stage(float2 (&a)[N]) {
#pragma unroll
  for (int i = 0; i < N; i += LEN) {
#pragma unroll
    for (int k = 0; k < LEN / 2; ++k) {
      const float c = kW32c[(32 / LEN) * k], s = kW32s[(32 / LEN) * k];  // w = c - j s
      const float2 v = a[i + k + LEN / 2];
      const float2 t = make_float2(v.x * c + v.y * s, v.y * c - v.x * s);
      const float2 u = a[i + k];
      a[i + k] = make_float2(u.x + t.x, u.y + t.y);
      a[i + k + LEN / 2] = make_float2(u.x - t.x, u.y - t.y);
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

struct MelWarpArgs {
  const void* wav;
  const int64_t* utt_wav_off;
  const int64_t* utt_frame_off;
  const int32_t* frame_utt;  // [n_frames]
  int64_t n_frames;
  int flen, shift, nfft, n;  // n = nfft/2 = 32 * R2
  const float* ham;          // [flen]
  const float2* tw4;         // [R2][32]  W_n^(k1 * lane)
  const float2* unpackA;     // [n+1] exp(-2 pi i k / nfft)
  const float* fbank;        // [J, n+1]
  const int32_t* band_lo;
  const int32_t* band_hi;
  float* feats;              // [n_frames, J]
  double* cmvn;              // [2, J+1] or null
  int J, power;
};

constexpr int kMelWarps = 8;
constexpr int kMelMaxBandsPerLane = 4;  // J <= 128

template <int R2, typename S>
__global__ void __launch_bounds__(kMelWarps * 32, 2) mel_frames_warp_kernel(const MelWarpArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int n = 32 * R2;
  constexpr int TS = 33;  // transpose row stride (float2)
  // per warp: transpose tile [R2][33] float2, reused for Z[n] (natural order) once every lane holds its row in
  // registers, followed by mag[n+1] floats
  constexpr int kWarpF2 = R2 * TS + (n + 4) / 2;
  float2* tile = reinterpret_cast<float2*>(smem_raw) + (size_t)warp * kWarpF2;
  float2* zl = tile;
  float* mag = reinterpret_cast<float*>(tile + R2 * TS);
  const int ext = (a.flen % 2 == 0) ? a.flen / 2 - 1 : (a.flen - 1) / 2;
  const int used = a.flen < a.nfft ? a.flen : a.nfft;  // fft(x, nfft) truncates longer frames
  const int m1_hi = (used + 63) / 64;                   // rows m1 with a non-zero sample (m = 32 m1 + lane < used/2)
  double cs1[kMelMaxBandsPerLane], cs2[kMelMaxBandsPerLane];
#pragma unroll
  for (int i = 0; i < kMelMaxBandsPerLane; ++i) cs1[i] = cs2[i] = 0.0;
  long long nfr_done = 0;

  const long long wstride = (long long)gridDim.x * kMelWarps;
  for (long long fr = (long long)blockIdx.x * kMelWarps + warp; fr < a.n_frames; fr += wstride) {
    const int u = a.frame_utt[fr];
    const int kf = (int)(fr - a.utt_frame_off[u]);
    const int64_t woff = a.utt_wav_off[u];
    const int64_t T = a.utt_wav_off[u + 1] - woff;
    const S* sig = reinterpret_cast<const S*>(a.wav) + woff;
    const int64_t q0 = (int64_t)kf * a.shift - ext;
    float2 v[R2];
#pragma unroll
    for (int m1 = 0; m1 < R2; ++m1) {
      float2 z = make_float2(0.f, 0.f);
      if (m1 < m1_hi) {
        const int i0 = 2 * (32 * m1 + lane);
        if (i0 < used) z.x = (float)sig[reflect_index(q0 + i0, T)] * __ldg(a.ham + i0);
        if (i0 + 1 < used) z.y = (float)sig[reflect_index(q0 + i0 + 1, T)] * __ldg(a.ham + i0 + 1);
      }
      v[m1] = z;
    }
    fft_reg<R2>(v);
#pragma unroll
    for (int k1 = 0; k1 < R2; ++k1) {
      const float2 w = __ldg(a.tw4 + k1 * 32 + lane);
      const float2 y = make_float2(v[k1].x * w.x - v[k1].y * w.y, v[k1].x * w.y + v[k1].y * w.x);
      tile[k1 * TS + lane] = y;
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
    // unpack the packed real transform: X[k] = E - j A D,  E = (Z[k] + conj Z[n-k]) / 2,  D = (Z[k] - conj Z[n-k]) / 2
    for (int kq = lane; kq <= n; kq += 32) {
      const float2 zk = zl[(kq == n) ? 0 : kq];
      float2 zc = zl[(kq == 0) ? 0 : n - kq];
      zc.y = -zc.y;
      const float ex = 0.5f * (zk.x + zc.x), ey = 0.5f * (zk.y + zc.y);
      const float dx = 0.5f * (zk.x - zc.x), dy = 0.5f * (zk.y - zc.y);
      const float2 A = __ldg(a.unpackA + kq);
      const float adx = A.x * dx - A.y * dy, ady = A.x * dy + A.y * dx;
      const float vx = ex + ady, vy = ey - adx;
      mag[kq] = sqrtf(vx * vx + vy * vy);
    }
    __syncwarp();
#pragma unroll
    for (int i = 0; i < kMelMaxBandsPerLane; ++i) {
      const int j = lane + 32 * i;
      if (j < a.J) {
        const int lo = a.band_lo[j], hi = a.band_hi[j];
        const float* w = a.fbank + (size_t)j * (n + 1);
        float s0 = 0.f, s1 = 0.f;
        int b = lo;
        for (; b + 1 < hi; b += 2) { s0 = fmaf(mag[b], __ldg(w + b), s0); s1 = fmaf(mag[b + 1], __ldg(w + b + 1), s1); }
        if (b < hi) s0 = fmaf(mag[b], __ldg(w + b), s0);
        const float s = s0 + s1;
        const float o = a.power ? s * s : log10f(s);
        a.feats[fr * a.J + j] = o;
        cs1[i] += (double)o;
        cs2[i] += (double)o * (double)o;
      }
    }
    ++nfr_done;
    __syncwarp();
  }
  if (a.cmvn && nfr_done > 0) {
#pragma unroll
    for (int i = 0; i < kMelMaxBandsPerLane; ++i) {
      const int j = lane + 32 * i;
      if (j < a.J) {
        atomicAdd(a.cmvn + j, cs1[i]);
        atomicAdd(a.cmvn + (a.J + 1) + j, cs2[i]);
      }
    }
    if (lane == 0) atomicAdd(a.cmvn + a.J, (double)nfr_done);
  }
}

}  // namespace fdlp
