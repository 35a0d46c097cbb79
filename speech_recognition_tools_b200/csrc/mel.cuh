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

}  // namespace fdlp
