// K0: SNR-scaled noise mixing coefficient per utterance.
//
// Reference: add_noise_to_wav  src/featgen/features.py:24-31
//     ns    = noise[rand_num : rand_num + len(sig)]
//     E_s   = mean(sig ** 2);  E_n = mean(ns ** 2)        <- computed IN THE INPUT DTYPE
//     alp   = sqrt(E_s / (E_n * 10 ** (snr / 10)))
//     out   = sig + alp * ns                               (float64)
// WAVs are read as int16 and never normalised (computeFDLPSpectrogram.py:156-157), so `sig ** 2` wraps
// modulo 2^16 in the reference.  FDLP_NOISE_INT16_WRAP reproduces that exactly with integer sums (the mean
// of the wrapped squares is an exact integer / T); FDLP_NOISE_FLOAT uses the true float64 energies.
// The mix itself (sig + alp * ns) is applied by the framing kernel while it loads samples (dct.cuh), so the
// mixed waveform is never written to memory.
#pragma once
#include "common.cuh"

namespace fdlp {

struct NoiseArgs {
  const void* wav;
  int wav_dtype;
  const int64_t* utt_wav_off;
  const int16_t* noise;
  const int64_t* utt_noise_off;
  const double* utt_snr_db;
  double* utt_alpha;  // out [n_utt]
  int mode;
};

__device__ __forceinline__ long long wrap_sq16(int v) {
  return (long long)(short)(v * v);  // int16 * int16 -> int16 with wrap-around, as numpy does
}

constexpr int kNoiseThreads = 512;

// sum of squares of T int16 samples at an arbitrarily aligned address: 16-byte loads (8 samples) over the aligned
// body, scalar head and tail.  Integer (wrapped) sums are exact in any order; the float sums are float64.
__device__ __forceinline__ void sum_sq_i16(const int16_t* __restrict__ p, const int64_t T, const bool wrap,
                                           long long& isum, double& fsum) {
  const int64_t mis = (int64_t)((16 - (reinterpret_cast<uintptr_t>(p) & 15)) & 15) >> 1;
  const int64_t head = (mis < T) ? mis : T;
  const int64_t nvec = (T - head) >> 3;
  const uint4* __restrict__ v = reinterpret_cast<const uint4*>(p + head);
  long long i0 = 0, i1 = 0;
  double f0 = 0.0, f1 = 0.0;
  for (int64_t i = threadIdx.x; i < nvec; i += kNoiseThreads) {
    const uint4 q = __ldg(v + i);
    const unsigned w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int lo = (short)(w[c] & 0xffffu), hi = (short)(w[c] >> 16);
      if (wrap) { i0 += wrap_sq16(lo); i1 += wrap_sq16(hi); }
      else { f0 += (double)(lo * lo); f1 += (double)(hi * hi); }
    }
  }
  for (int64_t i = threadIdx.x; i < head; i += kNoiseThreads) {
    const int sv = p[i];
    if (wrap) i0 += wrap_sq16(sv); else f0 += (double)(sv * sv);
  }
  for (int64_t i = head + 8 * nvec + threadIdx.x; i < T; i += kNoiseThreads) {
    const int sv = p[i];
    if (wrap) i1 += wrap_sq16(sv); else f1 += (double)(sv * sv);
  }
  isum += i0 + i1;
  fsum += f0 + f1;
}

// one CTA per utterance; HBM-bound (one read of the utterance and of its noise excerpt)
__global__ void __launch_bounds__(kNoiseThreads) noise_alpha_kernel(const NoiseArgs a) {
  const int u = blockIdx.x;
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  const int16_t* nz = a.noise + a.utt_noise_off[u];
  long long is = 0, in = 0;
  double fs = 0.0, fn = 0.0;
  const bool wrap = (a.mode == FDLP_NOISE_INT16_WRAP);
  sum_sq_i16(nz, T, wrap, in, fn);
  if (a.wav_dtype == FDLP_WAV_INT16) {
    sum_sq_i16(reinterpret_cast<const int16_t*>(a.wav) + woff, T, wrap, is, fs);
  } else {
    const float* s = reinterpret_cast<const float*>(a.wav) + woff;
    double f0 = 0.0, f1 = 0.0;
    int64_t i = threadIdx.x;
    for (; i + kNoiseThreads < T; i += 2 * kNoiseThreads) {
      const float s0 = __ldg(s + i), s1 = __ldg(s + i + kNoiseThreads);
      // float32 signal: numpy squares in float32 (rounded) before the float64 mean
      if (wrap) { f0 += (double)__fmul_rn(s0, s0); f1 += (double)__fmul_rn(s1, s1); }
      else { f0 += (double)s0 * s0; f1 += (double)s1 * s1; }
    }
    if (i < T) {
      const float s0 = __ldg(s + i);
      f0 += wrap ? (double)__fmul_rn(s0, s0) : (double)s0 * s0;
    }
    fs = f0 + f1;
  }
  constexpr int kW = kNoiseThreads / 32;
  __shared__ long long sh_i[2][kW];
  __shared__ double sh_f[2][kW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 16; o > 0; o >>= 1) {
    is += __shfl_xor_sync(0xffffffffu, is, o);
    in += __shfl_xor_sync(0xffffffffu, in, o);
    fs += __shfl_xor_sync(0xffffffffu, fs, o);
    fn += __shfl_xor_sync(0xffffffffu, fn, o);
  }
  if (lane == 0) { sh_i[0][warp] = is; sh_i[1][warp] = in; sh_f[0][warp] = fs; sh_f[1][warp] = fn; }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long tis = 0, tin = 0;
    double tfs = 0, tfn = 0;
    for (int w = 0; w < kW; ++w) { tis += sh_i[0][w]; tin += sh_i[1][w]; tfs += sh_f[0][w]; tfn += sh_f[1][w]; }
    const double Es = ((double)tis + tfs) / (double)T;
    const double En = ((double)tin + tfn) / (double)T;
    a.utt_alpha[u] = sqrt(Es / (En * pow(10.0, a.utt_snr_db[u] / 10.0)));
  }
}

// add_noise_to_wav's last line, features.py:31: sig + alpha * ns, as a waveform of its own (the FDLP path fuses this
// into the framing load; the mel path takes the mixed float32 waveform).  float64 sum rounded once.
struct NoiseMixArgs {
  const void* wav;
  int wav_dtype;
  const int64_t* utt_wav_off;   // [n_utt+1]
  const int16_t* noise;
  const int64_t* utt_noise_off; // [n_utt]
  const double* utt_alpha;      // [n_utt] from noise_alpha_kernel
  float* out;                   // packed like wav
};

__global__ void __launch_bounds__(256) noise_mix_kernel(const NoiseMixArgs a) {
  const int u = blockIdx.y;
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  const int16_t* nz = a.noise + a.utt_noise_off[u];
  const double alpha = a.utt_alpha[u];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < T; i += (int64_t)gridDim.x * blockDim.x) {
    const double s = a.wav_dtype == FDLP_WAV_INT16 ? (double)reinterpret_cast<const int16_t*>(a.wav)[woff + i]
                                                    : (double)reinterpret_cast<const float*>(a.wav)[woff + i];
    a.out[woff + i] = (float)(s + alpha * (double)nz[i]);
  }
}

// --add_noise diff (computeFDLPSpectrogram.py:162-164): signal = scipy.signal.convolve(signal, a, mode='same') with
// the 13-tap kernel a = [1, 2, 3, 2, 0, -2, -5, -2, 0, 2, 3, 2, 1]:  out[q] = sum_t a[t] x[q + 6 - t], x zero outside
// the utterance.  For int16 input the result is an integer below 2^20, so the float32 output is exact; for float32
// input the sum is formed in float64 and rounded once.  HBM-bound: one read and one write of the batch.
struct DiffArgs {
  const void* wav;
  int wav_dtype;
  const int64_t* utt_wav_off;  // [n_utt+1]
  int n_utt;
  float* out;                  // [total samples]
};

__global__ void __launch_bounds__(256) diff_fir_kernel(const DiffArgs a) {
  constexpr int kTaps = 13;
  constexpr int kA[kTaps] = {1, 2, 3, 2, 0, -2, -5, -2, 0, 2, 3, 2, 1};
  const int u = blockIdx.y;
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < T; q += (int64_t)gridDim.x * blockDim.x) {
    double acc = 0.0;
#pragma unroll
    for (int t = 0; t < kTaps; ++t) {
      const int64_t i = q + 6 - t;
      if (kA[t] != 0 && i >= 0 && i < T) {
        const double x = (a.wav_dtype == FDLP_WAV_INT16) ? (double)reinterpret_cast<const int16_t*>(a.wav)[woff + i]
                                                         : (double)reinterpret_cast<const float*>(a.wav)[woff + i];
        acc += (double)kA[t] * x;
      }
    }
    a.out[woff + q] = (float)acc;
  }
}

// Synthetic white utterances generated in place on the device (throughput sets are never stored):
// the same counter-based integer arithmetic as speech_recognition_tools_b200/synth.py white_int16 -- splitmix64 of
// (seed, utterance id, sample index), four 16-bit uniforms summed and centred, times scale_q16 / 65536 -- so the
// stream is bit-identical to the numpy one the parity tests use.
struct SynthArgs {
  int16_t* out;                 // packed batch
  const int64_t* utt_wav_off;   // [n_utt+1]
  const int64_t* utt_id;        // [n_utt] utterance number in the hash
  int n_utt;
  unsigned long long seed;
  int scale_q16;
};

__global__ void __launch_bounds__(256) synth_white_kernel(const SynthArgs a) {
  const int u = blockIdx.y;
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  const unsigned long long key = (a.seed * 0xD1342543DE82EF95ULL) ^ ((unsigned long long)a.utt_id[u] * 0x2545F4914F6CDD1DULL);
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < T; n += (int64_t)gridDim.x * blockDim.x) {
    unsigned long long z = key + (unsigned long long)n + 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z = z ^ (z >> 31);
    const long long s = (long long)((z & 0xFFFF) + ((z >> 16) & 0xFFFF) + ((z >> 32) & 0xFFFF) + (z >> 48)) - 2 * 65535;
    long long v = (s * a.scale_q16 + 32768) >> 16;
    v = v < -32768 ? -32768 : (v > 32767 ? 32767 : v);
    a.out[woff + n] = (int16_t)v;
  }
}

}  // namespace fdlp
