// Kr: --add_reverb, the room-impulse-response augmentation in front of the framing kernel.
//
// Reference: addReverb  src/featgen/features.py:110-115 (called at computeFDLPSpectrogram.py:168-170,
//            computeMelSpectrum.py:143-145, computeMfccFeatures.py:118-120)
//     out  = np.convolve(sig, reverb)              full convolution, T + R - 1 samples, float64
//     xxc  = np.correlate(sig, out, 'valid')       R lags: xxc[i] = sum_m sig[m] out[m + R-1-i]
//     indM = len(xxc) - np.argmax(xxc)             = d* + 1 with d* the best lag (first maximum in i = last in d)
//     out  = out[indM : indM + len(sig)]           (one sample short when indM == R)
// with reverb = rir[:, 1] / 2^15 (computeFDLPSpectrogram.py:76-87).
//
// B200 design.  Both the convolution (T x R multiply-adds per utterance, 2.6 G at 10 s x 1 s) and the R-lag
// cross-correlation are evaluated as uniformly partitioned block transforms in float64: blocks of B samples,
// F = 2B-point complex FFTs held in shared memory, decimation-in-frequency forward (natural in, digit-reversed
// out) and decimation-in-time inverse (digit-reversed in, natural out), so that spectra are only ever combined
// element-wise in the digit-reversed order and no permutation pass exists.
//   rv_fwd_kernel    X_b  = FFT(sig[(b-1)B .. (b+1)B))                        (overlap-save windows)
//   rv_conv_kernel   Y_b  = sum_q X_{b-q} . H_q ; full[bB .. bB+B) = Re IFFT(Y_b)[B .. 2B) / F ;
//                    O_b  = FFT([full block b, 0 .. 0])                       (fused: the block is in shared memory)
//   rv_corr_kernel   C_q  = sum_c conj(O_c) . X_{c-q} ; xxc at lags d = qB + B - e is Re IFFT(C_q)[e] / F, e = 1..B ;
//                    per-(utterance, q) maximum
//   rv_align_kernel  argmax over q, out[n] = full[indM + n]
// float64 throughout: the FDLP path needs the waveform to better than 32 bits (NOTES_r01.md), and for int16 input
// with an int16 / 2^15 impulse response the exact result is representable, so the error is the FFT's (1e-15 of peak).
// Bound: HBM/L2 traffic of the spectra (16 F bytes per block and pass), see DESIGN.md section 4.6.
#pragma once
#include "common.cuh"

namespace fdlp {

constexpr int kRvThreads = 512;

struct RvArgs {
  const void* wav;
  int wav_dtype;
  const int64_t* utt_wav_off;   // [n_utt+1]
  const int16_t* noise;         // optional noise mixing at load (features.py:24-31), as in the framing kernel
  const int64_t* utt_noise_off;
  const double* utt_alpha;
  const int32_t* utt_blk_off;   // [n_utt+1] first block of each utterance; nb_u = ceil((T_u + R - 1) / B)
  const int32_t* blk_utt;       // [n_blk]
  int R, B, F, logF, P;         // P = ceil(R / B) partitions of the impulse response
  const double2* tw;            // split twiddles: w_F^k = tw[k >> 6] * tw[F/64 + (k & 63)]
  const double2* H;             // [P, F] spectra of the impulse-response partitions
  double2* X;                   // [n_blk, F]
  double2* O;                   // [n_blk, F]
  double* full;                 // [n_blk * B]
  double* part_max;             // [n_utt, P]
  int32_t* part_arg;            // [n_utt, P]
  void* out;                    // packed like wav
  int out_dtype;                // FDLP_WAV_FLOAT32 / FDLP_WAV_FLOAT64
  int32_t* ind;                 // [n_utt] indM
  int64_t* out_len;             // [n_utt] samples written (T, or T - 1 when indM == R)
};

__device__ __forceinline__ double2 rv_cmul(double2 a, double2 b) {
  return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ double2 rv_cmulc(double2 a, double2 b) {  // a * conj(b)
  return make_double2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ double2 rv_add(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 rv_sub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }

// w_F^k from the split table staged in shared memory
__device__ __forceinline__ double2 rv_tw(const double2* stw, int na, int k) {
  return rv_cmul(stw[k >> 6], stw[na + (k & 63)]);
}

// In-place forward FFT of z[0..F): natural order in, digit-reversed order out (radix 4, one leading radix-2 pass
// when log2 F is odd).  Ends with a __syncthreads().
__device__ void rv_fft_dif(double2* z, const int F, const int logF, const double2* stw) {
  const int na = F >> 6;
  int L = F;
  if (logF & 1) {
    const int half = F >> 1;
    for (int j = threadIdx.x; j < half; j += blockDim.x) {
      const double2 a = z[j], b = z[j + half];
      z[j] = rv_add(a, b);
      z[j + half] = rv_cmul(rv_sub(a, b), rv_tw(stw, na, j));
    }
    L = half;
    __syncthreads();
  }
  while (L >= 4) {
    const int q = L >> 2, mul = F / L;
    for (int t = threadIdx.x; t < (F >> 2); t += blockDim.x) {
      const int j = t & (q - 1);
      const int base = (t - j) * 4 + j;  // group * L + j
      const double2 a0 = z[base], a1 = z[base + q], a2 = z[base + 2 * q], a3 = z[base + 3 * q];
      const double2 t0 = rv_add(a0, a2), t1 = rv_sub(a0, a2), t2 = rv_add(a1, a3);
      const double2 d = rv_sub(a1, a3);
      const double2 t3 = make_double2(d.y, -d.x);  // -i (a1 - a3)
      z[base] = rv_add(t0, t2);
      if (q == 1) {  // last pass: all twiddles are 1
        z[base + 1] = rv_add(t1, t3);
        z[base + 2] = rv_sub(t0, t2);
        z[base + 3] = rv_sub(t1, t3);
      } else {
        const int e = j * mul;
        z[base + q] = rv_cmul(rv_add(t1, t3), rv_tw(stw, na, e));
        z[base + 2 * q] = rv_cmul(rv_sub(t0, t2), rv_tw(stw, na, 2 * e));
        z[base + 3 * q] = rv_cmul(rv_sub(t1, t3), rv_tw(stw, na, 3 * e));
      }
    }
    L = q;
    __syncthreads();
  }
}

// In-place inverse of rv_fft_dif without the 1/F factor: digit-reversed order in, natural order out.
__device__ void rv_fft_dit_inv(double2* z, const int F, const int logF, const double2* stw) {
  const int na = F >> 6;
  const int Ltop = (logF & 1) ? (F >> 1) : F;
  for (int L = 4; L <= Ltop; L <<= 2) {
    const int q = L >> 2, mul = F / L;
    for (int t = threadIdx.x; t < (F >> 2); t += blockDim.x) {
      const int j = t & (q - 1);
      const int base = (t - j) * 4 + j;
      double2 u0 = z[base], u1 = z[base + q], u2 = z[base + 2 * q], u3 = z[base + 3 * q];
      if (q != 1) {
        const int e = j * mul;
        u1 = rv_cmulc(u1, rv_tw(stw, na, e));
        u2 = rv_cmulc(u2, rv_tw(stw, na, 2 * e));
        u3 = rv_cmulc(u3, rv_tw(stw, na, 3 * e));
      }
      const double2 s0 = rv_add(u0, u2), s1 = rv_sub(u0, u2), s2 = rv_add(u1, u3);
      const double2 d = rv_sub(u1, u3);
      const double2 s3 = make_double2(-d.y, d.x);  // i (u1 - u3)
      z[base] = rv_add(s0, s2);
      z[base + q] = rv_add(s1, s3);
      z[base + 2 * q] = rv_sub(s0, s2);
      z[base + 3 * q] = rv_sub(s1, s3);
    }
    __syncthreads();
  }
  if (logF & 1) {
    const int half = F >> 1;
    for (int j = threadIdx.x; j < half; j += blockDim.x) {
      const double2 y0 = z[j], y1 = rv_cmulc(z[j + half], rv_tw(stw, na, j));
      z[j] = rv_add(y0, y1);
      z[j + half] = rv_sub(y0, y1);
    }
    __syncthreads();
  }
}

__device__ __forceinline__ double2* rv_stage_tables(const RvArgs& a, unsigned char* smem_raw) {
  double2* stw = reinterpret_cast<double2*>(smem_raw + (size_t)a.F * sizeof(double2));
  const int ntw = (a.F >> 6) + 64;
  for (int i = threadIdx.x; i < ntw; i += blockDim.x) stw[i] = a.tw[i];
  return stw;
}

// X_b: spectrum of the overlap-save window sig[(b-1)B .. (b+1)B) of the (noise-mixed) utterance.
template <typename S>
__global__ void __launch_bounds__(kRvThreads) rv_fwd_kernel(const RvArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* z = reinterpret_cast<double2*>(smem_raw);
  const double2* stw = rv_stage_tables(a, smem_raw);
  const int blk = blockIdx.x;
  const int u = a.blk_utt[blk];
  const int b = blk - a.utt_blk_off[u];
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  const S* sig = reinterpret_cast<const S*>(a.wav) + woff;
  const int16_t* nz = a.noise ? a.noise + a.utt_noise_off[u] : nullptr;
  const double alpha = a.noise ? a.utt_alpha[u] : 0.0;
  const int64_t q0 = (int64_t)(b - 1) * a.B;
  for (int i = threadIdx.x; i < a.F; i += blockDim.x) {
    const int64_t q = q0 + i;
    double v = 0.0;
    if (q >= 0 && q < T) {
      v = (double)sig[q];
      if (nz) v += alpha * (double)nz[q];
    }
    z[i] = make_double2(v, 0.0);
  }
  __syncthreads();
  rv_fft_dif(z, a.F, a.logF, stw);
  double2* X = a.X + (size_t)blk * a.F;
  for (int i = threadIdx.x; i < a.F; i += blockDim.x) X[i] = z[i];
}

// spectra of the impulse-response partitions: H_q = FFT([h[qB .. qB+B), 0 .. 0]); rir is a device float64 array
__global__ void __launch_bounds__(kRvThreads) rv_rir_kernel(const RvArgs a, const double* __restrict__ rir, double2* H) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* z = reinterpret_cast<double2*>(smem_raw);
  const double2* stw = rv_stage_tables(a, smem_raw);
  const int q = blockIdx.x;
  for (int i = threadIdx.x; i < a.F; i += blockDim.x) {
    const int j = q * a.B + i;
    z[i] = make_double2((i < a.B && j < a.R) ? rir[j] : 0.0, 0.0);
  }
  __syncthreads();
  rv_fft_dif(z, a.F, a.logF, stw);
  for (int i = threadIdx.x; i < a.F; i += blockDim.x) H[(size_t)q * a.F + i] = z[i];
}

// Block b of the full convolution and its zero-padded spectrum.
__global__ void __launch_bounds__(kRvThreads) rv_conv_kernel(const RvArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* z = reinterpret_cast<double2*>(smem_raw);
  const double2* stw = rv_stage_tables(a, smem_raw);
  const int blk = blockIdx.x;
  const int u = a.blk_utt[blk];
  const int b = blk - a.utt_blk_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - a.utt_wav_off[u];
  const int F = a.F, B = a.B;
  const int nq = min(a.P, b + 1);
  const double2* __restrict__ Xb = a.X + (size_t)blk * F;
  for (int k = threadIdx.x; k < F; k += blockDim.x) {
    double2 acc0 = make_double2(0.0, 0.0), acc1 = acc0;
    int q = 0;
    for (; q + 1 < nq; q += 2) {
      const double2 x0 = __ldg(Xb - (size_t)q * F + k), h0 = __ldg(a.H + (size_t)q * F + k);
      const double2 x1 = __ldg(Xb - (size_t)(q + 1) * F + k), h1 = __ldg(a.H + (size_t)(q + 1) * F + k);
      acc0 = rv_add(acc0, rv_cmul(x0, h0));
      acc1 = rv_add(acc1, rv_cmul(x1, h1));
    }
    if (q < nq) acc0 = rv_add(acc0, rv_cmul(__ldg(Xb - (size_t)q * F + k), __ldg(a.H + (size_t)q * F + k)));
    z[k] = rv_add(acc0, acc1);
  }
  __syncthreads();
  rv_fft_dit_inv(z, F, a.logF, stw);
  const double inv = 1.0 / (double)F;
  const int64_t full_len = T + a.R - 1;
  double* full = a.full + (size_t)blk * B;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {  // a thread touches only z[i] and z[B + i]
    const int64_t n = (int64_t)b * B + i;
    const double v = (n < full_len) ? z[B + i].x * inv : 0.0;
    full[i] = v;
    z[i] = make_double2(v, 0.0);
    z[B + i] = make_double2(0.0, 0.0);
  }
  __syncthreads();
  rv_fft_dif(z, F, a.logF, stw);
  double2* O = a.O + (size_t)blk * F;
  for (int i = threadIdx.x; i < F; i += blockDim.x) O[i] = z[i];
}

// One (utterance, lag partition q): xxc at lags d in [qB, qB + B) and its maximum (ties: the largest d, which is
// the first index of the reference's xxc, i = R - 1 - d).
__global__ void __launch_bounds__(kRvThreads) rv_corr_kernel(const RvArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* z = reinterpret_cast<double2*>(smem_raw);
  const double2* stw = rv_stage_tables(a, smem_raw);
  const int q = blockIdx.x, u = blockIdx.y;
  const int blk0 = a.utt_blk_off[u];
  const int nb = a.utt_blk_off[u + 1] - blk0;
  const int F = a.F, B = a.B;
  const double2* __restrict__ O = a.O + (size_t)blk0 * F;
  const double2* __restrict__ X = a.X + (size_t)blk0 * F;
  for (int k = threadIdx.x; k < F; k += blockDim.x) {
    double2 acc0 = make_double2(0.0, 0.0), acc1 = acc0;
    int c = q;
    for (; c + 1 < nb; c += 2) {
      const double2 o0 = __ldg(O + (size_t)c * F + k), x0 = __ldg(X + (size_t)(c - q) * F + k);
      const double2 o1 = __ldg(O + (size_t)(c + 1) * F + k), x1 = __ldg(X + (size_t)(c + 1 - q) * F + k);
      acc0 = rv_add(acc0, rv_cmulc(x0, o0));
      acc1 = rv_add(acc1, rv_cmulc(x1, o1));
    }
    if (c < nb) acc0 = rv_add(acc0, rv_cmulc(__ldg(X + (size_t)(c - q) * F + k), __ldg(O + (size_t)c * F + k)));
    z[k] = rv_add(acc0, acc1);
  }
  __syncthreads();
  rv_fft_dit_inv(z, F, a.logF, stw);
  double best = -INFINITY;
  int best_d = -1;
  for (int e = 1 + threadIdx.x; e <= B; e += blockDim.x) {
    const int d = q * B + B - e;
    if (d < a.R) {
      const double v = z[e].x;
      if (v > best || (v == best && d > best_d)) { best = v; best_d = d; }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int od = __shfl_xor_sync(0xffffffffu, best_d, o);
    if (od >= 0 && (best_d < 0 || ov > best || (ov == best && od > best_d))) { best = ov; best_d = od; }
  }
  __shared__ double sh_v[kRvThreads / 32];
  __shared__ int sh_d[kRvThreads / 32];
  if ((threadIdx.x & 31) == 0) { sh_v[threadIdx.x >> 5] = best; sh_d[threadIdx.x >> 5] = best_d; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < kRvThreads / 32; ++w) {
      const double ov = sh_v[w];
      const int od = sh_d[w];
      if (od >= 0 && (best_d < 0 || ov > best || (ov == best && od > best_d))) { best = ov; best_d = od; }
    }
    a.part_max[(size_t)u * a.P + q] = best / (double)F;
    a.part_arg[(size_t)u * a.P + q] = best_d;
  }
}

// out[n] = full[indM + n], n < min(T, T + R - 1 - indM); HBM-bound copy
__global__ void __launch_bounds__(256) rv_align_kernel(const RvArgs a) {
  const int u = blockIdx.y;
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  double best = -INFINITY;
  int best_d = -1;
  for (int q = 0; q < a.P; ++q) {
    const double v = a.part_max[(size_t)u * a.P + q];
    const int d = a.part_arg[(size_t)u * a.P + q];
    if (d >= 0 && (best_d < 0 || v > best || (v == best && d > best_d))) { best = v; best_d = d; }
  }
  const int indM = best_d + 1;
  const int64_t full_len = T + a.R - 1;
  const int64_t n_out = min(T, full_len - indM);
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    if (a.ind) a.ind[u] = indM;
    if (a.out_len) a.out_len[u] = n_out;
  }
  const double* __restrict__ full = a.full + (size_t)a.utt_blk_off[u] * a.B + indM;
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < T; n += (int64_t)gridDim.x * blockDim.x) {
    const double v = (n < n_out) ? full[n] : 0.0;
    if (a.out_dtype == FDLP_WAV_FLOAT64) reinterpret_cast<double*>(a.out)[woff + n] = v;
    else reinterpret_cast<float*>(a.out)[woff + n] = (float)v;
  }
}

}  // namespace fdlp
