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
// Everything transformed is real, so two blocks share one complex transform: with D = half the blocks of the
// utterance, Z_j = X_j + i X_{j+D} and OO_b = O_b + i O_{b+D}.  sum_q Z_{b-q} H_q inverts to block b in the real and
// block b + D in the imaginary part (H_q is the spectrum of a real sequence); sum_c conj(OO_c) Z_{c-q} inverts to the
// full correlation sum in the real part (conj(O_c) X_{c-q} + conj(O_{c+D}) X_{c+D-q}, both wanted) with the cross terms
// of the packing in the imaginary part, which is dropped.  Nothing is ever unpacked.
// float64 throughout: the FDLP path needs the waveform to better than 32 bits (NOTES_r01.md), and for int16 input
// with an int16 / 2^15 impulse response the exact result is representable, so the error is the FFT's (1e-15 of peak).
// Bound: HBM/L2 traffic of the spectra (16 F bytes per block and pass), see DESIGN.md section 4.6.
#pragma once
#include "common.cuh"

namespace fdlp {

#ifndef FDLP_RV_THREADS
#define FDLP_RV_THREADS 512
#endif
constexpr int kRvThreads = FDLP_RV_THREADS;
#ifndef FDLP_RV_MINB
#define FDLP_RV_MINB 1
#endif
// The transforms exist in two forms: radix-4 passes (60 registers: two 512-thread CTAs per SM) and pairs of radix-4
// passes fused into in-register radix-16 passes (half the trips through shared memory, 128 registers).  Measured on the
// bench batch (tools/ab_reverb.sh): the window kernel is faster with radix 4 at 512 threads (2.12 vs 2.18 / 2.81 ms),
// the convolution kernel with radix 16 at 256 threads, two CTAs per SM (3.89 vs 4.52 ms); at F = 8192 radix 16 with
// 512 threads wins everywhere.  Both produce the same element order (a fused pass IS two radix-4 passes).

struct RvArgs {
  const void* wav;
  int wav_dtype;
  const int64_t* utt_wav_off;   // [n_utt+1]
  const int16_t* noise;         // optional noise mixing at load (features.py:24-31), as in the framing kernel
  const int64_t* utt_noise_off;
  const double* utt_alpha;
  const int32_t* utt_pair_off;  // [n_utt+1] first pair of each utterance; D_u = ceil(ceil((T_u + R - 1) / B) / 2)
  const int32_t* pair_utt;      // [n_pair]
  const int32_t* utt_conv_off;  // [n_utt+1] first convolution CTA of each utterance; ceil(D_u / pairs_per_cta) each
  const int32_t* conv_utt;      // [n_conv]
  int n_pair;
  int R, B, F, logF, P;         // P = ceil(R / B) partitions of the impulse response
  const double2* tw;            // split twiddles: w_F^k = tw[k >> 6] * tw[F/64 + (k & 63)]
  const double2* H;             // [P, F] spectra of the impulse-response partitions
  double2* X;                   // [n_pair + n_utt (P-1), F] packed window spectra Z_j
  double2* O;                   // [n_pair, F] packed spectra of the zero-padded convolution blocks
  double* full;                 // [2 n_pair B] full convolution, utterance u at 2 pair_off[u] B
  double2* C;                   // [n_utt, P, F] correlation spectra
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

// Shared-memory layout of a transform: element i lives at RVI(i) = i + (i >> 6), i.e. rows of 64 elements with a
// pitch of 65.  Passes whose butterflies span several rows run with consecutive lanes on consecutive elements of a
// row; passes inside a row (sub-transform length <= 64) run with consecutive lanes on consecutive ROWS, a stride of
// 65 x 16 bytes that spreads a quarter warp over all 32 banks, and all lanes of a warp share their twiddles.  (Without
// the pitch the last two radix-4 passes were 4-way and 2-way bank conflicted: 7.7 k conflicts per transform.)
#define RVI(i) ((i) + ((i) >> 6))

// butterfly t of a radix-4 pass of span L = 4q -> index of its first element
__device__ __forceinline__ void rv_bfly_index(const int t, const int q, const int logF, int& j, int& base) {
  if (q > 16) {
    j = t & (q - 1);
    base = (t - j) * 4 + j;  // group * L + j
  } else {
    const int lr = logF - 6;  // log2(rows)
    const int rho = t & ((1 << lr) - 1), beta = t >> lr;
    j = beta & (q - 1);
    base = (rho << 6) + (beta - j) * 4 + j;
  }
}

// ---- radix-16 pass = two consecutive radix-4 passes (spans L and L/4) on 16 registers: one trip through shared
// memory instead of two.  Element r = s + 4m of a butterfly sits at base + r q16 (q16 = L/16); stage A is the radix-4
// butterfly over m (constant twiddles w_16^{r1 s}), stage B the one over s; the j-dependent twiddles of both passes
// commute to the end: output (r1, r2), stored at base + (4 r1 + r2) q16, is multiplied by w_L^{(r1 + 4 r2) j}.
__device__ __forceinline__ double2 rv_w16(int k) {  // exp(-2 pi i k / 16), k in {1, 2, 3, 4, 6, 9}
  constexpr double c = 0.92387953251128674, s = 0.38268343236508978, h = 0.70710678118654752;
  switch (k) {
    case 1: return make_double2(c, -s);
    case 2: return make_double2(h, -h);
    case 3: return make_double2(s, -c);
    case 4: return make_double2(0.0, -1.0);
    case 6: return make_double2(-h, -h);
    case 9: return make_double2(-c, s);
    default: return make_double2(1.0, 0.0);
  }
}

__device__ __forceinline__ void rv_bfly4_fwd(double2& a0, double2& a1, double2& a2, double2& a3) {
  const double2 t0 = rv_add(a0, a2), t1 = rv_sub(a0, a2), t2 = rv_add(a1, a3);
  const double2 d = rv_sub(a1, a3);
  const double2 t3 = make_double2(d.y, -d.x);  // -i (a1 - a3)
  a0 = rv_add(t0, t2); a1 = rv_add(t1, t3); a2 = rv_sub(t0, t2); a3 = rv_sub(t1, t3);
}
__device__ __forceinline__ void rv_bfly4_inv(double2& u0, double2& u1, double2& u2, double2& u3) {
  const double2 s0 = rv_add(u0, u2), s1 = rv_sub(u0, u2), s2 = rv_add(u1, u3);
  const double2 d = rv_sub(u1, u3);
  const double2 s3 = make_double2(-d.y, d.x);  // i (u1 - u3)
  u0 = rv_add(s0, s2); u1 = rv_add(s1, s3); u2 = rv_sub(s0, s2); u3 = rv_sub(s1, s3);
}

// butterfly t of a radix-16 pass of span L = 16 q16 -> j and the index of its first element (lanes along a row when
// the butterfly spans rows, across rows when it lies inside one)
__device__ __forceinline__ void rv_bfly16_index(const int t, const int q16, const int logF, int& j, int& base) {
  if (q16 > 4) {
    j = t & (q16 - 1);
    base = (t - j) * 16 + j;
  } else {
    const int lr = logF - 6;
    const int rho = t & ((1 << lr) - 1), beta = t >> lr;
    j = beta & (q16 - 1);
    base = (rho << 6) + (beta - j) * 16 + j;
  }
}

// the 15 output twiddles w^(r1 + 4 r2), w = w_L^j: tw[4 r1 + r2]
__device__ __forceinline__ void rv_tw16(const double2 w1, double2* tw) {
  const double2 a2 = rv_cmul(w1, w1), a3 = rv_cmul(a2, w1);
  const double2 b1 = rv_cmul(a2, a2), b2 = rv_cmul(b1, b1), b3 = rv_cmul(b2, b1);
  tw[0] = make_double2(1.0, 0.0); tw[1] = b1; tw[2] = b2; tw[3] = b3;
  tw[4] = w1; tw[5] = rv_cmul(w1, b1); tw[6] = rv_cmul(w1, b2); tw[7] = rv_cmul(w1, b3);
  tw[8] = a2; tw[9] = rv_cmul(a2, b1); tw[10] = rv_cmul(a2, b2); tw[11] = rv_cmul(a2, b3);
  tw[12] = a3; tw[13] = rv_cmul(a3, b1); tw[14] = rv_cmul(a3, b2); tw[15] = rv_cmul(a3, b3);
}

__device__ __forceinline__ void rv_pass16_fwd(double2* z, const int F, const int logF, const int L, const double2* stw) {
  const int q16 = L >> 4, mul = F / L, na = F >> 6;
  for (int t = threadIdx.x; t < (F >> 4); t += blockDim.x) {
    int j, base;
    rv_bfly16_index(t, q16, logF, j, base);
    double2 v[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) v[r] = z[RVI(base + r * q16)];
#pragma unroll
    for (int sI = 0; sI < 4; ++sI) {  // stage A over m: v[sI + 4m]
      rv_bfly4_fwd(v[sI], v[sI + 4], v[sI + 8], v[sI + 12]);
      if (sI > 0) {
        v[sI + 4] = rv_cmul(v[sI + 4], rv_w16(1 * sI));
        v[sI + 8] = (2 * sI == 4) ? make_double2(v[sI + 8].y, -v[sI + 8].x) : rv_cmul(v[sI + 8], rv_w16(2 * sI));
        v[sI + 12] = rv_cmul(v[sI + 12], rv_w16(3 * sI));
      }
    }
    // after stage A, v[sI + 4 r1] = y[s][r1]; stage B over s for each r1
#pragma unroll
    for (int r1 = 0; r1 < 4; ++r1) rv_bfly4_fwd(v[4 * r1], v[4 * r1 + 1], v[4 * r1 + 2], v[4 * r1 + 3]);
    if (q16 > 1) {
      double2 tw[16];
      rv_tw16(rv_tw(stw, na, j * mul), tw);
#pragma unroll
      for (int r = 1; r < 16; ++r) v[r] = rv_cmul(v[r], tw[r]);
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) z[RVI(base + r * q16)] = v[r];
  }
  __syncthreads();
}

__device__ __forceinline__ void rv_pass16_inv(double2* z, const int F, const int logF, const int L, const double2* stw) {
  const int q16 = L >> 4, mul = F / L, na = F >> 6;
  for (int t = threadIdx.x; t < (F >> 4); t += blockDim.x) {
    int j, base;
    rv_bfly16_index(t, q16, logF, j, base);
    double2 v[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) v[r] = z[RVI(base + r * q16)];
    if (q16 > 1) {
      double2 tw[16];
      rv_tw16(rv_tw(stw, na, j * mul), tw);
#pragma unroll
      for (int r = 1; r < 16; ++r) v[r] = rv_cmulc(v[r], tw[r]);
    }
#pragma unroll
    for (int r1 = 0; r1 < 4; ++r1) rv_bfly4_inv(v[4 * r1], v[4 * r1 + 1], v[4 * r1 + 2], v[4 * r1 + 3]);
#pragma unroll
    for (int sI = 0; sI < 4; ++sI) {
      if (sI > 0) {
        v[sI + 4] = rv_cmulc(v[sI + 4], rv_w16(1 * sI));
        v[sI + 8] = (2 * sI == 4) ? make_double2(-v[sI + 8].y, v[sI + 8].x) : rv_cmulc(v[sI + 8], rv_w16(2 * sI));
        v[sI + 12] = rv_cmulc(v[sI + 12], rv_w16(3 * sI));
      }
      rv_bfly4_inv(v[sI], v[sI + 4], v[sI + 8], v[sI + 12]);
    }
#pragma unroll
    for (int r = 0; r < 16; ++r) z[RVI(base + r * q16)] = v[r];
  }
  __syncthreads();
}

// In-place forward FFT of z[0..F): natural order in, digit-reversed order out (radix 4, one leading radix-2 pass
// when log2 F is odd).  Ends with a __syncthreads().
template <bool kR16>
__device__ void rv_fft_dif(double2* z, const int F, const int logF, const double2* stw) {
  const int na = F >> 6;
  int L = F;
  if (logF & 1) {
    const int half = F >> 1;
    for (int j = threadIdx.x; j < half; j += blockDim.x) {
      const double2 a = z[RVI(j)], b = z[RVI(j + half)];
      z[RVI(j)] = rv_add(a, b);
      z[RVI(j + half)] = rv_cmul(rv_sub(a, b), rv_tw(stw, na, j));
    }
    L = half;
    __syncthreads();
  }
  while (kR16 && L >= 16) {
    rv_pass16_fwd(z, F, logF, L, stw);
    L >>= 4;
  }
  while (L >= 4) {
    const int q = L >> 2, mul = F / L;
    for (int t = threadIdx.x; t < (F >> 2); t += blockDim.x) {
      int j, base;
      rv_bfly_index(t, q, logF, j, base);
      const int i0 = RVI(base), i1 = RVI(base + q), i2 = RVI(base + 2 * q), i3 = RVI(base + 3 * q);
      const double2 a0 = z[i0], a1 = z[i1], a2 = z[i2], a3 = z[i3];
      const double2 t0 = rv_add(a0, a2), t1 = rv_sub(a0, a2), t2 = rv_add(a1, a3);
      const double2 d = rv_sub(a1, a3);
      const double2 t3 = make_double2(d.y, -d.x);  // -i (a1 - a3)
      z[i0] = rv_add(t0, t2);
      if (q == 1) {  // last pass: all twiddles are 1
        z[i1] = rv_add(t1, t3);
        z[i2] = rv_sub(t0, t2);
        z[i3] = rv_sub(t1, t3);
      } else {
        // w^2 and w^3 by multiplication: the same flop count as forming them from the split table, 4 shared-memory
        // loads fewer per butterfly (the transforms are bound by shared-memory bandwidth)
        const double2 w1 = rv_tw(stw, na, j * mul);
        const double2 w2 = rv_cmul(w1, w1);
        z[i1] = rv_cmul(rv_add(t1, t3), w1);
        z[i2] = rv_cmul(rv_sub(t0, t2), w2);
        z[i3] = rv_cmul(rv_sub(t1, t3), rv_cmul(w2, w1));
      }
    }
    L = q;
    __syncthreads();
  }
}

// In-place inverse of rv_fft_dif without the 1/F factor: digit-reversed order in, natural order out.
template <bool kR16>
__device__ void rv_fft_dit_inv(double2* z, const int F, const int logF, const double2* stw) {
  const int na = F >> 6;
  const int Ltop = (logF & 1) ? (F >> 1) : F;
  // the forward transform fuses pairs of radix-4 passes from the top: spans Ltop, Ltop/16, ...; a single radix-4 pass of
  // span 4 is left when Ltop is an odd power of 4
  int L0 = 4;
  if (kR16) {
    int Lf = Ltop;
    while (Lf >= 16) Lf >>= 4;  // Lf = 4 (a radix-4 pass of span 4 remains) or 1
    L0 = (Lf == 4) ? 8 : 1;     // 8: run the span-4 pass below, then the fused ones; 1: fused passes only
  }
  for (int L = 4; L <= (kR16 ? (L0 == 8 ? 4 : 0) : Ltop); L <<= 2) {
    const int q = L >> 2, mul = F / L;
    for (int t = threadIdx.x; t < (F >> 2); t += blockDim.x) {
      int j, base;
      rv_bfly_index(t, q, logF, j, base);
      const int i0 = RVI(base), i1 = RVI(base + q), i2 = RVI(base + 2 * q), i3 = RVI(base + 3 * q);
      double2 u0 = z[i0], u1 = z[i1], u2 = z[i2], u3 = z[i3];
      if (q != 1) {
        const double2 w1 = rv_tw(stw, na, j * mul);
        const double2 w2 = rv_cmul(w1, w1);
        u1 = rv_cmulc(u1, w1);
        u2 = rv_cmulc(u2, w2);
        u3 = rv_cmulc(u3, rv_cmul(w2, w1));
      }
      const double2 s0 = rv_add(u0, u2), s1 = rv_sub(u0, u2), s2 = rv_add(u1, u3);
      const double2 d = rv_sub(u1, u3);
      const double2 s3 = make_double2(-d.y, d.x);  // i (u1 - u3)
      z[i0] = rv_add(s0, s2);
      z[i1] = rv_add(s1, s3);
      z[i2] = rv_sub(s0, s2);
      z[i3] = rv_sub(s1, s3);
    }
    __syncthreads();
  }
  if (kR16) {
    for (int L = (L0 == 8) ? 64 : 16; L <= Ltop; L <<= 4) rv_pass16_inv(z, F, logF, L, stw);
  }
  if (logF & 1) {
    const int half = F >> 1;
    for (int j = threadIdx.x; j < half; j += blockDim.x) {
      const double2 y0 = z[RVI(j)], y1 = rv_cmulc(z[RVI(j + half)], rv_tw(stw, na, j));
      z[RVI(j)] = rv_add(y0, y1);
      z[RVI(j + half)] = rv_sub(y0, y1);
    }
    __syncthreads();
  }
}

__device__ __forceinline__ double2* rv_stage_tables(const RvArgs& a, unsigned char* smem_raw) {
  double2* stw = reinterpret_cast<double2*>(smem_raw + (size_t)(a.F + (a.F >> 6)) * sizeof(double2));
  const int ntw = (a.F >> 6) + 64;
  for (int i = threadIdx.x; i < ntw; i += blockDim.x) stw[i] = a.tw[i];
  return stw;
}

// Z_j = X_j + i X_{j+D}: spectrum of two overlap-save windows of the (noise-mixed) utterance packed into one complex
// transform, X_j = FFT(sig[(j-1)B .. (j+1)B)), D = pairs of the utterance (half its blocks, rounded up),
// j = -(P-1) .. D-1.  The impulse-response partitions and the blocks of the convolution are real sequences, so every
// later step is linear over the two packed halves and nothing is ever unpacked (header comment).
// CTAs [0, n_pair) take j >= 0; the remaining n_utt (P-1) CTAs the P-1 windows before the utterance, whose real half
// is zero.  Row of Z_j of utterance u: pair_off[u] + u (P-1) + j + P-1.
template <typename S>
__global__ void __launch_bounds__(kRvThreads, FDLP_RV_MINB) rv_fwd_kernel(const RvArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* z = reinterpret_cast<double2*>(smem_raw);
  const double2* stw = rv_stage_tables(a, smem_raw);
  int u, j;
  if ((int)blockIdx.x < a.n_pair) {
    u = a.pair_utt[blockIdx.x];
    j = blockIdx.x - a.utt_pair_off[u];
  } else {
    const int r = blockIdx.x - a.n_pair;
    u = r / (a.P - 1);
    j = -1 - (r % (a.P - 1));
  }
  const int D = a.utt_pair_off[u + 1] - a.utt_pair_off[u];
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - woff;
  const S* sig = reinterpret_cast<const S*>(a.wav) + woff;
  const int16_t* nz = a.noise ? a.noise + a.utt_noise_off[u] : nullptr;
  const double alpha = a.noise ? a.utt_alpha[u] : 0.0;
  const int64_t q0 = (int64_t)(j - 1) * a.B, q1 = (int64_t)(j + D - 1) * a.B;
  for (int i = threadIdx.x; i < a.F; i += blockDim.x) {
    const int64_t qa = q0 + i, qb = q1 + i;
    double va = 0.0, vb = 0.0;
    if (qa >= 0 && qa < T) {
      va = (double)sig[qa];
      if (nz) va += alpha * (double)nz[qa];
    }
    if (qb >= 0 && qb < T) {
      vb = (double)sig[qb];
      if (nz) vb += alpha * (double)nz[qb];
    }
    z[RVI(i)] = make_double2(va, vb);
  }
  __syncthreads();
  rv_fft_dif<false>(z, a.F, a.logF, stw);
  double2* Z = a.X + ((size_t)a.utt_pair_off[u] + (size_t)u * (a.P - 1) + (size_t)(j + a.P - 1)) * a.F;
  for (int i = threadIdx.x; i < a.F; i += blockDim.x) Z[i] = z[RVI(i)];
}

// spectra of the impulse-response partitions: H_q = FFT([h[qB .. qB+B), 0 .. 0]); rir is a device float64 array
__global__ void __launch_bounds__(kRvThreads, FDLP_RV_MINB) rv_rir_kernel(const RvArgs a, const double* __restrict__ rir, double2* H) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* z = reinterpret_cast<double2*>(smem_raw);
  const double2* stw = rv_stage_tables(a, smem_raw);
  const int q = blockIdx.x;
  for (int i = threadIdx.x; i < a.F; i += blockDim.x) {
    const int j = q * a.B + i;
    z[RVI(i)] = make_double2((i < a.B && j < a.R) ? rir[j] : 0.0, 0.0);
  }
  __syncthreads();
  rv_fft_dif<false>(z, a.F, a.logF, stw);
  for (int i = threadIdx.x; i < a.F; i += blockDim.x) H[(size_t)q * a.F + i] = z[RVI(i)];
}

// Blocks b and b + D of the full convolution (real and imaginary half of one transform) and the packed spectrum of
// the zero-padded pair.  A CTA takes kM (1 or 2) consecutive pairs: the P + kM - 1 spectra Z they share are read
// once per bin, the accumulators of the later pairs wait in registers while the first goes through the transforms.
constexpr int kRvMaxBins = 8192 / kRvThreads;  // F <= 8192

template <bool kR16>
__device__ __forceinline__ void rv_conv_finish(const RvArgs& a, double2* z, const double2* stw, const int u, const int b,
                                               const int D, const int64_t full_len) {
  const int F = a.F, B = a.B;
  rv_fft_dit_inv<kR16>(z, F, a.logF, stw);
  const double inv = 1.0 / (double)F;
  double* full = a.full + (size_t)2 * a.utt_pair_off[u] * B;
  for (int i = threadIdx.x; i < B; i += blockDim.x) {  // a thread touches only z[i] and z[B + i]
    const int64_t n0 = (int64_t)b * B + i, n1 = (int64_t)(b + D) * B + i;
    const double2 y = z[RVI(B + i)];
    const double v0 = (n0 < full_len) ? y.x * inv : 0.0;
    const double v1 = (n1 < full_len) ? y.y * inv : 0.0;
    full[n0] = v0;
    full[n1] = v1;
    z[RVI(i)] = make_double2(v0, v1);
    z[RVI(B + i)] = make_double2(0.0, 0.0);
  }
  __syncthreads();
  rv_fft_dif<kR16>(z, F, a.logF, stw);
  double2* O = a.O + ((size_t)a.utt_pair_off[u] + b) * F;
  for (int i = threadIdx.x; i < F; i += blockDim.x) O[i] = z[RVI(i)];
}

template <int kBins, int kM>
__global__ void __launch_bounds__(kRvThreads, FDLP_RV_MINB) rv_conv_kernel(const RvArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* z = reinterpret_cast<double2*>(smem_raw);
  const double2* stw = rv_stage_tables(a, smem_raw);
  const int u = a.conv_utt[blockIdx.x];
  const int b0 = (blockIdx.x - a.utt_conv_off[u]) * kM;
  const int D = a.utt_pair_off[u + 1] - a.utt_pair_off[u];
  const int64_t T = a.utt_wav_off[u + 1] - a.utt_wav_off[u];
  const int64_t full_len = T + a.R - 1;
  const int F = a.F, P = a.P;
  const bool two = kM == 2 && b0 + 1 < D;
  // row of Z_{b0 - q}, q = -1 .. P-1 (q = -1 is the newest window, used by pair b0 + 1 only)
  const double2* __restrict__ Zb = a.X + ((size_t)a.utt_pair_off[u] + (size_t)u * (P - 1) + (size_t)(b0 + P - 1)) * F;
  double2 held[kM == 2 ? kBins : 1];
  if constexpr (kM == 1) {  // any thread count (launched with 256 threads at F <= 4096)
    for (int k = threadIdx.x; k < F; k += blockDim.x) {
      double2 acc0 = make_double2(0.0, 0.0), acc1 = acc0;
      int q = 0;
      for (; q + 1 < P; q += 2) {
        const double2 h0 = __ldg(a.H + (size_t)q * F + k), z0 = __ldg(Zb - (size_t)q * F + k);
        const double2 h1 = __ldg(a.H + (size_t)(q + 1) * F + k), z1 = __ldg(Zb - (size_t)(q + 1) * F + k);
        acc0 = rv_add(acc0, rv_cmul(z0, h0));
        acc1 = rv_add(acc1, rv_cmul(z1, h1));
      }
      if (q < P) acc0 = rv_add(acc0, rv_cmul(__ldg(Zb - (size_t)q * F + k), __ldg(a.H + (size_t)q * F + k)));
      z[RVI(k)] = rv_add(acc0, acc1);
    }
  } else {
#pragma unroll
    for (int i = 0; i < kBins; ++i) {
      const int k = threadIdx.x + i * kRvThreads;
      if (k < F) {
        double2 acc0 = make_double2(0.0, 0.0), acc1 = acc0;
        double2 znew = two ? __ldg(Zb + (size_t)F + k) : make_double2(0.0, 0.0);
        for (int q = 0; q < P; ++q) {
          const double2 h = __ldg(a.H + (size_t)q * F + k);
          const double2 zc = __ldg(Zb - (size_t)q * F + k);
          acc1 = rv_add(acc1, rv_cmul(znew, h));  // pair b0 + 1: Z_{b0 + 1 - q}
          acc0 = rv_add(acc0, rv_cmul(zc, h));    // pair b0:     Z_{b0 - q}
          znew = zc;
        }
        held[i] = acc1;
        z[RVI(k)] = acc0;
      }
    }
  }
  __syncthreads();
  rv_conv_finish<kM == 1>(a, z, stw, u, b0, D, full_len);
  if constexpr (kM == 2) {
    if (two) {
      __syncthreads();
#pragma unroll
      for (int i = 0; i < kBins; ++i) {
        const int k = threadIdx.x + i * kRvThreads;
        if (k < F) z[RVI(k)] = held[i];
      }
      __syncthreads();
      rv_conv_finish<false>(a, z, stw, u, b0 + 1, D, full_len);
    }
  }
}

// Correlation spectra C_q = sum_c conj(OO_c) Z_{c-q} for kQ consecutive lag partitions at once: a thread owns one
// bin and walks the pairs of the utterance, keeping the last kQ window spectra of its bin in a register ring, so every
// row of OO and Z is read once per group of kQ partitions instead of once per partition.  The inverse transform has the
// wanted correlation (both halves of every pair) in the real part and the cross terms of the packing in the imaginary
// part.  grid = (ceil(F / threads), ceil(P / kQ), n_utt); L2/HBM-bound streaming of 2 D rows per utterance and group.
constexpr int kRvQ = 8;

__global__ void __launch_bounds__(kRvThreads, FDLP_RV_MINB) rv_corr_mac_kernel(const RvArgs a) {
  const int k = blockIdx.x * kRvThreads + threadIdx.x;
  const int q0 = blockIdx.y * kRvQ, u = blockIdx.z;
  const int F = a.F, P = a.P;
  if (k >= F) return;
  const int p0 = a.utt_pair_off[u];
  const int D = a.utt_pair_off[u + 1] - p0;
  const double2* __restrict__ O = a.O + (size_t)p0 * F + k;
  const double2* __restrict__ Z = a.X + ((size_t)p0 + (size_t)u * (P - 1) + (size_t)(P - 1)) * F + k;  // Z_0
  double2 acc[kRvQ], ring[kRvQ];
#pragma unroll
  for (int i = 0; i < kRvQ; ++i) {
    acc[i] = make_double2(0.0, 0.0);
    // ring[i] = Z_{c - q0 - i} at step c; before the first step the entries i >= 1 hold Z_{-1 - q0 - (i-1)}
    ring[i] = (i >= 1 && q0 + i <= P - 1) ? __ldg(Z - (ptrdiff_t)(q0 + i) * F) : make_double2(0.0, 0.0);
  }
#pragma unroll 2
  for (int c = 0; c < D; ++c) {
    const double2 o = __ldg(O + (size_t)c * F);
    ring[0] = __ldg(Z + (ptrdiff_t)(c - q0) * F);
#pragma unroll
    for (int i = 0; i < kRvQ; ++i) acc[i] = rv_add(acc[i], rv_cmulc(ring[i], o));
#pragma unroll
    for (int i = kRvQ - 1; i >= 1; --i) ring[i] = ring[i - 1];
  }
#pragma unroll
  for (int i = 0; i < kRvQ; ++i)
    if (q0 + i < P) a.C[((size_t)u * P + q0 + i) * F + k] = acc[i];
}

// One (utterance, lag partition q): xxc at lags d in [qB, qB + B) from C_q and its maximum (ties: the largest d, which
// is the first index of the reference's xxc, i = R - 1 - d).
__global__ void __launch_bounds__(kRvThreads, FDLP_RV_MINB) rv_corr_kernel(const RvArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* z = reinterpret_cast<double2*>(smem_raw);
  const double2* stw = rv_stage_tables(a, smem_raw);
  const int q = blockIdx.x, u = blockIdx.y;
  const int F = a.F, B = a.B;
  const double2* __restrict__ C = a.C + ((size_t)u * a.P + q) * F;
  for (int k = threadIdx.x; k < F; k += blockDim.x) z[RVI(k)] = C[k];
  __syncthreads();
  rv_fft_dit_inv<false>(z, F, a.logF, stw);
  double best = -INFINITY;
  int best_d = -1;
  for (int e = 1 + threadIdx.x; e <= B; e += blockDim.x) {
    const int d = q * B + B - e;
    if (d < a.R) {
      const double v = z[RVI(e)].x;
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
  const double* __restrict__ full = a.full + (size_t)2 * a.utt_pair_off[u] * a.B + indM;
  for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < T; n += (int64_t)gridDim.x * blockDim.x) {
    const double v = (n < n_out) ? full[n] : 0.0;
    if (a.out_dtype == FDLP_WAV_FLOAT64) reinterpret_cast<double*>(a.out)[woff + n] = v;
    else reinterpret_cast<float*>(a.out)[woff + n] = (float)v;
  }
}

}  // namespace fdlp
