// K1+K2: segment framing (reflect padding, Hamming) fused with the long DCT-II.
//
// Reference: getFrames (src/featgen/features.py:118-154) and
//            scipy.fftpack.dct(time_frames) / sqrt(2N) (src/featgen/computeFDLPSpectrogram.py:174-178).
//
// One CTA per segment.  The DCT-II of length N is computed with Makhoul's reordering as a real FFT of
// length N, itself done as an N/2-point complex FFT held entirely in shared memory (float64: 8N bytes,
// 187.5 KB at N = 24000; float32 when 8N does not fit).  The FFT is an in-place decimation-in-frequency
// mixed-radix (4/2/3/5) transform; its digit-reversed output order is absorbed by the unpack step through
// a precomputed permutation.  Samples are fetched with 16-byte vector loads aligned on the global
// address (8 int16 or 4 float32 per load); only segments touching an utterance edge take the scalar
// reflect path.
#pragma once
#include "common.cuh"

namespace fdlp {

constexpr int kMaxRadixPasses = 16;

struct DctArgs {
  const void* wav;
  const int64_t* utt_wav_off;
  const int32_t* utt_seg_off;
  const int32_t* seg_utt;
  // optional noise mixing (features.py:24-31): sample = sig + alpha[u] * noise[noise_off[u] + q]
  const int16_t* noise;
  const int64_t* utt_noise_off;
  const double* utt_alpha;
  int N, S, n;  // n = N/2
  int npass;
  int radix[kMaxRadixPasses];
  const double* ham;       // [N] np.hamming(N)
  const double2* tw;       // [n] exp(-2 pi i k / n) (mel CTA kernel)
  // split twiddle tables of the radix passes, staged in shared memory behind the data: for pass p with
  // sub-transform length L_p, w_L^j = A_p[j >> 6] * B_p[j & 63], A_p[a] = exp(-2 pi i 64 a / L_p),
  // B_p[b] = exp(-2 pi i b / L_p); tws holds A_p then B_p for every pass, tw_off[p] = start of A_p
  const double2* tws;
  int tws_n;                         // entries of tws
  int tw_off[kMaxRadixPasses];
  int tw_na[kMaxRadixPasses];        // entries of A_p
  unsigned magic[kMaxRadixPasses];   // ceil(2^32 / (L_p / R_p)): b / sub == __umulhi(b, magic) for b < 2^20
  int fuse[kMaxRadixPasses];         // pass p and p + 1 run as one in-register pass of radix R_p * R_{p+1} (fft_pass2_s)
  // split tables of the unpack / post-twiddle step, staged behind tws: unpackA[k] = UA_hi[k >> 6] * UA_lo[k & 63],
  // postB[k] = UB_hi[k >> 6] * UB_lo[k & 63] (scale in UB_lo); layout UA_hi[unp_na] UA_lo[64] UB_hi[unp_na] UB_lo[64]
  const double2* unp;
  int unp_na;
  // split Hamming table, staged behind unp: np.hamming(N)[i] = 0.54 - 0.46 Re(HC[i >> 6] * HD[i & 63]),
  // HC[a] = exp(j theta 64 a), HD[b] = exp(j theta b), theta = 2 pi / (N - 1); layout HC[ham_na] HD[64]
  const double2* hamt;
  int ham_na;
  double win_a0, win_a1;  // window = a0 - a1 cos(theta i): hamming 0.54/0.46, hanning 0.5/0.5, square 1/0
  const int32_t* perm;     // [n] position of X[k] after the DIF passes
  const double2* unpackA;  // [n+1] exp(-2 pi i k / N)
  const double2* postB;    // [N]  exp(-i pi k / (2N)) * 2 / sqrt(2N)
  double* out;             // [n_seg, out_ld]
  int out_ld;              // row stride of out (>= N)
  int out_pad;             // coefficients 0..out_pad-1 are repeated at N..N+out_pad-1 (circular extension)
  int vec_ok;              // wav base pointer is 16-byte aligned
  int seg0;                // first segment of this launch (chunked launches); out is indexed by blockIdx.x
  // dct_half_kernel: split table of the first radix-2 stage, w_n^j = T1A[j >> 6] * T1B[j & 63], layout T1A[tw1_na] T1B[64]
  const double2* tw1;
  int tw1_na;
};

template <typename T>
__device__ __forceinline__ cplx<T> load_tw(const double2* tw, int idx) {
  double2 w = __ldg(tw + idx);
  return {(T)w.x, (T)w.y};
}

template <typename T, int R>
__device__ __forceinline__ void dft_small(cplx<T>* a) {
  if constexpr (R == 2) {
    cplx<T> t = a[0];
    a[0] = cadd(t, a[1]);
    a[1] = csub(t, a[1]);
  } else if constexpr (R == 4) {
    cplx<T> t0 = cadd(a[0], a[2]), t1 = csub(a[0], a[2]);
    cplx<T> t2 = cadd(a[1], a[3]), t3 = mulmj(csub(a[1], a[3]));
    a[0] = cadd(t0, t2);
    a[2] = csub(t0, t2);
    a[1] = cadd(t1, t3);
    a[3] = csub(t1, t3);
  } else if constexpr (R == 3) {
    const T s = (T)0.86602540378443864676;
    cplx<T> t1 = cadd(a[1], a[2]);
    cplx<T> t2 = {a[0].x - (T)0.5 * t1.x, a[0].y - (T)0.5 * t1.y};
    cplx<T> t3 = mulmj(cscale(csub(a[1], a[2]), s));
    a[0] = cadd(a[0], t1);
    a[1] = cadd(t2, t3);
    a[2] = csub(t2, t3);
  } else if constexpr (R == 5) {
    const T c1 = (T)0.30901699437494742410, c2 = (T)-0.80901699437494742410;
    const T s1 = (T)0.95105651629515357212, s2 = (T)0.58778525229247312917;
    cplx<T> t1 = cadd(a[1], a[4]), t2 = cadd(a[2], a[3]);
    cplx<T> t3 = csub(a[1], a[4]), t4 = csub(a[2], a[3]);
    cplx<T> m1 = {a[0].x + c1 * t1.x + c2 * t2.x, a[0].y + c1 * t1.y + c2 * t2.y};
    cplx<T> m2 = {a[0].x + c2 * t1.x + c1 * t2.x, a[0].y + c2 * t1.y + c1 * t2.y};
    cplx<T> n1 = mulmj(cplx<T>{s1 * t3.x + s2 * t4.x, s1 * t3.y + s2 * t4.y});
    cplx<T> n2 = mulmj(cplx<T>{s2 * t3.x - s1 * t4.x, s2 * t3.y - s1 * t4.y});
    a[0] = cadd(a[0], cadd(t1, t2));
    a[1] = cadd(m1, n1);
    a[4] = csub(m1, n1);
    a[2] = cadd(m2, n2);
    a[3] = csub(m2, n2);
  }
}

// One decimation-in-frequency pass of radix R over sub-transforms of length L (in place).
template <typename T, int R>
__device__ __forceinline__ void fft_pass(cplx<T>* z, int n, int L, const double2* tw) {
  const int sub = L / R;
  const int nb = n / R;
  const int twstep = n / L;
  for (int b = threadIdx.x; b < nb; b += blockDim.x) {
    const int blk = b / sub, j = b - blk * sub;
    cplx<T>* p = z + blk * L + j;
    cplx<T> a[R];
#pragma unroll
    for (int t = 0; t < R; ++t) a[t] = p[t * sub];
    dft_small<T, R>(a);
    p[0] = a[0];
#pragma unroll
    for (int u = 1; u < R; ++u) {
      cplx<T> w = load_tw<T>(tw, j * u * twstep);
      p[u * sub] = cmul(a[u], w);
    }
  }
  __syncthreads();
}

// Same pass with every twiddle taken from the split tables in shared memory (no global loads inside the passes:
// with the single strided table in global memory the kernel spent 46 % of its issue slots waiting on those
// loads) and the powers w^2 .. w^(R-1) formed from w by multiplication.
template <typename T, int R>
__device__ __forceinline__ void fft_pass_s(cplx<T>* z, int n, int L, const double2* twA, const double2* twB,
                                           unsigned magic) {
  const int sub = L / R;
  const int nb = n / R;
  for (int b = threadIdx.x; b < nb; b += blockDim.x) {
    int blk = b, j = 0;
    if (sub > 1) { blk = (int)__umulhi((unsigned)b, magic); j = b - blk * sub; }
    cplx<T>* p = z + blk * L + j;
    cplx<T> a[R];
#pragma unroll
    for (int t = 0; t < R; ++t) a[t] = p[t * sub];
    dft_small<T, R>(a);
    p[0] = a[0];
    if (sub > 1) {
      const double2 wa = twA[j >> 6], wb = twB[j & 63];
      const cplx<double> w1 = {wa.x * wb.x - wa.y * wb.y, wa.x * wb.y + wa.y * wb.x};
      cplx<double> w = w1;
#pragma unroll
      for (int u = 1; u < R; ++u) {
        p[u * sub] = cmul(a[u], cplx<T>{(T)w.x, (T)w.y});
        if (u + 1 < R) w = cmul(w, w1);
      }
    } else {
#pragma unroll
      for (int u = 1; u < R; ++u) p[u] = a[u];
    }
  }
  __syncthreads();
}

// cos / sin of 2 pi m / R as compile-time constants (octant reduction on the integer fraction, then 12 Taylor terms
// on [0, pi/4]: below 1e-17)
constexpr double kTwoPi = 6.283185307179586476925286766559;
constexpr double cx_sin_small(double x) {
  double term = x, sum = x;
  const double x2 = x * x;
  for (int i = 1; i < 12; ++i) { term *= -x2 / (double)((2 * i) * (2 * i + 1)); sum += term; }
  return sum;
}
constexpr double cx_cos_small(double x) {
  double term = 1.0, sum = 1.0;
  const double x2 = x * x;
  for (int i = 1; i < 12; ++i) { term *= -x2 / (double)((2 * i - 1) * (2 * i)); sum += term; }
  return sum;
}
// which: 0 -> cos(2 pi m / R), 1 -> sin(2 pi m / R)
constexpr double cx_trig(int m, int R, int which) {
  m %= R;
  if (m < 0) m += R;
  int q = 4 * m;  // angle = 2 pi q / (4 R); quarter turn = R, half turn = 2 R
  double sc = 1.0, ss = 1.0;
  if (q > 2 * R) { q = 4 * R - q; ss = -1.0; }
  if (q > R) { q = 2 * R - q; sc = -1.0; }
  bool swap = false;
  if (2 * q > R) { q = R - q; swap = true; }
  const double x = kTwoPi * (double)q / (double)(4 * R);
  const double c = swap ? cx_sin_small(x) : cx_cos_small(x);
  const double sn = swap ? cx_cos_small(x) : cx_sin_small(x);
  return which == 0 ? sc * c : ss * sn;
}

// Two consecutive decimation-in-frequency passes (radix R1 over length L, then radix R2 over length L / R1) done by one
// thread on R1 * R2 elements in registers: the data layout, the twiddle tables and the output permutation are
// those of the two separate passes, but shared memory is read and written once instead of twice.
// Element (t, t2) sits at j + t * sub1 + t2 * sub2; the first-stage butterfly of column t2 is the one at
// j + t2 * sub2, whose twiddle w_L^((j + t2 sub2) u) = w_L^(j u) * w_(R1 R2)^(t2 u) -- the second factor a constant.
template <typename T, int R1, int R2>
__device__ __forceinline__ void fft_pass2_s(cplx<T>* z, int n, int L, const double2* twA1, const double2* twB1,
                                            const double2* twA2, const double2* twB2, unsigned magic2) {
  constexpr int R = R1 * R2;
  const int sub1 = L / R1, sub2 = sub1 / R2;
  const int nb = n / R;
  for (int b = threadIdx.x; b < nb; b += blockDim.x) {
    int blk = b, j = 0;
    if (sub2 > 1) { blk = (int)__umulhi((unsigned)b, magic2); j = b - blk * sub2; }
    cplx<T>* p = z + blk * L + j;
    cplx<T> a[R2][R1];
#pragma unroll
    for (int t2 = 0; t2 < R2; ++t2)
#pragma unroll
      for (int t = 0; t < R1; ++t) a[t2][t] = p[t * sub1 + t2 * sub2];
    cplx<double> pw[R1];
    {
      const double2 wa = twA1[j >> 6], wb = twB1[j & 63];
      pw[1] = {wa.x * wb.x - wa.y * wb.y, wa.x * wb.y + wa.y * wb.x};
#pragma unroll
      for (int u = 2; u < R1; ++u) pw[u] = cmul(pw[u - 1], pw[1]);
    }
#pragma unroll
    for (int t2 = 0; t2 < R2; ++t2) {
      dft_small<T, R1>(a[t2]);
#pragma unroll
      for (int u = 1; u < R1; ++u) {
        cplx<double> w = pw[u];
        if (t2 > 0) {
          constexpr int kR = R;
          const cplx<double> c = {cx_trig(t2 * u, kR, 0), -cx_trig(t2 * u, kR, 1)};
          w = cmul(w, c);
        }
        a[t2][u] = cmul(a[t2][u], cplx<T>{(T)w.x, (T)w.y});
      }
    }
    cplx<double> pw2[R2];
    if (sub2 > 1) {
      const double2 wa = twA2[j >> 6], wb = twB2[j & 63];
      pw2[1] = {wa.x * wb.x - wa.y * wb.y, wa.x * wb.y + wa.y * wb.x};
#pragma unroll
      for (int u2 = 2; u2 < R2; ++u2) pw2[u2] = cmul(pw2[u2 - 1], pw2[1]);
    }
#pragma unroll
    for (int u = 0; u < R1; ++u) {
      cplx<T> c[R2];
#pragma unroll
      for (int t2 = 0; t2 < R2; ++t2) c[t2] = a[t2][u];
      dft_small<T, R2>(c);
      p[u * sub1] = c[0];
#pragma unroll
      for (int u2 = 1; u2 < R2; ++u2)
        p[u * sub1 + u2 * sub2] = (sub2 > 1) ? cmul(c[u2], cplx<T>{(T)pw2[u2].x, (T)pw2[u2].y}) : c[u2];
    }
  }
  __syncthreads();
}

// all passes of a plan (single and fused), tables in shared memory at stw
template <typename T>
__device__ __forceinline__ void fft_run_passes(cplx<T>* z, const int n, const DctArgs& a, const double2* stw) {
  int L = n;
  for (int p = 0; p < a.npass;) {
    const int R = a.radix[p];
    const double2* twA = stw + a.tw_off[p];
    const double2* twB = twA + a.tw_na[p];
    if (a.fuse[p]) {
      const int R2 = a.radix[p + 1];
      const double2* twA2 = stw + a.tw_off[p + 1];
      const double2* twB2 = twA2 + a.tw_na[p + 1];
      const unsigned mg = a.magic[p + 1];
      switch (R * 8 + R2) {
        case 4 * 8 + 4: fft_pass2_s<T, 4, 4>(z, n, L, twA, twB, twA2, twB2, mg); break;
        case 5 * 8 + 3: fft_pass2_s<T, 5, 3>(z, n, L, twA, twB, twA2, twB2, mg); break;
        case 5 * 8 + 2: fft_pass2_s<T, 5, 2>(z, n, L, twA, twB, twA2, twB2, mg); break;
        case 4 * 8 + 3: fft_pass2_s<T, 4, 3>(z, n, L, twA, twB, twA2, twB2, mg); break;
        case 4 * 8 + 2: fft_pass2_s<T, 4, 2>(z, n, L, twA, twB, twA2, twB2, mg); break;
        case 3 * 8 + 3: fft_pass2_s<T, 3, 3>(z, n, L, twA, twB, twA2, twB2, mg); break;
        default: fft_pass2_s<T, 3, 2>(z, n, L, twA, twB, twA2, twB2, mg); break;
      }
      L /= R * R2;
      p += 2;
    } else {
      if (R == 4) fft_pass_s<T, 4>(z, n, L, twA, twB, a.magic[p]);
      else if (R == 2) fft_pass_s<T, 2>(z, n, L, twA, twB, a.magic[p]);
      else if (R == 3) fft_pass_s<T, 3>(z, n, L, twA, twB, a.magic[p]);
      else fft_pass_s<T, 5>(z, n, L, twA, twB, a.magic[p]);
      L /= R;
      ++p;
    }
  }
}

template <typename S>
__device__ __forceinline__ double sample_to_double(S v) { return (double)v; }

// Scalar sample fetch with reflect padding and optional noise mixing.
template <typename S>
__device__ __forceinline__ double fetch_sample(const DctArgs& a, const S* sig, int64_t q, int64_t T, int u,
                                               double alpha, const int16_t* nz) {
  q = reflect_index(q, T);
  double v = (double)sig[q];
  if (nz) v += alpha * (double)nz[q];
  return v;
}

#ifndef FDLP_DCT_MAXT
#define FDLP_DCT_MAXT 512
#endif
template <typename T, typename S>
__global__ void __launch_bounds__(FDLP_DCT_MAXT) dct_segments_kernel(const DctArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* sv = reinterpret_cast<T*>(smem_raw);  // v[] (real view) == z[] (complex view)
  cplx<T>* z = reinterpret_cast<cplx<T>*>(smem_raw);
  double2* stw = reinterpret_cast<double2*>(smem_raw + (size_t)a.N * sizeof(T));  // N * sizeof(T) is a multiple of 16
  for (int i = threadIdx.x; i < a.tws_n; i += blockDim.x) stw[i] = a.tws[i];
  double2* sunp = stw + a.tws_n;
  for (int i = threadIdx.x; i < 2 * (a.unp_na + 64); i += blockDim.x) sunp[i] = a.unp[i];
  double2* sham = sunp + 2 * (a.unp_na + 64);
  for (int i = threadIdx.x; i < a.ham_na + 64; i += blockDim.x) sham[i] = a.hamt[i];
  __syncthreads();

  const int seg = a.seg0 + blockIdx.x;
  const int u = a.seg_utt[seg];
  const int k = seg - a.utt_seg_off[u];
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T_utt = a.utt_wav_off[u + 1] - woff;
  const S* sig = reinterpret_cast<const S*>(a.wav) + woff;
  const int N = a.N;
  const int ext = (N % 2 == 0) ? N / 2 - 1 : (N - 1) / 2;
  const int64_t q_base = (int64_t)k * a.S - ext;  // utterance-relative index of window sample 0
  const int16_t* nz = a.noise ? a.noise + a.utt_noise_off[u] : nullptr;
  const double alpha = a.noise ? a.utt_alpha[u] : 0.0;

  // ---- load: the windowed segment, scattered into Makhoul order ----
  // A lane owns the sample pair (i, i+1), i = 2 * pair - sh, where sh makes the pair's global address even
  // (4-byte / 8-byte aligned loads, 128 / 256 contiguous bytes per warp).  Consecutive lanes then write
  // consecutive doubles of the ascending (even i) and of the descending (odd i) half of v[]: no bank conflicts.
  // The loads of kLdBatch pairs are issued together (the segment comes from HBM: one load per trip left the
  // kernel waiting on memory latency for a quarter of its time), and the window is formed from the split table
  // in shared memory; i & 63 is the same for every pair of a thread (512 threads x 2 samples = 0 mod 64).
  {
    constexpr int kLdBatch = 8;
    const int64_t g0 = woff + q_base;  // global sample index of window sample 0 (may be < 0)
    const int sh = (int)(((g0 % 2) + 2) % 2);
    const int npair = (N + sh + 1) / 2;
    const bool pair_ok = a.vec_ok;  // with noise mixing the noise excerpt (arbitrary offset) is read with scalar loads
    const double2 hd0 = sham[a.ham_na + ((2 * (int)threadIdx.x - sh) & 63)];
    const double2 hd1 = sham[a.ham_na + ((2 * (int)threadIdx.x - sh + 1) & 63)];
    for (int pb = threadIdx.x; pb < npair; pb += blockDim.x * kLdBatch) {
      double v0[kLdBatch], v1[kLdBatch];
      bool fast[kLdBatch];
      if constexpr (sizeof(S) == 2) {
        uint32_t raw[kLdBatch];
#pragma unroll
        for (int e = 0; e < kLdBatch; ++e) {
          const int i0 = 2 * (pb + e * (int)blockDim.x) - sh;
          const int64_t q0 = q_base + i0;
          fast[e] = pair_ok && i0 >= 0 && i0 + 2 <= N && q0 >= 0 && q0 + 2 <= T_utt;
          raw[e] = fast[e] ? __ldg(reinterpret_cast<const uint32_t*>(sig + q0)) : 0u;
        }
#pragma unroll
        for (int e = 0; e < kLdBatch; ++e) {
          v0[e] = (double)(int16_t)(raw[e] & 0xffffu);
          v1[e] = (double)(int16_t)(raw[e] >> 16);
        }
      } else if constexpr (sizeof(S) == 4) {
        float2 raw[kLdBatch];
#pragma unroll
        for (int e = 0; e < kLdBatch; ++e) {
          const int i0 = 2 * (pb + e * (int)blockDim.x) - sh;
          const int64_t q0 = q_base + i0;
          fast[e] = pair_ok && i0 >= 0 && i0 + 2 <= N && q0 >= 0 && q0 + 2 <= T_utt;
          raw[e] = fast[e] ? __ldg(reinterpret_cast<const float2*>(sig + q0)) : make_float2(0.f, 0.f);
        }
#pragma unroll
        for (int e = 0; e < kLdBatch; ++e) { v0[e] = (double)raw[e].x; v1[e] = (double)raw[e].y; }
      } else {  // float64 samples (the reverberated waveform, fdlp_reverb_forward)
#pragma unroll
        for (int e = 0; e < kLdBatch; ++e) {
          const int i0 = 2 * (pb + e * (int)blockDim.x) - sh;
          const int64_t q0 = q_base + i0;
          fast[e] = pair_ok && i0 >= 0 && i0 + 2 <= N && q0 >= 0 && q0 + 2 <= T_utt;
          const double2 raw = fast[e] ? __ldg(reinterpret_cast<const double2*>(sig + q0)) : make_double2(0.0, 0.0);
          v0[e] = raw.x; v1[e] = raw.y;
        }
      }
      if (nz) {  // sig + alpha * noise (features.py:29-31) on the fast pairs; the loads of a batch are issued together
        int n0[kLdBatch], n1[kLdBatch];
#pragma unroll
        for (int e = 0; e < kLdBatch; ++e) {
          const int64_t q0 = q_base + 2 * (pb + e * (int)blockDim.x) - sh;
          n0[e] = fast[e] ? (int)__ldg(nz + q0) : 0;
          n1[e] = fast[e] ? (int)__ldg(nz + q0 + 1) : 0;
        }
#pragma unroll
        for (int e = 0; e < kLdBatch; ++e) { v0[e] += alpha * (double)n0[e]; v1[e] += alpha * (double)n1[e]; }
      }
#pragma unroll
      for (int e = 0; e < kLdBatch; ++e) {
        const int pi = pb + e * (int)blockDim.x;
        if (pi < npair) {
          const int i0 = 2 * pi - sh, i1 = i0 + 1;
          const int64_t q0 = q_base + i0;
          if (!fast[e]) {
            v0[e] = (i0 >= 0 && i0 < N) ? fetch_sample<S>(a, sig, q0, T_utt, u, alpha, nz) : 0.0;
            v1[e] = (i1 >= 0 && i1 < N) ? fetch_sample<S>(a, sig, q0 + 1, T_utt, u, alpha, nz) : 0.0;
          }
          if (i0 >= 0 && i0 < N) {
            const double2 hc = sham[i0 >> 6];
            const double w = a.win_a0 - a.win_a1 * (hc.x * hd0.x - hc.y * hd0.y);
            sv[(i0 & 1) ? (N - 1 - (i0 >> 1)) : (i0 >> 1)] = (T)(v0[e] * w);
          }
          if (i1 >= 0 && i1 < N) {
            const double2 hc = sham[i1 >> 6];
            const double w = a.win_a0 - a.win_a1 * (hc.x * hd1.x - hc.y * hd1.y);
            sv[(i1 & 1) ? (N - 1 - (i1 >> 1)) : (i1 >> 1)] = (T)(v1[e] * w);
          }
        }
      }
    }
  }
  __syncthreads();

  // ---- N/2-point complex FFT, in place, decimation in frequency ----
  fft_run_passes<T>(z, a.n, a, stw);

  // ---- unpack to the length-N real FFT, Makhoul post-twiddle, scale, store ----
  double* out = a.out + (size_t)blockIdx.x * a.out_ld;
  const int n = a.n;
  constexpr int kUnpBatch = 4;
  for (int kb = threadIdx.x; kb <= n; kb += blockDim.x * kUnpBatch) {
   int p1[kUnpBatch], p2[kUnpBatch];
#pragma unroll
   for (int e = 0; e < kUnpBatch; ++e) {  // the permutation indices come from L2: issue the loads of a batch together
     const int kq = kb + e * (int)blockDim.x;
     const bool in = kq <= n;
     p1[e] = in ? __ldg(a.perm + ((kq == n) ? 0 : kq)) : 0;
     p2[e] = in ? __ldg(a.perm + ((kq == 0) ? 0 : n - kq)) : 0;
   }
#pragma unroll
   for (int e = 0; e < kUnpBatch; ++e) {
    const int kq = kb + e * (int)blockDim.x;
    if (kq > n) break;
    const cplx<T> zk = z[p1[e]];
    cplx<T> zc = z[p2[e]];
    zc.y = -zc.y;
    const cplx<double> Zk = {(double)zk.x, (double)zk.y}, Zc = {(double)zc.x, (double)zc.y};
    const cplx<double> E = {0.5 * (Zk.x + Zc.x), 0.5 * (Zk.y + Zc.y)};
    const cplx<double> D = {0.5 * (Zk.x - Zc.x), 0.5 * (Zk.y - Zc.y)};
    const double2 ah = sunp[kq >> 6], al = sunp[a.unp_na + (kq & 63)];
    const double2 A = make_double2(ah.x * al.x - ah.y * al.y, ah.x * al.y + ah.y * al.x);
    // V = E - j * A * D
    const cplx<double> AD = cmul(cplx<double>{A.x, A.y}, D);
    const cplx<double> V = {E.x + AD.y, E.y - AD.x};
    const double2 bh = sunp[a.unp_na + 64 + (kq >> 6)], bl = sunp[2 * a.unp_na + 64 + (kq & 63)];
    const double2 B = make_double2(bh.x * bl.x - bh.y * bl.y, bh.x * bl.y + bh.y * bl.x);
    const double yk = B.x * V.x - B.y * V.y;  // Re(B * V), B carries 2/sqrt(2N)
    out[kq] = yk;
    if (kq < a.out_pad) out[N + kq] = yk;
    if (kq > 0 && kq < n) {
      // postB[N - kq] = -j conj(postB[kq]) = (-B.y, -B.x);  Re(B2 * conj(V)) = B2.x V.x + B2.y V.y
      out[N - kq] = -(B.y * V.x + B.x * V.y);
    }
   }
  }
}

// Segments whose N float64 values do not fit in shared memory (N > ~28 k): two CTAs per segment, each holding one
// half of the first decimation-in-frequency stage of the N/2-point transform,
//     e[j] = z[j] + z[j + n/2]  -> Z[2k],      o[j] = (z[j] - z[j + n/2]) w_n^j  -> Z[2k+1],     j < n/2,
// formed while the samples are loaded (first the samples that land in z[j], then, after a barrier, those of
// z[j + n/2]: every slot is touched by one thread per phase).  The unpack step pairs Z[k] with Z[n - k], which have
// the same parity (n is even), so each CTA finishes the DCT coefficients k = h (mod 2) on its own: no exchange
// between the halves.  The passes, tables and permutation are those of an n/2-point plan.
template <typename S>
__global__ void __launch_bounds__(512) dct_half_kernel(const DctArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int N = a.N, n = a.n, n2 = a.n / 2;
  double* sv = reinterpret_cast<double*>(smem_raw);
  cplx<double>* z = reinterpret_cast<cplx<double>*>(smem_raw);
  double2* stw = reinterpret_cast<double2*>(smem_raw + (size_t)n2 * sizeof(cplx<double>));
  for (int i = threadIdx.x; i < a.tws_n; i += blockDim.x) stw[i] = a.tws[i];
  double2* sunp = stw + a.tws_n;
  for (int i = threadIdx.x; i < 2 * (a.unp_na + 64); i += blockDim.x) sunp[i] = a.unp[i];
  double2* sham = sunp + 2 * (a.unp_na + 64);
  for (int i = threadIdx.x; i < a.ham_na + 64; i += blockDim.x) sham[i] = a.hamt[i];
  double2* st1 = sham + a.ham_na + 64;
  for (int i = threadIdx.x; i < a.tw1_na + 64; i += blockDim.x) st1[i] = a.tw1[i];

  const int seg_local = blockIdx.x >> 1, h = blockIdx.x & 1;
  const int seg = a.seg0 + seg_local;
  const int u = a.seg_utt[seg];
  const int k = seg - a.utt_seg_off[u];
  const int64_t woff = a.utt_wav_off[u];
  const int64_t T_utt = a.utt_wav_off[u + 1] - woff;
  const S* sig = reinterpret_cast<const S*>(a.wav) + woff;
  const int ext = (N % 2 == 0) ? N / 2 - 1 : (N - 1) / 2;
  const int64_t q_base = (int64_t)k * a.S - ext;
  const int16_t* nz = a.noise ? a.noise + a.utt_noise_off[u] : nullptr;
  const double alpha = a.noise ? a.utt_alpha[u] : 0.0;
  __syncthreads();

  for (int phase = 0; phase < 2; ++phase) {
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
      const int idx = (i & 1) ? (N - 1 - (i >> 1)) : (i >> 1);  // Makhoul order: v[idx] = x[i]
      const int m = idx >> 1;                                   // z[m] = v[2m] + j v[2m+1]
      const int role = (m >= n2) ? 1 : 0;
      if (role != phase) continue;
      const int64_t q = q_base + i;
      double v;
      if (q >= 0 && q < T_utt) {
        v = (double)sig[q];
        if (nz) v += alpha * (double)nz[q];
      } else {
        v = fetch_sample<S>(a, sig, q, T_utt, u, alpha, nz);
      }
      const double2 hc = sham[i >> 6], hd = sham[a.ham_na + (i & 63)];
      v *= a.win_a0 - a.win_a1 * (hc.x * hd.x - hc.y * hd.y);
      const int slot = 2 * (m - role * n2) + (idx & 1);
      if (phase == 0) sv[slot] = v;
      else sv[slot] = h ? sv[slot] - v : sv[slot] + v;
    }
    __syncthreads();
  }
  if (h) {
    for (int j = threadIdx.x; j < n2; j += blockDim.x) {
      const double2 wa = st1[j >> 6], wb = st1[a.tw1_na + (j & 63)];
      z[j] = cmul(z[j], cplx<double>{wa.x * wb.x - wa.y * wb.y, wa.x * wb.y + wa.y * wb.x});
    }
    __syncthreads();
  }

  fft_run_passes<double>(z, n2, a, stw);

  // ---- unpack: the coefficients kq = h (mod 2), same arithmetic as dct_segments_kernel ----
  double* out = a.out + (size_t)seg_local * a.out_ld;
  for (int kq = 2 * (int)threadIdx.x + h; kq <= n; kq += 2 * (int)blockDim.x) {
    const int k1 = (kq == n) ? 0 : kq, k2 = (kq == 0) ? 0 : n - kq;
    const cplx<double> Zk = z[__ldg(a.perm + ((k1 - h) >> 1))];
    cplx<double> Zc = z[__ldg(a.perm + ((k2 - h) >> 1))];
    Zc.y = -Zc.y;
    const cplx<double> E = {0.5 * (Zk.x + Zc.x), 0.5 * (Zk.y + Zc.y)};
    const cplx<double> D = {0.5 * (Zk.x - Zc.x), 0.5 * (Zk.y - Zc.y)};
    const double2 ah = sunp[kq >> 6], al = sunp[a.unp_na + (kq & 63)];
    const double2 A = make_double2(ah.x * al.x - ah.y * al.y, ah.x * al.y + ah.y * al.x);
    const cplx<double> AD = cmul(cplx<double>{A.x, A.y}, D);
    const cplx<double> V = {E.x + AD.y, E.y - AD.x};
    const double2 bh = sunp[a.unp_na + 64 + (kq >> 6)], bl = sunp[2 * a.unp_na + 64 + (kq & 63)];
    const double2 B = make_double2(bh.x * bl.x - bh.y * bl.y, bh.x * bl.y + bh.y * bl.x);
    const double yk = B.x * V.x - B.y * V.y;
    out[kq] = yk;
    if (kq < a.out_pad) out[N + kq] = yk;
    if (kq > 0 && kq < n) out[N - kq] = -(B.y * V.x + B.x * V.y);
  }
}

}  // namespace fdlp
