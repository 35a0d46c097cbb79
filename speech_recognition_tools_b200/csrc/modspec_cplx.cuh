// Complex-modulation path of the modulation-spectrum extractor (--complex_modulation).
//
// Reference: src/featgen/computeModulationSpectrum.py:45-48 (filterbank over int(fduration*srate) points),
// :143-152 (hanning / square frames, ifft, first half), :162-173 (band = filt * spectrum, computeLpcFast with
// keepreal=False, computeModSpecFromLpc, |c| or real parts then imaginary parts); features.py:222-246.
//
//   window_frames_kernel   the windowed, reflect-padded frames [n_seg, N] float64 (the caller takes their inverse
//                          DFT; the first N/2 bins are the kernel's input below)
//   modspec_complex_kernel one CTA per (frame, band): the band's non-zero taps times the spectrum staged in shared
//                          memory, the circular autocorrelation r[0..p+1] = sum_n x[(n+k) mod M] conj(x[n]) by
//                          direct sums (one lag per thread; the reference's fft / ifft pair gives the same numbers
//                          to rounding), then warp 0: Levinson-Durbin on the Hermitian Toeplitz system
//                          (scipy.linalg.solve_toeplitz(y[0:p], -y[1:p+1]) takes the first row as conj(column)),
//                          gain, complex cepstrum recursion, selection and scaling.
#pragma once
#include "common.cuh"

namespace fdlp {

struct FramesArgs {
  const void* wav;
  const int64_t* utt_wav_off;
  const int32_t* utt_seg_off;
  const int32_t* seg_utt;
  int N, S;
  double a0, a1;  // window = a0 - a1 cos(2 pi i / (N - 1))
  double* out;    // [n_seg, N]
  int64_t n_seg;
};

template <typename S>
__global__ void __launch_bounds__(256) window_frames_kernel(const FramesArgs a) {
  const int N = a.N;
  const int ext = (N % 2 == 0) ? N / 2 - 1 : (N - 1) / 2;
  const double inv = N > 1 ? 2.0 / (double)(N - 1) : 0.0;
  for (int64_t seg = blockIdx.x; seg < a.n_seg; seg += gridDim.x) {
    const int u = a.seg_utt[seg];
    const int k = (int)(seg - a.utt_seg_off[u]);
    const int64_t woff = a.utt_wav_off[u];
    const int64_t T = a.utt_wav_off[u + 1] - woff;
    const S* sig = reinterpret_cast<const S*>(a.wav) + woff;
    const int64_t q_base = (int64_t)k * a.S - ext;
    double* o = a.out + seg * (int64_t)N;
    for (int i = threadIdx.x; i < N; i += blockDim.x) {
      const double w = a.a0 - a.a1 * cospi(inv * (double)i);
      o[i] = w * (double)sig[reflect_index(q_base + i, T)];
    }
  }
}

struct CplxModspecArgs {
  const double2* spec;  // [n_seg, ld] complex; bins 0..M-1 are read
  int64_t ld;
  const double* fbank;  // [J, M]
  const int32_t* row_lo;
  const int32_t* row_hi;  // taps [lo, hi) of band j are non-zero
  int M, J, order, lim, n_sel, absolute, max_width;
  const int32_t* sel_idx;
  const double* sel_scale;
  float* out;  // [n_seg, J * feat_len], feat_len = absolute ? n_sel : 2 * n_sel
  int32_t* status;
  const int32_t* seg_utt;
  int64_t n_tasks;
};

__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
  return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ double2 cmul_conj(double2 a, double2 b) {  // a * conj(b)
  return make_double2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y);
}
__device__ __forceinline__ double2 warp_sum2(double2 v) {
  return make_double2(warp_sum(v.x), warp_sum(v.y));
}

__global__ void modspec_complex_kernel(const CplxModspecArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* x = reinterpret_cast<double2*>(smem_raw);  // [max_width]
  double2* r = x + a.max_width;                       // [p + 2]
  double2* lp = r + (a.order + 2);                    // [max(p, lim) + 2]: 1, a_1 .. a_p, zeros
  double2* cep = lp + (max(a.order, a.lim) + 2);      // [lim]
  const int p = a.order, M = a.M, lim = a.lim;
  const int lane = threadIdx.x & 31;
  const int feat_len = a.absolute ? a.n_sel : 2 * a.n_sel;

  for (int64_t task = blockIdx.x; task < a.n_tasks; task += gridDim.x) {
    const int64_t seg = task / a.J;
    const int j = (int)(task - seg * a.J);
    const int lo = a.row_lo[j], hi = a.row_hi[j];
    const double2* sp = a.spec + seg * a.ld;
    const double* fb = a.fbank + (int64_t)j * M;
    for (int n = lo + threadIdx.x; n < hi; n += blockDim.x) {
      const double f = fb[n];
      const double2 s = sp[n];
      x[n - lo] = make_double2(f * s.x, f * s.y);
    }
    __syncthreads();
    // ---- circular autocorrelation, lag k on thread k ----
    if ((int)threadIdx.x < p + 2) {
      const int k = threadIdx.x;
      double2 acc = make_double2(0.0, 0.0);
      for (int n = lo; n < hi - k; ++n) {
        const double2 v = cmul_conj(x[n + k - lo], x[n - lo]);
        acc.x += v.x; acc.y += v.y;
      }
      for (int n = max(lo, M + lo - k); n < hi; ++n) {  // (n + k) mod M falls back into [lo, hi)
        const int t = n + k - M;
        if (t < hi) {
          const double2 v = cmul_conj(x[t - lo], x[n - lo]);
          acc.x += v.x; acc.y += v.y;
        }
      }
      r[k] = acc;
    }
    __syncthreads();
    if (threadIdx.x < 32) {
      // ---- Levinson-Durbin, Hermitian Toeplitz ----
      for (int i = lane; i < max(p, lim) + 2; i += 32) lp[i] = make_double2(i == 0 ? 1.0 : 0.0, 0.0);
      __syncwarp();
      double E = r[0].x;
      int st = 0;
      for (int m = 1; m <= p; ++m) {
        double2 acc = make_double2(0.0, 0.0);
        for (int i = 1 + lane; i < m; i += 32) {
          const double2 v = cmul(lp[i], r[m - i]);
          acc.x += v.x; acc.y += v.y;
        }
        acc = warp_sum2(acc);
        acc.x += r[m].x; acc.y += r[m].y;
        if (!(fabs(E) > 0.0) || !isfinite(E)) st |= 1;
        const double2 kk = make_double2(-acc.x / E, -acc.y / E);
        for (int i = 1 + lane; 2 * i <= m; i += 32) {
          const double2 ai = lp[i], am = lp[m - i];
          const double2 ui = cmul_conj(kk, am), um = cmul_conj(kk, ai);
          lp[i] = make_double2(ai.x + ui.x, ai.y + ui.y);
          if (2 * i != m) lp[m - i] = make_double2(am.x + um.x, am.y + um.y);
        }
        if (lane == 0) lp[m] = kk;
        E *= 1.0 - (kk.x * kk.x + kk.y * kk.y);
        __syncwarp();
      }
      // gg = y[0] + sum(xlpc * y[1:p+2])
      double2 g = make_double2(0.0, 0.0);
      for (int i = lane; i <= p; i += 32) {
        const double2 v = cmul(lp[i], r[i + 1]);
        g.x += v.x; g.y += v.y;
      }
      g = warp_sum2(g);
      g.x += r[0].x; g.y += r[0].y;
      // ---- cepstrum: xlpc[1:] = -xlpc[1:], c0 = log(sqrt(gg)), c_n = sum_{i<n} (i/n) x[n-i] c_i + x[n] ----
      for (int i = 1 + lane; i <= p; i += 32) lp[i] = make_double2(-lp[i].x, -lp[i].y);
      if (lane == 0) {
        const double mag = hypot(g.x, g.y);
        const double sr = sqrt(0.5 * (mag + g.x)), si = copysign(sqrt(0.5 * (mag - g.x)), g.y);  // principal sqrt
        cep[0] = make_double2(log(hypot(sr, si)), atan2(si, sr));
      }
      __syncwarp();
      if (lane == 0 && lim > 1) cep[1] = lp[1];
      __syncwarp();
      for (int n = 2; n < lim; ++n) {
        double2 acc = make_double2(0.0, 0.0);
        const double invn = 1.0 / (double)n;
        for (int i = max(1, n - p) + lane; i < n; i += 32) {
          const double2 xv = lp[n - i];
          const double w = (double)i * invn;
          const double2 v = cmul(make_double2(w * xv.x, w * xv.y), cep[i]);
          acc.x += v.x; acc.y += v.y;
        }
        acc = warp_sum2(acc);
        if (lane == 0) cep[n] = make_double2(acc.x + lp[n].x, acc.y + lp[n].y);
        __syncwarp();
      }
      // ---- selection ----
      float* o = a.out + seg * (int64_t)(a.J * feat_len) + (int64_t)j * feat_len;
      for (int q = lane; q < a.n_sel; q += 32) {
        const double2 c = cep[a.sel_idx[q]];
        const double s = a.sel_scale[q];
        const double cr = c.x * s, ci = c.y * s;
        if (!isfinite(cr) || !isfinite(ci)) st |= 2;
        if (a.absolute) o[q] = (float)hypot(cr, ci);
        else { o[q] = (float)cr; o[a.n_sel + q] = (float)ci; }
      }
      st = __reduce_or_sync(0xffffffffu, st);
      if (lane == 0 && st && a.status) atomicOr(a.status + a.seg_utt[seg], st);
    }
    __syncthreads();
  }
}

}  // namespace fdlp
