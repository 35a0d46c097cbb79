// K4 + K5: per-(segment, band) Levinson-Durbin LPC, LPC gain, LPC->cepstrum recursion, cepstral weighting
// and the Hilbert-envelope evaluation -- one warp per (segment, band), float64 throughout.
//
// Reference: computeLpcFast        src/featgen/features.py:226-228  (solve_toeplitz + the gain quirk
//                                  gg = y[0] + sum_{k=0..p} xlpc[k] * y[k+1])
//            computeModSpecFromLpc src/featgen/features.py:233-246
//            weighting             src/featgen/computeFDLPSpectrogram.py:194-200
//            envelope              src/featgen/computeFDLPSpectrogram.py:201-205
//              abs(exp(fft(ms, M)))[t] == exp(sum_n ms[n] cos(2 pi n t / M)) because ms is real, so the
//              envelope is a real contraction of the weighted cepstrum against a cosine basis, then exp,
//              then the hanning(kk)/hamming(kk) window ratio.
#pragma once
#include "common.cuh"

namespace fdlp {

struct LpcArgs {
  const double* r;        // [n_tasks, nlags]
  int n_tasks, J;
  int order, nlags, L, M, kk;
  const double* weights;  // [L]
  const double* cosM;     // [M] cos(2 pi q / M)
  const double* wratio;   // [kk] np.hanning(kk) / np.hamming(kk)
  double* lpc_out;        // [n_tasks, order+1] or null
  double* gain_out;       // [n_tasks] or null
  double* cep_out;        // [n_tasks, L] or null (before weighting)
  double* env;            // [n_tasks, kk]
  int32_t* status;        // status words to OR FDLP_ST_* bits into, or null
  const int32_t* seg_utt; // [n_seg] segment -> status index (utterance); null: index by segment
};

constexpr int kLpcWarps = 8;

__host__ __device__ inline size_t lpc_smem_doubles_per_warp(int order, int nlags, int L) {
  return (size_t)nlags + 2 * (size_t)(order + 1) + 2 * (size_t)L;
}

__global__ void __launch_bounds__(kLpcWarps * 32) lpc_envelope_kernel(const LpcArgs a) {
  extern __shared__ __align__(16) double sm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int p = a.order, L = a.L, M = a.M;

  double* s_cos = sm;  // [M]
  double* base = sm + M + (size_t)warp * lpc_smem_doubles_per_warp(p, a.nlags, L);
  double* s_r = base;                 // [nlags]
  double* s_a0 = s_r + a.nlags;       // [p+1]
  double* s_a1 = s_a0 + (p + 1);      // [p+1]
  double* s_cep = s_a1 + (p + 1);     // [L]
  double* s_mc = s_cep + L;           // [L]  m * cep[m]

  for (int q = threadIdx.x; q < M; q += blockDim.x) s_cos[q] = a.cosM[q];
  __syncthreads();

  const int task = blockIdx.x * kLpcWarps + warp;
  if (task >= a.n_tasks) return;

  const double* rg = a.r + (size_t)task * a.nlags;
  for (int i = lane; i < a.nlags; i += 32) s_r[i] = rg[i];
  for (int i = lane; i <= p; i += 32) { s_a0[i] = (i == 0) ? 1.0 : 0.0; s_a1[i] = (i == 0) ? 1.0 : 0.0; }
  __syncwarp();

  // ---- Levinson-Durbin: T a = -r[1..p] ----
  double* cur = s_a0;
  double* nxt = s_a1;
  double E = s_r[0];
  int bad = 0;
  for (int m = 1; m <= p; ++m) {
    double part = 0.0;
    for (int i = 1 + lane; i < m; i += 32) part = fma(cur[i], s_r[m - i], part);
    const double acc = warp_sum(part) + s_r[m];
    if (E == 0.0 || !isfinite(E)) bad = FDLP_ST_SINGULAR;
    const double k = -acc / E;
    for (int i = 1 + lane; i < m; i += 32) nxt[i] = fma(k, cur[m - i], cur[i]);
    if (lane == 0) nxt[m] = k;
    E *= (1.0 - k * k);
    __syncwarp();
    double* t = cur; cur = nxt; nxt = t;
  }
  // cur[0..p] = xlpc (cur[0] == 1)

  // ---- gain: gg = r[0] + sum_{k=0..p} xlpc[k] * r[k+1]  (features.py:228) ----
  double gpart = 0.0;
  for (int i = lane; i <= p; i += 32) gpart = fma(cur[i], s_r[i + 1], gpart);
  const double gg = s_r[0] + warp_sum(gpart);
  if (!(gg > 0.0) || !isfinite(gg)) bad |= FDLP_ST_NONFINITE;

  if (a.lpc_out) for (int i = lane; i <= p; i += 32) a.lpc_out[(size_t)task * (p + 1) + i] = cur[i];
  if (a.gain_out && lane == 0) a.gain_out[task] = gg;

  // ---- LPC -> cepstrum (features.py:233-246); x[0] = 1, x[i] = -a[i], zero beyond the order ----
  auto xval = [&](int i) -> double { return (i <= p) ? -cur[i] : 0.0; };
  if (lane == 0) {
    s_cep[0] = log(sqrt(gg));
    s_mc[0] = 0.0;
    if (L > 1) { s_cep[1] = xval(1); s_mc[1] = s_cep[1]; }
  }
  __syncwarp();
  for (int n = 2; n < L; ++n) {
    double part = 0.0;
    for (int m = 1 + lane; m < n; m += 32) part = fma(xval(n - m), s_mc[m], part);
    const double v = warp_sum(part) / (double)n + xval(n);
    if (lane == 0) { s_cep[n] = v; s_mc[n] = v * (double)n; }
    __syncwarp();
  }
  if (a.cep_out) for (int i = lane; i < L; i += 32) a.cep_out[(size_t)task * L + i] = s_cep[i];
  __syncwarp();

  // ---- weighting (mask * lifter * gamma, odd-zero folded in) ----
  for (int i = lane; i < L; i += 32) s_mc[i] = s_cep[i] * a.weights[i];
  __syncwarp();

  // ---- envelope: exp(sum_n ms[n] cos(2 pi n t / M)) * hanning/hamming ----
  double* eout = a.env + (size_t)task * a.kk;
  for (int t = lane; t < a.kk; t += 32) {
    double s = 0.0;
    int idx = 0;
    const int step = t % M;
    for (int n = 0; n < L; ++n) {
      s = fma(s_mc[n], s_cos[idx], s);
      idx += step;
      if (idx >= M) idx -= M;
    }
    eout[t] = exp(s) * a.wratio[t];
  }
  if (bad && a.status && lane == 0) {
    const int seg = task / a.J;
    atomicOr(a.status + (a.seg_utt ? a.seg_utt[seg] : seg), bad);
  }
}

}  // namespace fdlp
