// K4 + K5: per-(segment, band) Levinson-Durbin LPC, LPC gain, LPC->cepstrum recursion and cepstral
// weighting (K4), then the Hilbert-envelope evaluation as a dense FP64 tensor-core contraction (K5).
//
// Reference: computeLpcFast        src/featgen/features.py:226-228  (solve_toeplitz + the gain quirk
//                                  gg = y[0] + sum_{k=0..p} xlpc[k] * y[k+1])
//            computeModSpecFromLpc src/featgen/features.py:233-246
//            weighting             src/featgen/computeFDLPSpectrogram.py:194-200
//            envelope              src/featgen/computeFDLPSpectrogram.py:201-205
//              abs(exp(fft(ms, M)))[t] == exp(sum_n ms[n] cos(2 pi n t / M)) because ms is real, so the
//              envelope is a real contraction of the weighted cepstrum against a cosine basis, then exp,
//              then the hanning(kk)/hamming(kk) window ratio.
//
// K4 is a chain of order-long serial recursions, so it is instruction-issue bound, not flop bound.  Each
// (segment, band) is owned by a GROUP of 8 lanes (4 solves per warp): the length-m dot products and
// updates keep most lanes busy for m up to 150..300, the tree reduction is 3 shuffle stages instead of 5,
// and the reciprocal of the prediction error is computed once for 4 solves.  All state lives in shared
// memory; the Levinson update is done in place on the symmetric pairs (i, m-i).
// K5 is a GEMM [tasks x L] x [L x kk] on DMMA.8x8x4: float64 holds the tolerance by construction (the
// TF32 / BF16x3 alternatives named in the north star are evaluated numerically in
// tests/test_envelope_precision.py: TF32 5.4e-3 fails, BF16x3 1.6e-5 holds but buys nothing at 0.4 % of the flops).
#pragma once
#include "autocorr.cuh"
#include "common.cuh"

namespace fdlp {

struct LpcArgs {
  const double* r;        // [n_tasks, nlags]
  int n_tasks, J;
  int order, nlags, L, Lp;  // Lp = L rounded up to a multiple of 4 (row stride of ms)
  const double* weights;  // [L]
  const double* inv_n;    // [L] 1/n (inv_n[0] unused)
  double* lpc_out;        // [n_tasks, order+1] or null
  double* gain_out;       // [n_tasks] or null
  double* cep_out;        // [n_tasks, L] or null (before weighting)
  double* ms;             // [n_tasks_padded, Lp] weighted cepstrum (zero padded), input of K5
  int32_t* status;        // status words to OR FDLP_ST_* bits into, or null
  const int32_t* seg_utt; // [n_seg] segment -> status index (utterance); null: index by segment
};

constexpr int kLpcMaxThreads = 256;
constexpr int kLpcGroup = 8;  // lanes per (segment, band); a CTA of T threads owns T/8 solves

// per-task shared memory (doubles): r | a (length max(p+1, L)) | cep/mc ; stride == 8 (mod 16) so that the
// two groups of a half-warp fall in disjoint banks
__host__ __device__ inline int lpc_task_stride(int order, int nlags, int L) {
  const int alen = (order + 1 > L) ? order + 1 : L;
  int s = nlags + alen + L;
  s += (8 - (s % 16) + 16) % 16;
  return s;
}

__device__ __forceinline__ double group_sum8(double v) {
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v;
}

__global__ void __launch_bounds__(kLpcMaxThreads) lpc_cepstrum_kernel(const LpcArgs a) {
  extern __shared__ __align__(16) double sm[];
  const int g = threadIdx.x & (kLpcGroup - 1);
  const int slot = threadIdx.x / kLpcGroup;
  const int p = a.order, L = a.L;
  const int alen = (p + 1 > L) ? p + 1 : L;
  double* s_r = sm + (size_t)slot * lpc_task_stride(p, a.nlags, L);
  double* s_a = s_r + a.nlags;
  double* s_mc = s_a + alen;

  int task = blockIdx.x * (blockDim.x / kLpcGroup) + slot;
  const bool live = task < a.n_tasks;
  if (!live) task = a.n_tasks - 1;  // idle groups shadow the last task (keeps warp-level shuffles uniform)

  const double* rg = a.r + (size_t)task * a.nlags;
  for (int i = g; i < a.nlags; i += kLpcGroup) s_r[i] = rg[i];
  for (int i = g; i < alen; i += kLpcGroup) s_a[i] = (i == 0) ? 1.0 : 0.0;
  __syncwarp();

  // ---- Levinson-Durbin: T a = -r[1..p], in place ----
  // One pass per order m: the symmetric pairs (i, m-i) are updated in place and, with the new values still in
  // registers, multiplied into the NEXT order's inner product acc_{m+1} = r[m+1] + sum_i a'[i] r[m+1-i], so every
  // coefficient is read and written once per order.  k = -acc * (1/E) with the reciprocal from __drcp_rn.
  double E = s_r[0];
  int bad = 0;
  double acc = s_r[1];
  for (int m = 1; m <= p; ++m) {
    if (E == 0.0 || !isfinite(E)) bad = FDLP_ST_SINGULAR;
    const double k = -acc * __drcp_rn(E);
    const double* rn = s_r + 1;  // rn[i] = r[i+1]
    double part = 0.0;
#pragma unroll 2
    for (int i = 1 + g; 2 * i < m; i += kLpcGroup) {
      const int jx = m - i;
      const double ai = s_a[i], aj = s_a[jx];
      const double ni = fma(k, aj, ai), nj = fma(k, ai, aj);
      s_a[i] = ni;
      s_a[jx] = nj;
      part = fma(ni, rn[jx], part);
      part = fma(nj, rn[i], part);
    }
    if (g == 0) {
      if ((m & 1) == 0) {
        const int h = m >> 1;
        const double ah = s_a[h];
        const double nh = fma(k, ah, ah);
        s_a[h] = nh;
        part = fma(nh, rn[h], part);
      }
      s_a[m] = k;
      part = fma(k, s_r[1], part);
    }
    acc = group_sum8(part) + s_r[m + 1];  // nlags = p + 2, so r[p+1] exists (the value is unused after m = p)
    E *= (1.0 - k * k);
    __syncwarp();
  }

  // ---- gain: gg = r[0] + sum_{k=0..p} xlpc[k] * r[k+1]  (features.py:228) ----
  double gpart = 0.0;
  for (int i = g; i <= p; i += kLpcGroup) gpart = fma(s_a[i], s_r[i + 1], gpart);
  const double gg = s_r[0] + group_sum8(gpart);
  if (!(gg > 0.0) || !isfinite(gg)) bad |= FDLP_ST_NONFINITE;

  if (live) {
    if (a.lpc_out) for (int i = g; i <= p; i += kLpcGroup) a.lpc_out[(size_t)task * (p + 1) + i] = s_a[i];
    if (a.gain_out && g == 0) a.gain_out[task] = gg;
  }
  __syncwarp();

  // ---- LPC -> cepstrum (features.py:233-246): x[0] = 1, x[i] = -a[i] (zero beyond the order) ----
  for (int i = 1 + g; i <= p; i += kLpcGroup) s_a[i] = -s_a[i];
  // the recursion only needs m*cep[m] (kept in s_mc); the plain values go straight to global memory
  double* msg = a.ms + (size_t)task * a.Lp;
  __syncwarp();
  const double c0 = log(sqrt(gg));
  const double c1 = (L > 1) ? s_a[1] : 0.0;
  if (g == 0) {
    s_mc[0] = 0.0;
    if (L > 1) s_mc[1] = c1;
    if (live) {
      msg[0] = c0 * a.weights[0];
      if (L > 1) msg[1] = c1 * a.weights[1];
      if (a.cep_out) { a.cep_out[(size_t)task * L] = c0; if (L > 1) a.cep_out[(size_t)task * L + 1] = c1; }
    }
  }
  __syncwarp();
  for (int n = 2; n < L; ++n) {
    double part = 0.0;
    for (int m = 1 + g; m < n; m += kLpcGroup) part = fma(s_a[n - m], s_mc[m], part);
    const double v = fma(group_sum8(part), a.inv_n[n], s_a[n]);
    if (g == 0) {
      s_mc[n] = v * (double)n;
      if (live) {
        msg[n] = v * a.weights[n];
        if (a.cep_out) a.cep_out[(size_t)task * L + n] = v;
      }
    }
    __syncwarp();
  }
  if (live && g == 0) for (int n = L; n < a.Lp; ++n) msg[n] = 0.0;

  if (bad && live && a.status && g == 0) {
    const int seg = task / a.J;
    atomicOr(a.status + (a.seg_utt ? a.seg_utt[seg] : seg), bad);
  }
}

// ---------------------------------------------------------------------------------------------------
// K4, register-resident variant (orders up to 8*NS - 2, L <= 8*NS): the shared-memory kernel above is bound by
// shared-memory bandwidth (every order reads and writes the whole coefficient vector and reads r).  Here the
// 8 lanes of a group hold everything in registers, element i in lane i%8, slot i/8, and the Levinson step is
// written with the forward vector f_i = a_i and the backward vector b_i = a_{m-i}:
//     sb = shift(b) (sb_i = b_{i-1});  f' = f + k sb;  b' = sb + k f;  acc_{m+1} = sum_j b'_j r[j+1]
// so nothing is ever reversed: the shift is one lane rotation per slot (lane 0 takes the previous slot) and
// r[j+1] sits at a FIXED element index, i.e. in registers too.  The LPC->cepstrum recursion uses the same
// rotation on y_m = x[n-m].  The slot loops are unrolled per phase (orders 8S..8S+7 touch slots 0..S).
template <int NS>
__global__ void __launch_bounds__(128, (NS > 8) ? 3 : 5) lpc_cepstrum_reg_kernel(const LpcArgs a) {
  const int lane = threadIdx.x & 31;
  const int g = lane & 7;
  const int gbase = lane & ~7;
  const int src_prev = gbase | ((g + 7) & 7);  // lane holding element i-1
  const int p = a.order, L = a.L;
  // grid-stride over blocks of 16 solves: a capped grid (one CTA per SM) lets this kernel share the SMs with
  // the DMMA-bound autocorrelation kernels of the next chunk when fdlp_forward pipelines chunks over two streams
  const int n_vblocks = (a.n_tasks + 15) >> 4;
  for (int vb = blockIdx.x; vb < n_vblocks; vb += gridDim.x) {
  int task = (vb * (int)blockDim.x + (int)threadIdx.x) >> 3;
  const bool live = task < a.n_tasks;
  if (!live) task = a.n_tasks - 1;  // idle groups shadow the last task (keeps the shuffles uniform)

  const double* rg = a.r + (size_t)task * a.nlags;
  double f[NS], b[NS], rr[NS];
#pragma unroll
  for (int s = 0; s < NS; ++s) {
    const int i = 8 * s + g;
    rr[s] = (i + 1 < a.nlags) ? __ldg(rg + i + 1) : 0.0;
    f[s] = 0.0;
    b[s] = 0.0;
  }
  const double r0 = __ldg(rg);
  if (g == 0) { f[0] = 1.0; b[0] = 1.0; }

  // ---- Levinson-Durbin ----
  double E = r0;
  int bad = 0;
  double acc = __shfl_sync(0xffffffffu, rr[0], gbase);  // r[1]
#pragma unroll
  for (int S = 0; S < NS; ++S) {
    const int m_lo = (8 * S > 1) ? 8 * S : 1;
    const int m_hi = (8 * S + 7 < p) ? 8 * S + 7 : p;
    for (int m = m_lo; m <= m_hi; ++m) {
      if (E == 0.0 || !isfinite(E)) bad = FDLP_ST_SINGULAR;
      const double k = -acc * __drcp_rn(E);
      double part = 0.0, part1 = 0.0;
      double t_hi = __shfl_sync(0xffffffffu, b[S], src_prev);
#pragma unroll
      for (int s = S; s >= 0; --s) {
        const double t_lo = (s > 0) ? __shfl_sync(0xffffffffu, b[s - 1], src_prev) : 0.0;
        const double sb = (g == 0) ? t_lo : t_hi;
        const double fo = f[s];
        f[s] = fma(k, sb, fo);
        const double nb = fma(k, fo, sb);
        b[s] = nb;
        if (s & 1) part1 = fma(nb, rr[s], part1);
        else part = fma(nb, rr[s], part);
        t_hi = t_lo;
      }
      acc = group_sum8(part + part1);
      E *= (1.0 - k * k);
    }
  }

  // ---- gain: gg = r[0] + sum_{i=0..p} a[i] * r[i+1]  (features.py:228) ----
  double gpart = 0.0;
#pragma unroll
  for (int s = 0; s < NS; ++s) gpart = fma(f[s], rr[s], gpart);
  const double gg = r0 + group_sum8(gpart);
  if (!(gg > 0.0) || !isfinite(gg)) bad |= FDLP_ST_NONFINITE;
  if (live) {
    if (a.lpc_out) {
#pragma unroll
      for (int s = 0; s < NS; ++s)
        if (8 * s + g <= p) a.lpc_out[(size_t)task * (p + 1) + 8 * s + g] = f[s];
    }
    if (a.gain_out && g == 0) a.gain_out[task] = gg;
  }

  // ---- LPC -> cepstrum (features.py:233-246): x[i] = -a[i] (i >= 1), zero beyond the order ----
  // mc[s]: m * cep[m] at element m (element 0 stays 0);  y[s]: y_m = x[n - m], rotated by one element per n
  double* msg = a.ms + (size_t)task * a.Lp;
#pragma unroll
  for (int s = 0; s < NS; ++s) { f[s] = -f[s]; b[s] = 0.0; rr[s] = 0.0; }  // f := x, b := y, rr := mc
  const double c0 = log(sqrt(gg));
  const double c1 = (L > 1) ? __shfl_sync(0xffffffffu, f[0], gbase | 1) : 0.0;  // x[1]
  if (g == 1) rr[0] = c1;  // 1 * cep[1]
  if (g == 0) b[0] = c1;   // y^(1): element 0 = x[1]
  if (g == 0 && live) {
    msg[0] = c0 * a.weights[0];
    if (L > 1) msg[1] = c1 * a.weights[1];
    if (a.cep_out) { a.cep_out[(size_t)task * L] = c0; if (L > 1) a.cep_out[(size_t)task * L + 1] = c1; }
  }
#pragma unroll
  for (int S = 0; S < NS; ++S) {
    const int n_lo = (8 * S > 2) ? 8 * S : 2;
    const int n_hi = (8 * S + 7 < L - 1) ? 8 * S + 7 : L - 1;
    for (int n = n_lo; n <= n_hi; ++n) {
      const double xn = (n <= p) ? __shfl_sync(0xffffffffu, f[S], gbase | (n & 7)) : 0.0;
      double part = 0.0, part1 = 0.0;
      double t_hi = __shfl_sync(0xffffffffu, b[S], src_prev);
#pragma unroll
      for (int s = S; s >= 0; --s) {
        const double t_lo = (s > 0) ? __shfl_sync(0xffffffffu, b[s - 1], src_prev) : 0.0;
        double y = (g == 0) ? t_lo : t_hi;
        if (s == 0 && g == 0) y = xn;  // element 0 of y^(n) is x[n]
        b[s] = y;
        if (s & 1) part1 = fma(rr[s], y, part1);
        else part = fma(rr[s], y, part);
        t_hi = t_lo;
      }
      const double v = fma(group_sum8(part + part1), __ldg(a.inv_n + n), xn);
      if (g == (n & 7)) rr[S] = v * (double)n;
      if (g == 0 && live) {
        msg[n] = v * __ldg(a.weights + n);
        if (a.cep_out) a.cep_out[(size_t)task * L + n] = v;
      }
    }
  }
  if (live && g == 0) for (int n = L; n < a.Lp; ++n) msg[n] = 0.0;

  if (bad && live && a.status && g == 0) {
    const int seg = task / a.J;
    atomicOr(a.status + (a.seg_utt ? a.seg_utt[seg] : seg), bad);
  }
  }  // grid-stride loop
}

// K4, blocked-cyclic register layout.  The rotation kernel above moves every register of the shifting vector
// through a shuffle at every order (2 SHFL + 2 FSEL per 3 DFMA: the LSU pipe and the issue slots, not the FP64
// pipe, set its time).  Here element i of a vector lives in group i / 40, lane (i % 40) / 5, register i % 5, so
// a shift by one element stays inside the lane for four registers out of five.  Those four do not move at
// all: the five registers of a (group, lane) form a circular buffer whose origin advances by one per order,
// and because the order loop is unrolled by five every index is a compile-time constant.  Only the register
// that leaves the lane goes through a shuffle: 2 SHFL + 2 FSEL per 15 DFMA.
// Physical register q of the shifting vector holds logical position (q + rho) mod 5 after rho shifts.
// 64-bit shuffles as two explicit 32-bit halves
__device__ __forceinline__ double shfl_idx_d(const double v, const int src) {
  const int lo = __shfl_sync(0xffffffffu, __double2loint(v), src);
  const int hi = __shfl_sync(0xffffffffu, __double2hiint(v), src);
  return __hiloint2double(hi, lo);
}
__device__ __forceinline__ double shfl_xor_d(const double v, const int mask) {
  const int lo = __shfl_xor_sync(0xffffffffu, __double2loint(v), mask);
  const int hi = __shfl_xor_sync(0xffffffffu, __double2hiint(v), mask);
  return __hiloint2double(hi, lo);
}
// sum over the LS (8 or 16) lanes of a solve, bit-identical in every lane (each level adds a symmetric pair)
template <int LS>
__device__ __forceinline__ double group_sum_d(double v) {
  if constexpr (LS == 16) v += shfl_xor_d(v, 8);
  v += shfl_xor_d(v, 4);
  v += shfl_xor_d(v, 2);
  v += shfl_xor_d(v, 1);
  return v;
}

// 1 / x for normal-range x: the MUFU seed and the two Newton steps of the library reciprocal without its
// special-case branch.  x == 0, denormal, infinite or NaN give a non-finite result, which the recursion
// carries to the end where it is reported as FDLP_ST_SINGULAR.
__device__ __forceinline__ double rcp_newton(const double x) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double e = fma(-x, y, 1.0);
  e = fma(e, e, e);
  y = fma(y, e, y);
  e = fma(-x, y, 1.0);
  return fma(y, e, y);
}

// One Levinson order m (T = m % 5, groups 0..PH hold elements 0..m):
//   f' = f + k shift(b),  b' = shift(b) + k f,  acc' = sum_j b'_j r[j+1]
// (Forming acc' as S1 + k S2 from lane-local sums that do not need k shortens the serial chain but costs five
// more DFMA per group and order; measured, the FP64 pipe then sets the time: 5.6 ms against 5.98 ms.)
template <int NG, int LS, int PH, int T>
__device__ __forceinline__ void lev_order(double (&f)[NG][5], double (&b)[NG][5], const double (&rr)[NG][5],
                                          double& acc, double& E, const bool g0, const int src_prev) {
  constexpr int C = 5;
  constexpr int qo = (C - T) % C;  // m - 1 shifts so far: logical 4 sits in register (4 - (m - 1)) mod 5
  double tin[PH + 1];
#pragma unroll
  for (int G = 0; G <= PH; ++G) tin[G] = shfl_idx_d(b[G][qo], src_prev);
#pragma unroll
  for (int G = PH; G >= 0; --G) {
    const double from_prev_group = (G > 0) ? tin[G > 0 ? G - 1 : 0] : 0.0;
    b[G][qo] = g0 ? from_prev_group : tin[G];
  }
  const double k = -acc * rcp_newton(E);
  E *= (1.0 - k * k);
  double pa[PH + 1], pb[PH + 1];  // two partial sums per group: 3 + 2 terms
#pragma unroll
  for (int G = 0; G <= PH; ++G) {
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int q = (c - T + C) % C;  // after the shift logical c sits in register (c - m) mod 5
      const double sb = b[G][q];
      const double fo = f[G][c];
      f[G][c] = fma(k, sb, fo);
      const double nb = fma(k, fo, sb);
      b[G][q] = nb;
      if (c == 0) pa[G] = nb * rr[G][c];
      else if (c == 1) pb[G] = nb * rr[G][c];
      else if (c & 1) pb[G] = fma(nb, rr[G][c], pb[G]);
      else pa[G] = fma(nb, rr[G][c], pa[G]);
    }
  }
  // balanced sum of the 2 (PH + 1) partials
  double t0 = pa[0] + pb[0];
  if constexpr (PH >= 1) { const double t1 = pa[1] + pb[1]; t0 += t1; }
  if constexpr (PH >= 2) {
    double t2 = pa[2] + pb[2];
    if constexpr (PH >= 3) { const double t3 = pa[3] + pb[3]; t2 += t3; }
    t0 += t2;
  }
  acc = group_sum_d<LS>(t0);
}

// One cepstrum step n (T = n % 5): y shifts by one with x[n] entering at element 0, v = x[n] + (1/n) sum mc y.
// x[n] is not needed after this step, so cep[n] = v takes its place in x: the step has no store and no
// divergent branch, and the kernel writes the whole cepstrum at the end.
// xg: group that holds element n of x (compile-time PH in the main loop, run-time in the tail).
// WRAP (n >= 40 NG, which is > order): the vectors are circular over the 40 NG slots -- y_m = x[n - m] is non-zero
// only for n - order <= m < n, so the 150 live values of y and of mc fit the ring; x[n] = 0 and the element that
// leaves the last slot re-enters at slot 0.
template <int NG, int LS, int PH, int T, bool WRAP = false>
__device__ __forceinline__ void cep_step(double (&x)[NG][5], double (&y)[NG][5], double (&mc)[NG][5], const int n,
                                         const int xg, const LpcArgs& a, const int g, const int gbase,
                                         const int src_prev) {
  constexpr int C = 5;
  double xv = x[PH][T];
#pragma unroll
  for (int G = 0; G < PH; ++G) xv = (G == xg) ? x[G][T] : xv;
  const int xl = (n / C) & (LS - 1);
  double xn = 0.0;
  if constexpr (!WRAP) {
    xn = shfl_idx_d(xv, gbase | xl);
    xn = (n <= a.order) ? xn : 0.0;
  }
  constexpr int qo = (1 - T + C) % C;  // n - 2 shifts so far
  double tin[PH + 1];
#pragma unroll
  for (int G = 0; G <= PH; ++G) tin[G] = shfl_idx_d(y[G][qo], src_prev);
#pragma unroll
  for (int G = PH; G >= 0; --G) {
    const double from_prev_group = (G > 0) ? tin[G > 0 ? G - 1 : 0] : (WRAP ? tin[PH] : xn);
    y[G][qo] = (g == 0) ? from_prev_group : tin[G];
  }
  double part[PH + 1];
#pragma unroll
  for (int G = 0; G <= PH; ++G) {
    part[G] = 0.0;
#pragma unroll
    for (int cc = 0; cc < C; ++cc) {  // element n - 1 (register (T - 1) mod 5), written by the previous step, goes last
      const int c = (T + cc) % C;
      part[G] = fma(mc[G][c], y[G][(c - T + 1 + C) % C], part[G]);
    }
  }
  double tot = part[0];
#pragma unroll
  for (int G = 1; G <= PH; ++G) tot += part[G];
  const double v = fma(group_sum_d<LS>(tot), __ldg(a.inv_n + n), xn);
  const double vn = v * (double)n;
#pragma unroll
  for (int G = 0; G <= PH; ++G) {
    const bool here = (g == xl) && (G == xg);
    mc[G][T] = here ? vn : mc[G][T];
    x[G][T] = here ? v : x[G][T];
  }
}

// blocks of five unguarded orders m = 5 qb + t for the blocks qb of phase PH (8 PH <= qb <= 8 PH + 7), then
// the next phase
template <int NG, int LS, int PH>
__device__ __forceinline__ void lev_phase(double (&f)[NG][5], double (&b)[NG][5], const double (&rr)[NG][5],
                                          double& acc, double& E, const bool g0, const int src_prev,
                                          const int qb_full) {
  const int qb_lo = (PH == 0) ? 1 : LS * PH;
  const int qb_hi = (LS * PH + LS - 1 < qb_full) ? LS * PH + LS - 1 : qb_full;
#pragma unroll 1
  for (int qb = qb_lo; qb <= qb_hi; ++qb) {
    lev_order<NG, LS, PH, 0>(f, b, rr, acc, E, g0, src_prev);
    lev_order<NG, LS, PH, 1>(f, b, rr, acc, E, g0, src_prev);
    lev_order<NG, LS, PH, 2>(f, b, rr, acc, E, g0, src_prev);
    lev_order<NG, LS, PH, 3>(f, b, rr, acc, E, g0, src_prev);
    lev_order<NG, LS, PH, 4>(f, b, rr, acc, E, g0, src_prev);
  }
  if constexpr (PH + 1 < NG) lev_phase<NG, LS, PH + 1>(f, b, rr, acc, E, g0, src_prev, qb_full);
}

// Five cepstrum steps n = n0 .. n0 + 4 (n0 a multiple of 5, >= 5).  The sum of step n splits into the terms
// m < n0, which need nothing from this block: the five shifts of y, the five dot products against the mc of
// earlier blocks and the five 8-lane sums are independent and pipeline; and the at most four terms m >= n0,
// (m cep[m]) x[n - m] with n - m <= 4, which every lane evaluates redundantly from x[1..4] (xs) in a short
// serial chain without any shuffle.  mc and cep of the block are written to their owner lane at the end.
template <int NG, int LS, int PH, bool WRAP = false>
__device__ __forceinline__ void cep_block(double (&x)[NG][5], double (&y)[NG][5], double (&mc)[NG][5], const int n0,
                                          const double (&xs)[4], const LpcArgs& a, const int g, const int gbase,
                                          const int src_prev) {
  constexpr int C = 5;
  constexpr int AG = WRAP ? NG - 1 : PH;  // last active group
  const int xl = (n0 / C) & (LS - 1);
  double S[C], xnv[C];
#pragma unroll
  for (int T = 0; T < C; ++T) {
    double xn = 0.0;
    if constexpr (!WRAP) {
      xn = shfl_idx_d(x[PH][T], gbase | xl);
      xn = (n0 + T <= a.order) ? xn : 0.0;
    }
    xnv[T] = xn;
    const int qo = (1 - T + C) % C;  // n - 2 shifts so far
    double tin[AG + 1];
#pragma unroll
    for (int G = 0; G <= AG; ++G) tin[G] = shfl_idx_d(y[G][qo], src_prev);
#pragma unroll
    for (int G = AG; G >= 0; --G) {
      const double from_prev_group = (G > 0) ? tin[G > 0 ? G - 1 : 0] : (WRAP ? tin[AG] : xn);
      y[G][qo] = (g == 0) ? from_prev_group : tin[G];
    }
    double pa[AG + 1], pb[AG + 1];
#pragma unroll
    for (int G = 0; G <= AG; ++G) {
      pa[G] = mc[G][0] * y[G][(0 - T + 1 + C) % C];
      pb[G] = mc[G][1] * y[G][(1 - T + 1 + C) % C];
      pa[G] = fma(mc[G][2], y[G][(2 - T + 1 + C) % C], pa[G]);
      pb[G] = fma(mc[G][3], y[G][(3 - T + 1 + C) % C], pb[G]);
      pa[G] = fma(mc[G][4], y[G][(4 - T + 1 + C) % C], pa[G]);
    }
    double tot = pa[0] + pb[0];
#pragma unroll
    for (int G = 1; G <= AG; ++G) tot += pa[G] + pb[G];
    S[T] = group_sum_d<LS>(tot);
    // circular: slot n0 + T still holds mc[n0 + T - 40 NG].  The later steps of this block meet it against
    // y = x[1..4] (the ring aliases x[d] and x[d + 40 NG]); that product belongs to the in-block terms below, so
    // the slot is cleared once step T has passed it (steps before T need it: it pairs with x[40 NG - (T - T')]).
    if constexpr (WRAP) mc[PH][T] = (g == xl) ? 0.0 : mc[PH][T];
  }
  double v[C], vn[C];
#pragma unroll
  for (int T = 0; T < C; ++T) {
    double fresh = S[T];
#pragma unroll
    for (int d = T; d >= 1; --d) fresh = fma(vn[T - d], xs[d - 1], fresh);  // newest term (d = 1) last
    v[T] = fma(fresh, __ldg(a.inv_n + n0 + T), xnv[T]);
    vn[T] = v[T] * (double)(n0 + T);
  }
  const bool here = g == xl;
#pragma unroll
  for (int T = 0; T < C; ++T) {
    mc[PH][T] = here ? vn[T] : mc[PH][T];
    x[PH][T] = here ? v[T] : x[PH][T];
  }
}

template <int NG, int LS, int PH>
__device__ __forceinline__ void cep_phase(double (&x)[NG][5], double (&y)[NG][5], double (&mc)[NG][5],
                                          const int nb_full, const double (&xs)[4], const LpcArgs& a, const int g,
                                          const int gbase, const int src_prev) {
  const int qb_lo = (PH == 0) ? 1 : LS * PH;
  const int qb_hi = (LS * PH + LS - 1 < nb_full) ? LS * PH + LS - 1 : nb_full;
#pragma unroll 1
  for (int qb = qb_lo; qb <= qb_hi; ++qb) cep_block<NG, LS, PH>(x, y, mc, 5 * qb, xs, a, g, gbase, src_prev);
  if constexpr (PH + 1 < NG) cep_phase<NG, LS, PH + 1>(x, y, mc, nb_full, xs, a, g, gbase, src_prev);
}

// one lap of the circular phase: blocks n0 = nbase + 5 qb', qb' = 8 PH .. 8 PH + 7, while the block ends before n_end
template <int NG, int LS, int PH>
__device__ __forceinline__ void cep_lap_phase(double (&x)[NG][5], double (&y)[NG][5], double (&mc)[NG][5],
                                              const int nbase, const int n_end, const double (&xs)[4],
                                              const LpcArgs& a, const int g, const int gbase, const int src_prev) {
#pragma unroll 1
  for (int qb = LS * PH; qb <= LS * PH + LS - 1; ++qb) {
    const int n0 = nbase + 5 * qb;
    if (n0 + 4 >= n_end) break;
    cep_block<NG, LS, PH, true>(x, y, mc, n0, xs, a, g, gbase, src_prev);
  }
  if constexpr (PH + 1 < NG) cep_lap_phase<NG, LS, PH + 1>(x, y, mc, nbase, n_end, xs, a, g, gbase, src_prev);
}

// cep[nbase + i], i = element index, for the elements below n_end: weighted copy for K5, plain copy on request
template <int NG, int LS>
__device__ __forceinline__ void cep_flush(const double (&x)[NG][5], const int nbase, const int n_end, const LpcArgs& a,
                                          double* msg, const int task, const int g) {
#pragma unroll
  for (int G = 0; G < NG; ++G)
#pragma unroll
    for (int c = 0; c < 5; ++c) {
      const int n = nbase + 5 * LS * G + 5 * g + c;
      if (n < n_end) {
        msg[n] = x[G][c] * __ldg(a.weights + n);
        if (a.cep_out) a.cep_out[(size_t)task * a.L + n] = x[G][c];
      }
    }
}

template <int NG, int LS>
__global__ void __launch_bounds__(128, (NG > 2) ? 3 : 5) lpc_cepstrum_blk_kernel(const LpcArgs a) {
  constexpr int C = 5, GE = LS * C;   // elements per group
  constexpr int kLog = (LS == 16) ? 4 : 3;
  constexpr int kTasks = 128 / LS;    // solves per CTA
  const int lane = threadIdx.x & 31;
  const int g = lane & (LS - 1);
  const bool g0 = g == 0;
  const int gbase = lane & ~(LS - 1);
  const int src_prev = gbase | ((g + LS - 1) & (LS - 1));
  const int p = a.order, L = a.L;
  const int n_vblocks = (a.n_tasks + kTasks - 1) / kTasks;
  for (int vb = blockIdx.x; vb < n_vblocks; vb += gridDim.x) {
  int task = (vb * (int)blockDim.x + (int)threadIdx.x) >> kLog;
  const bool live = task < a.n_tasks;
  if (!live) task = a.n_tasks - 1;  // idle groups shadow the last task (keeps the shuffles uniform)

  const double* rg = a.r + (size_t)task * a.nlags;
  double f[NG][C], b[NG][C], rr[NG][C];
#pragma unroll
  for (int G = 0; G < NG; ++G)
#pragma unroll
    for (int c = 0; c < C; ++c) {
      const int i = GE * G + C * g + c;
      rr[G][c] = (i + 1 < a.nlags) ? __ldg(rg + i + 1) : 0.0;
      f[G][c] = 0.0;
      b[G][c] = 0.0;
    }
  const double r0 = __ldg(rg);
  if (g0) { f[0][0] = 1.0; b[0][0] = 1.0; }

  // ---- Levinson-Durbin ----
  // orders 1..4 peeled, then unguarded blocks of five orders m = 5 qb + t (no branch between the orders, so the
  // registers stay where they are), then the p % 5 leftover orders
  double E = r0;
  int bad = 0;
  double acc = __shfl_sync(0xffffffffu, rr[0][0], gbase);  // r[1]
  if (p >= 1) lev_order<NG, LS, 0, 1>(f, b, rr, acc, E, g0, src_prev);
  if (p >= 2) lev_order<NG, LS, 0, 2>(f, b, rr, acc, E, g0, src_prev);
  if (p >= 3) lev_order<NG, LS, 0, 3>(f, b, rr, acc, E, g0, src_prev);
  if (p >= 4) lev_order<NG, LS, 0, 4>(f, b, rr, acc, E, g0, src_prev);
  const int qb_full = (p >= 4) ? (p - 4) / C : 0;  // last block whose five orders are all <= p
  lev_phase<NG, LS, 0>(f, b, rr, acc, E, g0, src_prev, qb_full);
  {
    const int m0 = C * (qb_full + 1);
    if (p >= 4 && m0 <= p) lev_order<NG, LS, NG - 1, 0>(f, b, rr, acc, E, g0, src_prev);
    if (p >= 4 && m0 + 1 <= p) lev_order<NG, LS, NG - 1, 1>(f, b, rr, acc, E, g0, src_prev);
    if (p >= 4 && m0 + 2 <= p) lev_order<NG, LS, NG - 1, 2>(f, b, rr, acc, E, g0, src_prev);
    if (p >= 4 && m0 + 3 <= p) lev_order<NG, LS, NG - 1, 3>(f, b, rr, acc, E, g0, src_prev);
  }

  // a zero or non-finite prediction error at any order turns k, and with it E, into NaN for good
  if (!isfinite(E) || r0 == 0.0) bad = FDLP_ST_SINGULAR;

  // ---- gain: gg = r[0] + sum_{i=0..p} a[i] * r[i+1]  (features.py:228) ----
  double gpart = 0.0;
#pragma unroll
  for (int G = 0; G < NG; ++G)
#pragma unroll
    for (int c = 0; c < C; ++c) gpart = fma(f[G][c], rr[G][c], gpart);
  const double gg = r0 + group_sum_d<LS>(gpart);
  if (!(gg > 0.0) || !isfinite(gg)) bad |= FDLP_ST_NONFINITE;
  if (live) {
    if (a.lpc_out) {
#pragma unroll
      for (int G = 0; G < NG; ++G)
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const int i = GE * G + C * g + c;
          if (i <= p) a.lpc_out[(size_t)task * (p + 1) + i] = f[G][c];
        }
    }
    if (a.gain_out && g0) a.gain_out[task] = gg;
  }

  // ---- LPC -> cepstrum (features.py:233-246): x[i] = -a[i] (i >= 1), zero beyond the order ----
  // rr := mc (m * cep[m] at element m, element 0 stays 0);  b := y, y^(n)_m = x[n - m]
  double* msg = a.ms + (size_t)task * a.Lp;
#pragma unroll
  for (int G = 0; G < NG; ++G)
#pragma unroll
    for (int c = 0; c < C; ++c) { f[G][c] = -f[G][c]; b[G][c] = 0.0; rr[G][c] = 0.0; }
  const double c0 = log(sqrt(gg));
  const double c1 = (L > 1) ? __shfl_sync(0xffffffffu, f[0][1], gbase) : 0.0;  // x[1]
  double xs[4];  // x[1..4] in every lane (the in-block terms of cep_block)
#pragma unroll
  for (int d = 1; d <= 4; ++d) xs[d - 1] = (d <= p) ? shfl_idx_d(f[0][d], gbase) : 0.0;
  if (g0) { rr[0][1] = c1; b[0][0] = c1; f[0][0] = c0; f[0][1] = c1; }  // 1 * cep[1];  y^(1)_0 = x[1];  cep[0], cep[1]
  // lap 0: steps n = 2 .. L0 - 1 with the vectors growing from slot 0
  const int L0 = (L < GE * NG) ? L : GE * NG;
  if (L0 > 2) cep_step<NG, LS, 0, 2>(f, b, rr, 2, 0, a, g, gbase, src_prev);
  if (L0 > 3) cep_step<NG, LS, 0, 3>(f, b, rr, 3, 0, a, g, gbase, src_prev);
  if (L0 > 4) cep_step<NG, LS, 0, 4>(f, b, rr, 4, 0, a, g, gbase, src_prev);
  const int nb_full = (L0 >= 5) ? (L0 - 5) / C : 0;  // last block n = 5 qb .. 5 qb + 4 with 5 qb + 4 <= L0 - 1
  cep_phase<NG, LS, 0>(f, b, rr, nb_full, xs, a, g, gbase, src_prev);
  {
    const int n0 = C * (nb_full + 1), xg = (nb_full + 1) >> kLog;
    if (L0 >= 5 && n0 <= L0 - 1) cep_step<NG, LS, NG - 1, 0>(f, b, rr, n0, xg, a, g, gbase, src_prev);
    if (L0 >= 5 && n0 + 1 <= L0 - 1) cep_step<NG, LS, NG - 1, 1>(f, b, rr, n0 + 1, xg, a, g, gbase, src_prev);
    if (L0 >= 5 && n0 + 2 <= L0 - 1) cep_step<NG, LS, NG - 1, 2>(f, b, rr, n0 + 2, xg, a, g, gbase, src_prev);
    if (L0 >= 5 && n0 + 3 <= L0 - 1) cep_step<NG, LS, NG - 1, 3>(f, b, rr, n0 + 3, xg, a, g, gbase, src_prev);
  }
  // f now holds the cepstrum at elements 0 .. L0-1
  if (live) cep_flush<NG, LS>(f, 0, L0, a, msg, task, g);
  // further laps (coeff_num > 40 NG, e.g. the REVERB recipe's 300 coefficients): circular vectors
  for (int nbase = GE * NG; nbase < L; nbase += GE * NG) {
    const int n_end = (L < nbase + GE * NG) ? L : nbase + GE * NG;
    cep_lap_phase<NG, LS, 0>(f, b, rr, nbase, n_end, xs, a, g, gbase, src_prev);
    const int nblk = (n_end - nbase) / C;  // full blocks done above
    const int n0 = nbase + C * nblk, xg = nblk >> kLog;
    if (n0 < n_end) cep_step<NG, LS, NG - 1, 0, true>(f, b, rr, n0, xg, a, g, gbase, src_prev);
    if (n0 + 1 < n_end) cep_step<NG, LS, NG - 1, 1, true>(f, b, rr, n0 + 1, xg, a, g, gbase, src_prev);
    if (n0 + 2 < n_end) cep_step<NG, LS, NG - 1, 2, true>(f, b, rr, n0 + 2, xg, a, g, gbase, src_prev);
    if (n0 + 3 < n_end) cep_step<NG, LS, NG - 1, 3, true>(f, b, rr, n0 + 3, xg, a, g, gbase, src_prev);
    if (live) cep_flush<NG, LS>(f, nbase, n_end, a, msg, task, g);
  }
  if (live && g0) for (int n = L; n < a.Lp; ++n) msg[n] = 0.0;

  if (bad && live && a.status && g0) {
    const int seg = task / a.J;
    atomicOr(a.status + (a.seg_utt ? a.seg_utt[seg] : seg), bad);
  }
  }  // grid-stride loop
}

// ---------------------------------------------------------------------------------------------------
// K5: env[task, t] = exp( sum_n ms[task, n] * cos(2 pi n t / M) ) * wratio[t],  t < kk
// One warp owns 8 tasks (the DMMA M dimension) and NT blocks of 8 output frames; the cosine basis
// B[n][t] is a plan table ([Lp, ldb], zero rows beyond L) read through L1/L2.
struct EnvArgs {
  const double* ms;      // [n_tasks_padded, Lp]
  const double* basis;   // [Lp, ldb]
  const double* wratio;  // [kk]
  void* env;             // [n_tasks, kk], float64 (stage API) or float32 (hand-over to the overlap-add kernel)
  int n_tasks, Lp, ldb, kk;
  // float32 hand-over only: a log-envelope outside the float32 range of exp() raises FDLP_ST_ENV_RANGE for the utterance
  int32_t* status;         // [n_utt] or nullptr
  const int32_t* seg_utt;  // [n_seg]
  int J;
};

constexpr int kEnvWarps = 4;

constexpr int kEnvMA = 2;  // 8-task row blocks per warp: every basis fragment feeds kEnvMA DMMAs

template <int NT, typename ET>
__global__ void __launch_bounds__(kEnvWarps * 32) envelope_dmma_kernel(const EnvArgs a) {
  ET* __restrict__ env = reinterpret_cast<ET*>(a.env);
  constexpr int MA = kEnvMA;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lr = lane & 3, lc = lane >> 2;
  const int task0 = (blockIdx.x * kEnvWarps + warp) * 8 * MA;
  if (task0 >= a.n_tasks) return;
  const int t_base = blockIdx.y * NT * 8;
  double acc[MA][NT][2];
#pragma unroll
  for (int m = 0; m < MA; ++m)
#pragma unroll
    for (int i = 0; i < NT; ++i) acc[m][i][0] = acc[m][i][1] = 0.0;
  // A[row = lane/4][col = lane%4] = ms[task0 + 8 m + lane/4][k0 + lane%4]   (rows beyond n_tasks: clamped, discarded)
  const double* pa[MA];
#pragma unroll
  for (int m = 0; m < MA; ++m) pa[m] = a.ms + (size_t)min(task0 + 8 * m + lc, a.n_tasks - 1) * a.Lp + lr;
  // B[row = lane%4][col = lane/4] = basis[k0 + lane%4][t0 + lane/4]
  const double* pb = a.basis + (size_t)lr * a.ldb + t_base + lc;
  // software pipeline: the A fragments and the NT basis fragments of step k0+4 are loaded while step k0 computes
  double af[MA];
#pragma unroll
  for (int m = 0; m < MA; ++m) af[m] = __ldg(pa[m]);
  double bf[NT];
#pragma unroll
  for (int i = 0; i < NT; ++i) bf[i] = __ldg(pb + 8 * i);
  for (int k0 = 0; k0 < a.Lp; k0 += 4) {
    const int kn = (k0 + 4 < a.Lp) ? k0 + 4 : k0;
    double afn[MA];
#pragma unroll
    for (int m = 0; m < MA; ++m) afn[m] = __ldg(pa[m] + kn);
    const double* pbk = pb + (size_t)kn * a.ldb;
    double bn[NT];
#pragma unroll
    for (int i = 0; i < NT; ++i) bn[i] = __ldg(pbk + 8 * i);
#pragma unroll
    for (int i = 0; i < NT; ++i)
#pragma unroll
      for (int m = 0; m < MA; ++m) dmma884(acc[m][i][0], acc[m][i][1], af[m], bf[i]);
#pragma unroll
    for (int m = 0; m < MA; ++m) af[m] = afn[m];
#pragma unroll
    for (int i = 0; i < NT; ++i) bf[i] = bn[i];
  }
  // C[row = lane/4][col = 2*(lane%4) + {0,1}]
#pragma unroll
  for (int m = 0; m < MA; ++m) {
    const int task = task0 + 8 * m + lc;
    if (task < a.n_tasks) {
      double worst = 0.0;  // largest |log envelope| of this lane's values (NaN compares false: reported by K4 already)
#pragma unroll
      for (int i = 0; i < NT; ++i) {
        const int t = t_base + 8 * i + 2 * lr;
        if (t < a.kk) env[(size_t)task * a.kk + t] = (ET)(exp(acc[m][i][0]) * __ldg(a.wratio + t));
        if (t + 1 < a.kk) env[(size_t)task * a.kk + t + 1] = (ET)(exp(acc[m][i][1]) * __ldg(a.wratio + t + 1));
        if (sizeof(ET) == 4) {
          if (t < a.kk) worst = fmax(worst, fabs(acc[m][i][0]));
          if (t + 1 < a.kk) worst = fmax(worst, fabs(acc[m][i][1]));
        }
      }
      // exp() leaves the normal float32 range beyond +-87: the float64 value is fine, the hand-over is not
      if (sizeof(ET) == 4 && worst > 87.0 && a.status) atomicOr(a.status + (a.seg_utt ? a.seg_utt[task / a.J] : task / a.J), FDLP_ST_ENV_RANGE);
    }
  }
}

}  // namespace fdlp
