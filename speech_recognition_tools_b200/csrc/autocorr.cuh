// K3: sub-band windowing fused with the circular autocorrelation of the windowed DCT coefficients.
//
// Reference: band_dct = fbank[j, 0:-1] * cos_trans[i, :]   (src/featgen/computeFDLPSpectrogram.py:190-191)
//            y = real(ifft(fft(x) * conj(fft(x))))          (src/featgen/features.py:223-225, computeLpcFast)
//            of which only lags 0..order+1 are consumed     (features.py:226-228).
//
// r[k] = sum_n x[n] x[(n+k) mod N] is evaluated in direct form on the FP64 tensor path (DMMA.8x8x4):
// view x as a matrix X[m][c] = x[8m + c] (N/8 rows of 8).  Then
//     C_a[c][c'] = sum_m X[m][c] * X[(m+a) mod rows][c']        (an 8 x 8 x (N/8) GEMM per block a)
// holds every product with lag 8a + (c' - c), so r[8a+d] is the sum of the d-th super-diagonal of C_a and
// the (d-8)-th sub-diagonal of C_{a+1}.  Both DMMA operands are the SAME fragment pattern
//     frag(r)[lane] = x[8 * (r + lane%4) + lane/4],
// a contiguous 256-byte read per warp, so the windowed coefficients x = F_j * c are formed on the fly from
// two coalesced global loads (no shared-memory staging) and kept in a register ring: a step of 4 rows
// issues NA DMMAs (one per lag block) and fetches only 4 new fragments.  One warp owns one
// (segment, band, lag-group); FP64 products and accumulation are mandatory.
#pragma once
#include "common.cuh"

namespace fdlp {

struct AcArgs {
  const double* dct;     // [n_seg, ld]  ld = N + pad: the first `pad` coefficients repeated after the N-th
  const double* fbank;   // [J, ld]      (circular extension, so fragment reads never wrap)
  const int32_t* row_lo; // [J] first 8-sample row (multiple of 4) with a non-zero tap
  const int32_t* row_hi; // [J] one past the last such row (multiple of 4)
  double* r;             // [n_seg, J, nlags]
  int N, J, nlags, rows, n_seg;
  int ld;                // row stride of dct / fbank in doubles
  int a_groups;          // lag groups per band (each NA blocks of 8 lags)
  int na_total;          // a_groups * NA
};

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

constexpr int kAcWarps = 4;

template <int NA, bool MULTI>
__global__ void __launch_bounds__(kAcWarps * 32, 3) autocorr_dmma_kernel(const AcArgs a) {
  constexpr int RS = NA + 4;  // register ring of B fragments
  constexpr int U = RS / 4;   // steps per unrolled group (ring indices become compile-time)
  extern __shared__ __align__(16) double cs[];  // [bands_per_cta][na_total][8][8]

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int a_groups = MULTI ? a.a_groups : 1;
  const int bands_per_cta = kAcWarps / a_groups;
  const int nbg = (a.J + bands_per_cta - 1) / bands_per_cta;
  const int seg = blockIdx.x / nbg;
  const int bl = warp / a_groups;  // band slot in this CTA
  const int band = (blockIdx.x % nbg) * bands_per_cta + bl;
  const int a_base = MULTI ? (warp % a_groups) * NA : 0;
  const int lr = lane & 3, lc = lane >> 2;

  double acc[NA][2];
#pragma unroll
  for (int i = 0; i < NA; ++i) acc[i][0] = acc[i][1] = 0.0;

  if (band < a.J) {
    const int m_lo = a.row_lo[band], m_hi = a.row_hi[band];
    const int nsteps = (m_hi - m_lo) >> 2;
    // frag(r)[lane] = x[8 r + 8 (lane%4) + lane/4]; rows past N/8 read the circular extension, so the
    // fragment stream is two pointers advancing by 32 doubles per step with immediate offsets per load.
    const int lane_off = (lr << 3) + lc;
    const double* __restrict__ pc = a.dct + (size_t)seg * a.ld + (size_t)(m_lo + a_base) * 8 + lane_off;
    const double* __restrict__ pf = a.fbank + (size_t)band * a.ld + (size_t)(m_lo + a_base) * 8 + lane_off;

    double ring[RS];
#pragma unroll
    for (int i = 0; i < NA; ++i) ring[i] = __ldg(pf + 8 * i) * __ldg(pc + 8 * i);
    pc += 8 * NA;
    pf += 8 * NA;
    // raw operands of the 4 fragments the NEXT step adds: loaded one step ahead, multiplied when consumed,
    // so the L2 latency hides behind a full group of NA DMMAs
    double qf[4], qc[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) { qf[t] = __ldg(pf + 8 * t); qc[t] = __ldg(pc + 8 * t); }
#pragma unroll
    for (int i = NA; i < RS; ++i) ring[i] = 0.0;
    const double* __restrict__ ac = a.dct + (size_t)seg * a.ld + (size_t)m_lo * 8 + lane_off;
    const double* __restrict__ af = a.fbank + (size_t)band * a.ld + (size_t)m_lo * 8 + lane_off;
    double paf = 0.0, pac = 0.0;
    if (MULTI) { paf = __ldg(af); pac = __ldg(ac); }

    for (int s0 = 0; s0 < nsteps; s0 += U) {
#pragma unroll
      for (int s = 0; s < U; ++s) {
        if (s0 + s < nsteps) {
#pragma unroll
          for (int t = 0; t < 4; ++t) ring[(4 * s + NA + t) % RS] = qf[t] * qc[t];
          double afrag;
          if (MULTI) afrag = paf * pac;
          else afrag = ring[(4 * s) % RS];
          pc += 32;
          pf += 32;
#pragma unroll
          for (int t = 0; t < 4; ++t) { qf[t] = __ldg(pf + 8 * t); qc[t] = __ldg(pc + 8 * t); }
          if (MULTI) {
            ac += 32;
            af += 32;
            paf = __ldg(af);
            pac = __ldg(ac);
          }
#pragma unroll
          for (int i = 0; i < NA; ++i) dmma884(acc[i][0], acc[i][1], afrag, ring[(4 * s + i) % RS]);
        }
      }
    }
  }

  // ---- C fragments -> shared, then diagonal sums -> lags ----
  {
    double* my = cs + ((size_t)bl * a.na_total + a_base) * 64;
#pragma unroll
    for (int i = 0; i < NA; ++i) {
      // lane holds C[row = lane/4][col = 2*(lane%4) + {0,1}]
      *reinterpret_cast<double2*>(my + i * 64 + lc * 8 + 2 * lr) = make_double2(acc[i][0], acc[i][1]);
    }
  }
  __syncthreads();
  if (band < a.J) {
    const double* mine = cs + (size_t)bl * a.na_total * 64;
    double* rout = a.r + ((size_t)seg * a.J + band) * a.nlags;
    const int sub = warp % a_groups;  // the a_groups warps of a band split the lags
    for (int k = sub * 32 + lane; k < a.nlags; k += 32 * a_groups) {
      const int ab = k >> 3, d = k & 7;
      double s = 0.0;
      const double* c0 = mine + ab * 64;
      for (int row = 0; row + d < 8; ++row) s += c0[row * 8 + row + d];
      if (d > 0) {
        const double* c1 = c0 + 64;
        for (int row = 8 - d; row < 8; ++row) s += c1[row * 8 + row + d - 8];
      }
      rout[k] = s;
    }
  }
}

}  // namespace fdlp
