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

// ============================================================================================
// Family decomposition (used when the plan finds band-independent templates, e.g. cochlear filters).
//
// A cochlear filter is piecewise F_j[n] = w_j,g * T_g[n] with T_0 = 1 (pass band),
// T_1[n] = 10^(-bet * bark(n)) (upper skirt), T_2[n] = 10^(alp * bark(n)) (lower skirt) -- features.py:197-219.
// For a block of samples that lies, together with its p+1 look-ahead samples, inside ONE region of band j,
//     sum_{n in block} x_j[n] x_j[n+k] = w_j,g^2 * P_g[block][k],   P_g[block][k] = sum_{n in block} (T_g c)[n] (T_g c)[n+k]
// and P_g is shared by every band: it is computed once per segment (family_block_autocorr_kernel, the same
// DMMA scheme as above with accumulators flushed per block), combined per band (family_combine_kernel), and
// only the few blocks per band that straddle a region boundary (or wrap around N) go through the direct
// per-band kernel (autocorr_ranges_kernel).  The result is the exact autocorrelation of the sequence
// (w T_g c), a relative 1e-12 perturbation of F_j c, evaluated with the same FP64 products and sums.

// Diagonal fold of a warp's NA accumulator tiles into lags, through warp-private shared memory.
// C_i[row][col] (lane holds row = lane/4, cols 2*(lane%4)+{0,1}) carries lag 8i + col - row.  Elements are stored
// skewed: slot (i + (col >= row)), diagonal d = (col - row) & 7, position row ^ (d >> 1), so the 8 terms of lag
// 8*ab + d are the 8 doubles of slot ab+1, diagonal d (in some order).  The XOR on the store side and the
// per-lane rotation on the load side make both 2 wavefronts per instruction, the minimum for 32 x 8 bytes
// (searched exhaustively over swizzles; the plain layout costs 4 per store and 16 per load).
constexpr int kFoldStride = 64;  // doubles per slot
__host__ __device__ constexpr int fold_smem_doubles(int na) { return (na + 1) * kFoldStride; }

__device__ __forceinline__ void prefetch_l1(const void* ptr) { asm volatile("prefetch.global.L1 [%0];" ::"l"(ptr)); }
__device__ __forceinline__ void prefetch_l2(const void* ptr) { asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr)); }

// Lane-constant shared-memory addresses of the fold, computed once per kernel so that every store and load of
// a flush is [register + immediate].  Loads are 16-byte pairs; the pair order is rotated per lane
// ((u + (lane >> 1)) & 3), which spreads the 32 rows of an instruction over all banks (4 wavefronts for 512 B).
struct FoldCtx {
  double* s0;          // store address of acc[.][0] in slot 0
  double* s1;          // store address of acc[.][1] in slot 0
  const double2* q[4]; // load addresses of the 4 pairs of lag `lane`
};

__device__ __forceinline__ FoldCtx fold_ctx(double* my, int lane) {
  const int lr = lane & 3, lc = lane >> 2;
  const int c0 = 2 * lr, c1 = 2 * lr + 1;
  const int d0 = (c0 - lc) & 7, d1 = (c1 - lc) & 7;
  FoldCtx f;
  f.s0 = my + ((c0 >= lc) ? kFoldStride : 0) + (d0 << 3) + (lc ^ (d0 >> 1));
  f.s1 = my + ((c1 >= lc) ? kFoldStride : 0) + (d1 << 3) + (lc ^ (d1 >> 1));
  const double2* row = reinterpret_cast<const double2*>(my + kFoldStride + 8 * lane);  // lag k = lane: 8 (k + 8)
  const int rot = (lane >> 1) & 3;
#pragma unroll
  for (int u = 0; u < 4; ++u) f.q[u] = row + ((u + rot) & 3);
  return f;
}

template <int NA>
__device__ __forceinline__ void fold_store_tiles(const double (&acc)[NA][2], const FoldCtx& f) {
#pragma unroll
  for (int i = 0; i < NA; ++i) {
    f.s0[i * kFoldStride] = acc[i][0];
    f.s1[i * kFoldStride] = acc[i][1];
  }
}

// lag k = lane + 32 t
__device__ __forceinline__ double fold_lag(const FoldCtx& f, int t) {
  const double2 v0 = f.q[0][t * 128], v1 = f.q[1][t * 128], v2 = f.q[2][t * 128], v3 = f.q[3][t * 128];
  return ((v0.x + v0.y) + (v1.x + v1.y)) + ((v2.x + v2.y) + (v3.x + v3.y));
}

struct FamArgs {
  const double* dct;          // [n_seg, ld]
  const double* tmpl;         // [nfam, ld] templates, circular extension like fbank
  double* p;                  // [n_seg, nblk_total, nlags] block sums, blocks of all families numbered consecutively
  const int32_t* bend;        // [nblk_total] end step (exclusive, steps of 32 samples) of each block
  const int32_t* task_fam;    // [ntask] template of the task
  const int32_t* task_blk0;   // [ntask] first block
  const int32_t* task_blk1;   // [ntask] one past the last block
  const int32_t* task_step0;  // [ntask] first step (= start of the first block)
  int n_seg, ld, nlags, nblk_total, ntask;
  int lag_groups;             // MULTI: warps per (segment, task), each 8 (NA - 1) complete lags
  unsigned long long* next;   // work counter (zeroed by the launcher): warps draw (segment, task, lag group) items
};

// The warps of the autocorrelation kernels are persistent and draw their work items from a global counter: tasks
// differ in length (75 .. 225 steps), and with one item per warp a CTA kept its registers until its slowest warp
// was done (ncu, round 1: 9.8 of 12 resident warps active on average).
__device__ __forceinline__ unsigned long long next_item(unsigned long long* counter, int lane) {
  unsigned long long v = 0;
  if (lane == 0) v = atomicAdd(counter, 1ULL);
  return __shfl_sync(0xffffffffu, v, 0);
}

// Blocks have variable length: their boundaries are the steps where some band's filter enters or leaves a
// region of the template (plan analysis), so every band's range is a union of whole blocks and the direct
// remainder is only the look-ahead window around each region boundary.  A warp owns a run of consecutive blocks
// of one (segment, template); the accumulators are flushed whenever the step counter reaches a block end.
// MULTI (more than 8 (NA - 1) lags, e.g. LPC order 300): lag group h of a task is a warp of its own whose B ring
// starts NA - 1 rows ahead per group -- the groups overlap by one tile, so each holds both halves (super- and
// sub-diagonal) of its 8 (NA - 1) lags -- and whose A fragment is loaded separately, as in autocorr_dmma_kernel.
template <int NA, bool MULTI>
__global__ void __launch_bounds__(kAcWarps * 32, MULTI ? 2 : 3) family_block_autocorr_kernel(const FamArgs a) {
  constexpr int RS = NA + 4;
  constexpr int U = RS / 4;
  extern __shared__ __align__(16) double cs[];  // [kAcWarps][NA + 1][kFoldStride]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lg = MULTI ? a.lag_groups : 1;
  const unsigned long long n_items = (unsigned long long)a.n_seg * a.ntask * lg;
  const int lr = lane & 3, lc = lane >> 2;
  const int lane_off = (lr << 3) + lc;
  const FoldCtx fc = fold_ctx(cs + (size_t)warp * fold_smem_doubles(NA), lane);
  // warps are independent: no block-wide barrier below
  for (unsigned long long gw = next_item(a.next, lane); gw < n_items; gw = next_item(a.next, lane)) {
  const int h = MULTI ? (int)(gw % lg) : 0;
  const unsigned long long gt = MULTI ? gw / lg : gw;
  const int a_base = h * (NA - 1);
  const int lag0 = 8 * a_base;
  const int seg = (int)(gt / a.ntask);
  const int task = (int)(gt % a.ntask);  // tasks are sorted by first step: neighbouring items share rows of c
  const int fam = a.task_fam[task];
  const int blk1 = a.task_blk1[task];
  int blk = a.task_blk0[task];
  int step = a.task_step0[task];
  const int m_lo = step * 4;

  const double* __restrict__ pc = a.dct + (size_t)seg * a.ld + (size_t)(m_lo + a_base) * 8 + lane_off;
  // template 0 is the pass band (all ones): its fragments are read from the first rows of the template at every step,
  // which stay in L1, instead of streaming 192 KB of ones through it
  const int tstep = fam == 0 ? 0 : 32;
  const double* __restrict__ pf = a.tmpl + (size_t)fam * a.ld + (fam == 0 ? (size_t)0 : (size_t)(m_lo + a_base) * 8) + lane_off;
  const double* __restrict__ ac = a.dct + (size_t)seg * a.ld + (size_t)m_lo * 8 + lane_off;
  const double* __restrict__ af = a.tmpl + (size_t)fam * a.ld + (fam == 0 ? (size_t)0 : (size_t)m_lo * 8) + lane_off;
  double paf = 0.0, pac = 0.0;
  if (MULTI) { paf = __ldg(af); pac = __ldg(ac); }
  double ring[RS];
#pragma unroll
  for (int i = 0; i < NA; ++i) ring[i] = __ldg(pf + 8 * i) * __ldg(pc + 8 * i);
  pc += 8 * NA;
  pf += 8 * NA;
  double qf[4], qc[4];
#pragma unroll
  for (int t = 0; t < 4; ++t) { qf[t] = __ldg(pf + 8 * t); qc[t] = __ldg(pc + 8 * t); }
#pragma unroll
  for (int i = NA; i < RS; ++i) ring[i] = 0.0;

  // the rows of c come from HBM (the templates stay in L1/L2): pull them into L1 two groups of U steps ahead
  // (lanes 0..pf_lanes-1: rows of c; the next pf_lanes lanes: the same rows of a skirt template, which comes from L2)
  const int pf_lanes = (U * 32 * 8 + 127) / 128;
  const double* cpf = (lane < pf_lanes || fam == 0)
                          ? a.dct + (size_t)seg * a.ld + (size_t)(m_lo + a_base + NA + 4 + 8 * U) * 8 + lane * 16
                          : a.tmpl + (size_t)fam * a.ld + (size_t)(m_lo + a_base + NA + 4 + 8 * U) * 8 + (lane - pf_lanes) * 16;
  const int pf_active = fam == 0 ? pf_lanes : 2 * pf_lanes;
  int next_end = a.bend[blk];
  double acc[NA][2];
#pragma unroll
  for (int i = 0; i < NA; ++i) acc[i][0] = acc[i][1] = 0.0;
  while (blk < blk1) {
    if (lane < pf_active) prefetch_l1(cpf);
    cpf += U * 32;
#pragma unroll
    for (int s = 0; s < U; ++s) {
      if (blk < blk1) {
#pragma unroll
        for (int t = 0; t < 4; ++t) ring[(4 * s + NA + t) % RS] = qf[t] * qc[t];
        const double afrag = MULTI ? paf * pac : ring[(4 * s) % RS];
        pc += 32;
        pf += tstep;
#pragma unroll
        for (int t = 0; t < 4; ++t) { qf[t] = __ldg(pf + 8 * t); qc[t] = __ldg(pc + 8 * t); }
        if (MULTI) {
          ac += 32;
          af += tstep;
          paf = __ldg(af);
          pac = __ldg(ac);
        }
#pragma unroll
        for (int i = 0; i < NA; ++i) dmma884(acc[i][0], acc[i][1], afrag, ring[(4 * s + i) % RS]);
        ++step;
#ifdef FDLP_X_NOFLUSH  // timing experiment only (wrong results): the ceiling of the DMMA phase without per-block flushes
        if (step == next_end && blk + 1 < blk1) { ++blk; next_end = a.bend[blk]; } else
#endif
        if (step == next_end) {
          // flush: C tiles -> shared (warp-private), diagonal sums -> P[seg][blk][0..nlags)
          fold_store_tiles<NA>(acc, fc);
          __syncwarp();
          double* pout = a.p + ((size_t)seg * a.nblk_total + blk) * a.nlags + lag0 + lane;
          // all shared-memory loads of the fold are issued before the first sum (only the stores are predicated):
          // one load latency + one adder tree per flush instead of one per group of 32 lags
          double fv[NA / 4];
#pragma unroll
          for (int t = 0; t < NA / 4; ++t) fv[t] = fold_lag(fc, t);
#pragma unroll
          for (int t = 0; t < NA / 4; ++t)
            if ((!MULTI || lane + 32 * t < 8 * (NA - 1)) && lag0 + lane + 32 * t < a.nlags) pout[32 * t] = fv[t];
          __syncwarp();
#pragma unroll
          for (int i = 0; i < NA; ++i) acc[i][0] = acc[i][1] = 0.0;
          ++blk;
          if (blk < blk1) next_end = a.bend[blk];
        }
      }
    }
  }
  }  // work items
}

// ---- per-band combination of the family block sums ----
// Band j takes, from family g, the blocks [a, b) that lie inside one region of its filter.  For the lower
// skirt a = 0 for every band and for the upper skirt b is the common last block, so those sums are running
// (prefix / suffix) sums over the blocks, accumulated from the far, small end of the skirt towards the pass
// band -- pure additions in a fixed order, no differences of partial sums.  family_scan_kernel turns P into
// these running sums in place; the pass band is a short range of per-block values.  The per-band gather
// (one running-sum vector per skirt + the pass-band blocks) happens in the epilogue of autocorr_ranges_kernel.
constexpr int kCombRange = 0, kCombPrefix = 1, kCombSuffix = 2;
constexpr int kCombMaxFam = 3;

struct ScanArgs {
  double* p;                   // [n_seg, nblk_total, nlags], scanned in place for prefix / suffix families
  int32_t mode[kCombMaxFam];   // kComb* of the scanned families (indexed by scan slot)
  int32_t boff[kCombMaxFam];   // first block of the family
  int32_t nblk[kCombMaxFam];   // blocks of the family
  int n_scan, n_seg, nblk_total, nlags;
};

// grid (n_seg * n_scan), one thread per lag
__global__ void __launch_bounds__(640) family_scan_kernel(const ScanArgs a) {
  const int k = threadIdx.x;
  if (k >= a.nlags) return;
  const int seg = blockIdx.x / a.n_scan;
  const int which = blockIdx.x % a.n_scan;
  int mode = a.mode[0], boff = a.boff[0], nblk = a.nblk[0];
#pragma unroll
  for (int i = 1; i < kCombMaxFam; ++i)
    if (which == i) { mode = a.mode[i]; boff = a.boff[i]; nblk = a.nblk[i]; }
  double* pg = a.p + ((size_t)seg * a.nblk_total + boff) * a.nlags + k;
  const int end = nblk;
  const size_t st = a.nlags;
  double run = 0.0;
  if (mode == kCombPrefix) {
    for (int m0 = 0; m0 < nblk; m0 += 8) {
      double v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = (m0 + u < nblk) ? pg[(size_t)(m0 + u) * st] : 0.0;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        run += v[u];
        if (m0 + u < nblk) pg[(size_t)(m0 + u) * st] = run;
      }
    }
  } else {
    for (int m0 = end - 1; m0 >= 0; m0 -= 8) {
      double v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = (m0 - u >= 0) ? pg[(size_t)(m0 - u) * st] : 0.0;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        run += v[u];
        if (m0 - u >= 0) pg[(size_t)(m0 - u) * st] = run;
      }
    }
  }
}

// ---- direct per-band kernel restricted to row ranges (blocks that straddle a region boundary) ----
constexpr int kMaxRanges = 8;

struct AcRangeArgs {
  const double* dct;     // [n_seg, ld]
  const double* fbank;   // [J, ld]
  const int32_t* nrng;   // [J] number of row ranges of the band
  const int32_t* rng;    // [J, kMaxRanges, 2] (row_lo, row_hi), multiples of 4 rows
  double* r;             // [n_seg, J, nlags]  (written: direct remainder + family terms)
  int J, nlags, n_seg, ld;
  // family terms gathered in the epilogue
  const double* p;       // [n_seg, nblk_total, nlags] block sums (prefix / suffix families already scanned)
  const int32_t* ra;     // [nfam, J] first block of the band's range (family-local index)
  const int32_t* rb;     // [nfam, J] one past the last block
  const double* wsq;     // [J, nfam] squared template weights
  int32_t mode[kCombMaxFam];
  int32_t boff[kCombMaxFam];  // first block of each family in p
  int nfam, nblk_total;
  int lag_groups;             // MULTI: warps per (segment, band), see family_block_autocorr_kernel
  unsigned long long* next;   // work counter (zeroed by the launcher)
};

template <int NA, bool MULTI>
__global__ void __launch_bounds__(kAcWarps * 32, MULTI ? 2 : 3) autocorr_ranges_kernel(const AcRangeArgs a) {
  constexpr int RS = NA + 4;
  constexpr int U = RS / 4;
  extern __shared__ __align__(16) double cs[];  // [kAcWarps][NA + 1][kFoldStride]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int lg = MULTI ? a.lag_groups : 1;
  const unsigned long long n_items = (unsigned long long)a.n_seg * a.J * lg;
  const int lr = lane & 3, lc = lane >> 2;
  const int lane_off = (lr << 3) + lc;
  const FoldCtx fc = fold_ctx(cs + (size_t)warp * fold_smem_doubles(NA), lane);
  constexpr int NT = NA / 4;  // lags per lane: k = lag0 + lane + 32 t
  const int lines = (a.nlags * 8 + 127) / 128;

  // persistent warps, work items (segment, band, lag group) drawn from a counter; the NEXT item is drawn before the
  // current one is computed so that its rows of c (HBM) are on their way into L1 / L2 while this one computes
  unsigned long long gw = next_item(a.next, lane);
  while (gw < n_items) {
    const unsigned long long gw_next = next_item(a.next, lane);
    if (gw_next < n_items) {
      const unsigned long long gt = MULTI ? gw_next / lg : gw_next;
      const int nseg_ = (int)(gt / a.J), nband = (int)(gt % a.J);
      const int nnr = a.nrng[nband];
      const double* crow = a.dct + (size_t)nseg_ * a.ld;
      for (int q = 0; q < nnr; ++q) {
        const int m_lo = a.rng[(nband * kMaxRanges + q) * 2], m_hi = a.rng[(nband * kMaxRanges + q) * 2 + 1];
        const int nline = ((m_hi - m_lo + NA * lg + 4) * 64 + 127) / 128;  // rows of 64 bytes
        for (int i = lane; i < nline; i += 32) prefetch_l1(crow + (size_t)m_lo * 8 + i * 16);
      }
    }
    const int h = MULTI ? (int)(gw % lg) : 0;
    const unsigned long long gt = MULTI ? gw / lg : gw;
    const int seg = (int)(gt / a.J), band = (int)(gt % a.J);
    const int a_base = h * (NA - 1);
    const int lag0 = 8 * a_base;
    const int nr = a.nrng[band];
    // family terms of the epilogue: (first row, row count, weight) per family; their lines start towards L2 now
    int f_first[kCombMaxFam], f_cnt[kCombMaxFam];
    double f_w[kCombMaxFam];
#pragma unroll
    for (int g = 0; g < kCombMaxFam; ++g) {
      f_first[g] = 0; f_cnt[g] = 0; f_w[g] = 0.0;
      if (g < a.nfam) {
        const int b0 = a.ra[g * a.J + band], b1 = a.rb[g * a.J + band];
        if (b1 > b0) {
          f_first[g] = a.boff[g] + ((a.mode[g] == kCombPrefix) ? b1 - 1 : b0);
          f_cnt[g] = (a.mode[g] == kCombRange) ? b1 - b0 : 1;
          f_w[g] = a.wsq[(size_t)band * a.nfam + g];
          const double* pg = a.p + ((size_t)seg * a.nblk_total + f_first[g]) * a.nlags;
          for (int idx = lane; idx < f_cnt[g] * lines; idx += 32)
            prefetch_l2(pg + (size_t)(idx / lines) * a.nlags + (idx % lines) * 16);
        }
      }
    }

    double acc[NA][2];
#pragma unroll
    for (int i = 0; i < NA; ++i) acc[i][0] = acc[i][1] = 0.0;

    for (int q = 0; q < nr; ++q) {
      const int m_lo = a.rng[(band * kMaxRanges + q) * 2], m_hi = a.rng[(band * kMaxRanges + q) * 2 + 1];
      const int nsteps = (m_hi - m_lo) >> 2;
      const double* __restrict__ pc = a.dct + (size_t)seg * a.ld + (size_t)(m_lo + a_base) * 8 + lane_off;
      const double* __restrict__ pf = a.fbank + (size_t)band * a.ld + (size_t)(m_lo + a_base) * 8 + lane_off;
      const double* __restrict__ ac = a.dct + (size_t)seg * a.ld + (size_t)m_lo * 8 + lane_off;
      const double* __restrict__ af = a.fbank + (size_t)band * a.ld + (size_t)m_lo * 8 + lane_off;
      double paf = 0.0, pac = 0.0;
      if (MULTI) { paf = __ldg(af); pac = __ldg(ac); }
      double ring[RS];
#pragma unroll
      for (int i = 0; i < NA; ++i) ring[i] = __ldg(pf + 8 * i) * __ldg(pc + 8 * i);
      pc += 8 * NA;
      pf += 8 * NA;
      double qf[4], qc[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) { qf[t] = __ldg(pf + 8 * t); qc[t] = __ldg(pc + 8 * t); }
#pragma unroll
      for (int i = NA; i < RS; ++i) ring[i] = 0.0;
      for (int s0 = 0; s0 < nsteps; s0 += U) {
#pragma unroll
        for (int s = 0; s < U; ++s) {
          if (s0 + s < nsteps) {
#pragma unroll
            for (int t = 0; t < 4; ++t) ring[(4 * s + NA + t) % RS] = qf[t] * qc[t];
            const double afrag = MULTI ? paf * pac : ring[(4 * s) % RS];
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

    fold_store_tiles<NA>(acc, fc);
    __syncwarp();
    double* rout = a.r + ((size_t)seg * a.J + band) * a.nlags + lag0;
    const int nl = (MULTI && a.nlags - lag0 > 8 * (NA - 1)) ? 8 * (NA - 1) : a.nlags - lag0;  // lags of this warp
    double val[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) val[t] = fold_lag(fc, t);
    __syncwarp();  // the fold buffer is free for the next item
    // family terms: the rows of all families form one list, read kGatherRows rows at a time (all loads of a round are
    // independent, so a round costs one L2 latency) and added in list order: skirt sums first, then the pass-band blocks
    constexpr int kGatherRows = 8;
    const int c0 = f_cnt[0], c01 = c0 + f_cnt[1], total = c01 + f_cnt[2];
    const double* pbase = a.p + (size_t)seg * a.nblk_total * a.nlags + lag0 + lane;
    for (int i0 = 0; i0 < total; i0 += kGatherRows) {
      double v[kGatherRows][NT], w[kGatherRows];
#pragma unroll
      for (int u = 0; u < kGatherRows; ++u) {
        const int i = i0 + u;
        const bool ok = i < total;
        const int row = i < c0 ? f_first[0] + i : (i < c01 ? f_first[1] + (i - c0) : f_first[2] + (i - c01));
        w[u] = !ok ? 0.0 : (i < c0 ? f_w[0] : (i < c01 ? f_w[1] : f_w[2]));
        const double* rp = pbase + (size_t)row * a.nlags;
#pragma unroll
        for (int t = 0; t < NT; ++t) v[u][t] = (ok && lane + 32 * t < nl) ? rp[32 * t] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < kGatherRows; ++u)
#pragma unroll
        for (int t = 0; t < NT; ++t) val[t] = fma(w[u], v[u][t], val[t]);
    }
#pragma unroll
    for (int t = 0; t < NT; ++t)
      if (lane + 32 * t < nl) rout[lane + 32 * t] = val[t];
    gw = gw_next;
  }
}

}  // namespace fdlp
