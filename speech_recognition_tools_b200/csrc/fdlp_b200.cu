// libfdlp_b200.so -- C ABI (include/fdlp_b200.h) over the sm_100a kernels of the FDLP-spectrogram hot path.
// Host side: plan construction (filterbanks, windows, twiddles, cosine basis -- the one-time setup of the
// reference's getFeats, src/featgen/computeFDLPSpectrogram.py:43-118) and the stream-ordered launch sequence.
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <mutex>
#include <new>
#include <vector>

#include "autocorr.cuh"
#include "common.cuh"
#include "dct.cuh"
#include "lpc_env.cuh"
#include "mel.cuh"
#include "mel_mma.cuh"
#include "mfcc.cuh"
#include "mfcc_tc.cuh"
#include "modspec_cplx.cuh"
#include "noise.cuh"
#include "ola.cuh"
#include "reverb.cuh"
#include "wav_io.cuh"

namespace fdlp {
thread_local char g_err[512] = "";

// ---------------------------------------------------------------- numpy-compatible host helpers
static double np_round_half_even(double x) { return nearbyint(x); }  // default rounding mode = to nearest even

static std::vector<double> np_linspace(double start, double stop, int num) {
  std::vector<double> v(num);
  if (num == 1) { v[0] = start; return v; }
  const double step = (stop - start) / (num - 1);
  for (int i = 0; i < num; ++i) v[i] = start + i * step;
  v[num - 1] = stop;
  return v;
}

static std::vector<double> np_hamming(int M) {
  std::vector<double> w(M);
  if (M == 1) { w[0] = 1.0; return w; }
  for (int i = 0; i < M; ++i) { const double n = (1 - M) + 2.0 * i; w[i] = 0.54 + 0.46 * cos(M_PI * n / (M - 1)); }
  return w;
}
static std::vector<double> np_hanning(int M) {
  std::vector<double> w(M);
  if (M == 1) { w[0] = 1.0; return w; }
  for (int i = 0; i < M; ++i) { const double n = (1 - M) + 2.0 * i; w[i] = 0.5 + 0.5 * cos(M_PI * n / (M - 1)); }
  return w;
}

// createFbank, src/featgen/features.py:172-190
static std::vector<double> make_mel_fbank(int nfilters, int nfft, int srate, double warp) {
  const double mel_max = 2595 * log10(1 + (srate / warp) / 1400);
  std::vector<double> fw = np_linspace(0, mel_max, nfilters + 2);
  const int nbins = (int)floor(nfft / 2.0 + 1);
  std::vector<double> f((size_t)nfilters * nbins, 0.0);
  std::vector<double> bin(nfilters + 2);
  for (int i = 0; i < nfilters + 2; ++i) {
    const double hz = warp * (700 * (pow(10.0, fw[i] / 2595) - 1));
    bin[i] = floor((nfft + 1) * hz / srate);
  }
  for (int m = 1; m <= nfilters; ++m) {
    const int lo = (int)bin[m - 1], ce = (int)bin[m], hi = (int)bin[m + 1];
    for (int k = lo; k < ce && k < nbins; ++k) f[(size_t)(m - 1) * nbins + k] = (k - bin[m - 1]) / (bin[m] - bin[m - 1]);
    for (int k = ce; k < hi && k < nbins; ++k) f[(size_t)(m - 1) * nbins + k] = (bin[m + 1] - k) / (bin[m + 1] - bin[m]);
  }
  return f;
}

// createFbankCochlear + warp_func_bark, src/featgen/features.py:193-219
static std::vector<double> make_cochlear_fbank(int nfilters, int nfft, int srate, double om_w, double alp, int fixed,
                                               double bet, double warp) {
  const double f_max = srate / 2.0;
  auto bark = [&](double x) { return 6 * asinh((x / warp) / 600); };
  std::vector<double> cf = np_linspace(0, bark(f_max), nfilters);
  const int nbins = (int)floor(nfft / 2.0 + 1);
  std::vector<double> fl = np_linspace(0, f_max, nbins);
  std::vector<double> f((size_t)nfilters * nbins);
  for (int i = 0; i < nfilters; ++i) {
    const double fc = cf[i];
    const double a = (fixed == 1) ? alp : alp * exp(-0.1 * fc);
    for (int j = 0; j < nbins; ++j) {
      const double d = bark(fl[j]) - fc;
      double v;
      if (d <= -om_w / 2) v = pow(10.0, a * (d + om_w / 2));
      else if (d > -om_w / 2 && d < om_w / 2) v = 1.0;
      else v = pow(10.0, -bet * (d - om_w / 2));
      f[(size_t)i * nbins + j] = v;
    }
  }
  return f;
}

static int factor_radices(int n, int* radix, int* npass) {
  int e2 = 0, e3 = 0, e5 = 0;
  while (n % 2 == 0) { n /= 2; ++e2; }
  while (n % 3 == 0) { n /= 3; ++e3; }
  while (n % 5 == 0) { n /= 5; ++e5; }
  if (n != 1) return -1;
  int c = 0;
  for (int i = 0; i < e2 / 2; ++i) radix[c++] = 4;
  if (e2 % 2) radix[c++] = 2;
  for (int i = 0; i < e3; ++i) radix[c++] = 3;
  for (int i = 0; i < e5; ++i) radix[c++] = 5;
  if (c > kMaxRadixPasses) return -1;
  *npass = c;
  return 0;
}

// Orders the radices so that consecutive passes pair into in-register passes of radix R1 * R2 <= 16 (dct.cuh
// fft_pass2_s): repeatedly the largest remaining factor with the largest partner that keeps the product <= 16,
// fours first ((4,4) = 16).  fuse[p] = 1 marks pass p as fused with pass p + 1.  N/2 = 12000: (4,4) (5,3) (5,2) 5.
static void pair_radices(int* radix, int npass, int* fuse) {
  int cnt[6] = {0, 0, 0, 0, 0, 0};
  for (int i = 0; i < npass; ++i) ++cnt[radix[i]];
  int c = 0;
  for (int i = 0; i < kMaxRadixPasses; ++i) fuse[i] = 0;
  while (cnt[4] >= 2) { radix[c] = 4; radix[c + 1] = 4; fuse[c] = 1; c += 2; cnt[4] -= 2; }
  for (int f1 = 5; f1 >= 2; --f1)
    while (cnt[f1] > 0) {
      --cnt[f1];
      int f2 = 0;
      for (int q = f1; q >= 2; --q)
        if (cnt[q] > 0 && f1 * q <= 16 && !(f1 == 2 && q == 2) && !(f1 == 5 && q >= 4)) { f2 = q; break; }
      if (f2) { --cnt[f2]; radix[c] = f1; radix[c + 1] = f2; fuse[c] = 1; c += 2; }
      else radix[c++] = f1;
    }
}

struct FftTables {
  int n = 0, npass = 0;
  int radix[kMaxRadixPasses];
  int fuse[kMaxRadixPasses] = {};
  double2* d_tw = nullptr;
  int32_t* d_perm = nullptr;
  double2* d_unpackA = nullptr;  // [n+1] exp(-2 pi i k / (2n))
  // split per-pass twiddle tables (dct.cuh fft_pass_s)
  double2* d_tws = nullptr;
  int tws_n = 0;
  int tw_off[kMaxRadixPasses], tw_na[kMaxRadixPasses];
  unsigned magic[kMaxRadixPasses];
};

template <typename T>
static int upload(T** dptr, const std::vector<T>& h) {
  FDLP_CUDA(cudaMalloc((void**)dptr, h.size() * sizeof(T)));
  FDLP_CUDA(cudaMemcpy(*dptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  return 0;
}

// The dynamic shared-memory limit is an attribute of the KERNEL, not of a plan: several plans live in one process
// (different configurations), so the limit is only ever raised.
template <typename K>
static cudaError_t raise_smem_limit(K kern, size_t bytes) {
  cudaFuncAttributes fa;
  cudaError_t e = cudaFuncGetAttributes(&fa, kern);
  if (e != cudaSuccess) return e;
  if ((size_t)fa.maxDynamicSharedSizeBytes >= bytes) return cudaSuccess;
  return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

static int build_fft_tables(int n, FftTables* t) {
  t->n = n;
  if (factor_radices(n, t->radix, &t->npass) != 0)
    return set_err(FDLP_E_UNSUPPORTED, "FFT length %s%ld has a prime factor other than 2, 3, 5", "", n);
  static const bool no_fuse = getenv("FDLP_DCT_NOFUSE") != nullptr;  // A/B runs
  if (!no_fuse) pair_radices(t->radix, t->npass, t->fuse);
  std::vector<double2> tw(n), A(n + 1);
  for (int k = 0; k < n; ++k) {
    const long double ang = -2.0L * M_PIl * k / n;
    tw[k] = make_double2((double)cosl(ang), (double)sinl(ang));
  }
  for (int k = 0; k <= n; ++k) {
    const long double ang = -2.0L * M_PIl * k / (2.0L * n);
    A[k] = make_double2((double)cosl(ang), (double)sinl(ang));
  }
  std::vector<int32_t> perm(n);
  for (int k = 0; k < n; ++k) {
    int kk = k, pos = 0, L = n;
    for (int q = 0; q < t->npass; ++q) { const int r = t->radix[q]; L /= r; pos += (kk % r) * L; kk /= r; }
    perm[k] = pos;
  }
  std::vector<double2> tws;
  {
    int L = n;
    for (int q = 0; q < t->npass; ++q) {
      const int r = t->radix[q], sub = L / r;
      const int na = (sub + 63) / 64;
      t->tw_off[q] = (int)tws.size();
      t->tw_na[q] = na;
      t->magic[q] = sub > 1 ? (unsigned)((0x100000000ULL + (unsigned long long)sub - 1) / (unsigned long long)sub) : 0u;
      for (int a = 0; a < na; ++a) {
        const long double ang = -2.0L * M_PIl * (64.0L * a) / L;
        tws.push_back(make_double2((double)cosl(ang), (double)sinl(ang)));
      }
      for (int b = 0; b < 64; ++b) {
        const long double ang = -2.0L * M_PIl * b / L;
        tws.push_back(make_double2((double)cosl(ang), (double)sinl(ang)));
      }
      L = sub;
    }
  }
  t->tws_n = (int)tws.size();
  int rc;
  if ((rc = upload(&t->d_tw, tw))) return rc;
  if ((rc = upload(&t->d_tws, tws))) return rc;
  if ((rc = upload(&t->d_perm, perm))) return rc;
  if ((rc = upload(&t->d_unpackA, A))) return rc;
  return 0;
}
static void free_fft_tables(FftTables* t) {
  cudaFree(t->d_tw); cudaFree(t->d_perm); cudaFree(t->d_unpackA); cudaFree(t->d_tws);
  t->d_tw = nullptr; t->d_perm = nullptr; t->d_unpackA = nullptr; t->d_tws = nullptr;
}

}  // namespace fdlp

using namespace fdlp;

// ================================================================ FDLP plan
constexpr int kMaxChunks = 8;

struct fdlp_plan {
  fdlp_params prm;
  fdlp_geometry g;
  int device = 0;
  int na = 0, a_groups = 1;  // autocorrelation lag blocking
  int pad = 0, ld = 0;       // circular extension (doubles) and row stride of the DCT / filterbank buffers
  bool dct_fp64 = true;
  // envelope -> overlap-add hand-over in float32 (half the traffic of the largest intermediate; the output is float32
  // and a frame sums at most two envelope values, so this costs ~1e-7 in the log features); FDLP_ENV_F64=1 for A/B runs
  bool env_f32 = true;       // fdlp_params.env_f64 = 1 switches to the float64 hand-over
  int dct_threads = 512;     // threads per DCT CTA (chosen for the butterfly counts of the passes)
  bool dct_half = false;     // float64 DCT with two CTAs per segment (dct_half_kernel): N float64 exceed shared memory
  FftTables fft_half;        // plan of the N/4-point sub-transforms
  double2* d_tw1 = nullptr;  // split table of the first radix-2 stage
  int tw1_na = 0;
  size_t dct_smem = 0, ac_smem = 0, lpc_smem = 0, ola_smem = 0;
  std::vector<double> h_fbank;  // [J, N]
  FftTables fft;
  double* d_ham = nullptr;
  double2* d_postB = nullptr;
  double2* d_unp = nullptr;  // split unpack / post-twiddle tables
  int unp_na = 0;
  double* d_cep = nullptr;    // [tasks, L] cepstra of fdlp_modspec_forward (cepstrum_mode plans)
  size_t cap_cep = 0;
  double2* d_hamt = nullptr;  // split Hamming table
  int ham_na = 0;
  double* d_fbank = nullptr;
  int32_t *d_row_lo = nullptr, *d_row_hi = nullptr;
  double *d_weights = nullptr, *d_inv_n = nullptr, *d_basis = nullptr, *d_wratio = nullptr;
  int Lp = 0, ldb = 0, env_ny = 1, lpc_threads = 256;
  // family decomposition of the autocorrelation (autocorr.cuh): templates, per-band block classes, direct ranges
  bool fam_ok = false;       // the plan found templates and may use the family path
  int fam_na = 0, fam_groups = 1;  // its lag blocking: fam_groups warps of fam_na tiles, 8 (fam_na - 1) lags each
  int ac_mode = 0;           // 0: family path when available, 1: always the direct per-band kernel
  int nfam = 0, nblk = 0 /* blocks of all templates */, fam_ntask = 0;
  int32_t comb_mode[3] = {0, 0, 0}, comb_nband[3] = {0, 0, 0}, comb_boff[3] = {0, 0, 0}, comb_nblk[3] = {0, 0, 0};
  size_t fam_smem = 0;
  double *d_tmpl = nullptr, *d_wsq = nullptr, *d_p = nullptr;
  unsigned long long* d_next = nullptr;  // work counters of the persistent autocorrelation kernels
  int sms = 148;
  int32_t *d_bend = nullptr, *d_ra = nullptr, *d_rb = nullptr, *d_nrng = nullptr, *d_rng = nullptr;
  int32_t *d_task_fam = nullptr, *d_task_blk0 = nullptr, *d_task_blk1 = nullptr, *d_task_step0 = nullptr;
  int cap_seg_p = 0;
  cudaEvent_t ev_ac[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t ev_lpc = nullptr;  // between the LPC/cepstrum kernel and the envelope kernel
  // scratch (grow-only)
  int cap_seg = 0, cap_utt = 0;
  double *d_dct = nullptr, *d_r = nullptr, *d_env = nullptr, *d_alpha = nullptr, *d_ms = nullptr;
  double* d_cmvn_part = nullptr;  // per-CTA CMVN partial sums of the overlap-add kernel (ola.cuh), grow-only
  int64_t cap_cmvn_slots = 0;
  // optional per-stage timing (CUDA events on the launching stream)
  bool profiling = false;
  // two-stream pipeline: K4/K5 of chunk c run on `side` while K2/K3 of chunk c+1 run on the caller's stream
  // (measured on B200: K3 and K4 CTAs do not become co-resident -- 19.6 + 3.8 ms alone, 23.1 ms together -- so
  // the pipeline brings nothing yet and is off by default; fdlp_plan_set_overlap turns it on)
  bool overlap = false;
  int overlap_chunks = kMaxChunks;
  int overlap_cap = 148;  // CTAs of the LPC kernel when it shares the SMs with the next chunk's autocorrelation
  cudaStream_t side = nullptr;
  cudaEvent_t ev_start = nullptr, ev_side = nullptr, ev_chunk[kMaxChunks] = {};
  cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
};

static void plan_free(fdlp_plan* p) {
  if (!p) return;
  free_fft_tables(&p->fft);
  free_fft_tables(&p->fft_half);
  cudaFree(p->d_tw1);
  cudaFree(p->d_ham); cudaFree(p->d_postB); cudaFree(p->d_unp); cudaFree(p->d_hamt); cudaFree(p->d_cep); cudaFree(p->d_fbank); cudaFree(p->d_row_lo); cudaFree(p->d_row_hi);
  cudaFree(p->d_weights); cudaFree(p->d_inv_n); cudaFree(p->d_basis); cudaFree(p->d_wratio);
  cudaFree(p->d_dct); cudaFree(p->d_r); cudaFree(p->d_env); cudaFree(p->d_alpha); cudaFree(p->d_ms);
  cudaFree(p->d_cmvn_part);
  cudaFree(p->d_tmpl); cudaFree(p->d_wsq); cudaFree(p->d_p); cudaFree(p->d_next);
  cudaFree(p->d_bend); cudaFree(p->d_ra); cudaFree(p->d_rb); cudaFree(p->d_nrng); cudaFree(p->d_rng);
  cudaFree(p->d_task_fam); cudaFree(p->d_task_blk0); cudaFree(p->d_task_blk1); cudaFree(p->d_task_step0);
  for (int i = 0; i < 4; ++i) if (p->ev_ac[i]) cudaEventDestroy(p->ev_ac[i]);
  if (p->ev_lpc) cudaEventDestroy(p->ev_lpc);
  for (int i = 0; i < 5; ++i) if (p->ev[i]) cudaEventDestroy(p->ev[i]);
  for (int i = 0; i < kMaxChunks; ++i) if (p->ev_chunk[i]) cudaEventDestroy(p->ev_chunk[i]);
  if (p->ev_start) cudaEventDestroy(p->ev_start);
  if (p->ev_side) cudaEventDestroy(p->ev_side);
  if (p->side) cudaStreamDestroy(p->side);
  delete p;
}

static int ensure_scratch(fdlp_plan* p, int n_seg, int n_utt) {
  if (n_seg > p->cap_seg) {
    cudaFree(p->d_dct); cudaFree(p->d_r); cudaFree(p->d_env); cudaFree(p->d_ms);
    p->d_dct = p->d_r = p->d_env = p->d_ms = nullptr; p->cap_seg = 0;
    const size_t J = p->prm.nfilters;
    FDLP_CUDA(cudaMalloc((void**)&p->d_dct, (size_t)n_seg * p->ld * sizeof(double)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_r, (size_t)n_seg * J * p->g.nlags * sizeof(double)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_env, (size_t)n_seg * J * p->g.kk * sizeof(double)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_ms, ((size_t)n_seg * J + 16) * p->Lp * sizeof(double)));
    p->cap_seg = n_seg;
  }
  if (n_utt > p->cap_utt) {
    cudaFree(p->d_alpha); p->d_alpha = nullptr; p->cap_utt = 0;
    FDLP_CUDA(cudaMalloc((void**)&p->d_alpha, (size_t)n_utt * sizeof(double)));
    p->cap_utt = n_utt;
  }
  return 0;
}


// ---------------------------------------------------------------- family decomposition (host analysis)
// Finds, for every (band, block of blk_len samples), whether the filter restricted to the block and its
// look-ahead of 8*na samples is a constant multiple of one band-independent template.  The decision is made
// numerically on the filterbank the plan actually uses (built the reference's way), so a block that does not
// match simply stays on the direct path; the analytic form of createFbankCochlear (features.py:197-219) only
// supplies the candidate templates.
static int build_families(fdlp_plan* p, const std::vector<int32_t>& row_lo, const std::vector<int32_t>& row_hi) {
  const fdlp_params& prm = p->prm;
  const int N = p->g.N, J = prm.nfilters;
  p->fam_ok = false;
  if (prm.fbank_kind != FDLP_FBANK_COCHLEAR) return 0;
  // one lag group of p->na tiles when the general kernel needs one; else groups of 20 tiles overlapping by one
  p->fam_na = (p->a_groups == 1) ? p->na : 20;
  p->fam_groups = (p->a_groups == 1) ? 1 : (p->g.nlags + 8 * (p->fam_na - 1) - 1) / (8 * (p->fam_na - 1));
  if (p->fam_groups > 4) return 0;
  const double alp = prm.fbank_args[1], bet = prm.fbank_args[3], warp = prm.fbank_args[4];
  const int fixed = (int)prm.fbank_args[2];
  const int nfam = (fixed == 1) ? 3 : 2;
  const int ST = N / 32;       // steps of 32 samples (4 rows of 8): the granularity of the DMMA loops
  const int look = 8 * ((p->fam_groups - 1) * (p->fam_na - 1) + p->fam_na);  // lags the DMMA tiles evaluate
  const double tol = 1e-12;
  if (ST < 64) return 0;

  // templates (same expressions as make_cochlear_fbank)
  const double f_max = prm.srate / 2.0;
  std::vector<double> fl = np_linspace(0, f_max, N + 1);
  std::vector<double> tm((size_t)nfam * N);
  for (int n = 0; n < N; ++n) {
    const double b = 6 * asinh((fl[n] / warp) / 600);
    tm[n] = 1.0;
    tm[(size_t)N + n] = pow(10.0, -bet * b);
    if (nfam > 2) tm[(size_t)2 * N + n] = pow(10.0, alp * b);
  }

  // ---- per (band, template): the run of steps [ta, tb) whose samples, with their look-ahead, are w * T_g ----
  // (regardless of the tail_eps support: a skirt continues as the same template outside it, and the running
  // sums include those blocks for free)
  std::vector<int32_t> ta((size_t)nfam * J, 0), tb((size_t)nfam * J, 0);
  std::vector<double> wsq((size_t)J * nfam, 0.0);
  std::vector<int32_t> nrng(J, 0), rng((size_t)J * kMaxRanges * 2, 0);
  int64_t direct_taps = 0;
  std::vector<char> claimed(ST);
  for (int j = 0; j < J; ++j) {
    const double* F = &p->h_fbank[(size_t)j * N];
    std::fill(claimed.begin(), claimed.end(), 0);
    for (int g = 0; g < nfam; ++g) {
      const double* T = &tm[(size_t)g * N];
      double wref = 0.0;
      int first = -1, last = -1;
      for (int t = 0; t < ST; ++t) {
        const int s0 = 32 * t, e0 = s0 + 32;
        if (e0 + look > N) break;  // the look-ahead wraps around N: not a single region
        if (claimed[t]) continue;
        const double w = wref > 0 ? wref : F[s0] / T[s0];
        if (!(w > 0) || !std::isfinite(w)) continue;
        bool ok = true;
        for (int n = s0; n < e0 + look && ok; ++n) ok = fabs(F[n] - w * T[n]) <= tol * w * T[n];
        if (!ok) continue;
        if (first >= 0 && t != last + 1) return 0;  // not one contiguous run: keep the direct path
        if (first < 0) first = t;
        last = t;
        wref = w;
        claimed[t] = 1;
      }
      wsq[(size_t)j * nfam + g] = wref * wref;
      ta[(size_t)g * J + j] = first >= 0 ? first : 0;
      tb[(size_t)g * J + j] = first >= 0 ? last + 1 : 0;
    }
    // steps inside the support that no template covers: merged into row ranges for the direct kernel
    const int t_lo = row_lo[j] / 4, t_hi = row_hi[j] / 4;
    int nr = 0;
    for (int t = t_lo; t < t_hi;) {
      if (claimed[t]) { ++t; continue; }
      int t2 = t;
      while (t2 < t_hi && !claimed[t2]) ++t2;
      if (nr >= kMaxRanges) return 0;  // too fragmented: keep the direct path for the whole plan
      rng[((size_t)j * kMaxRanges + nr) * 2] = 4 * t;
      rng[((size_t)j * kMaxRanges + nr) * 2 + 1] = 4 * t2;
      direct_taps += 32 * (t2 - t);
      ++nr;
      t = t2;
    }
    nrng[j] = nr;
  }

  // ---- blocks: per template, the intervals between consecutive run boundaries of any band ----
  std::vector<int32_t> bend, ra((size_t)nfam * J, 0), rb((size_t)nfam * J, 0);
  int32_t mode[kCombMaxFam] = {0, 0, 0}, nband[kCombMaxFam] = {0, 0, 0}, boff[kCombMaxFam] = {0, 0, 0},
          nblk_g[kCombMaxFam] = {0, 0, 0};
  std::vector<int32_t> task_fam, task_blk0, task_blk1, task_step0;
  int64_t fam_taps = 0;
  int chunk_steps = 150;
  if (const char* e = getenv("FDLP_FAM_CHUNK")) chunk_steps = std::max(8, atoi(e));
  for (int g = 0; g < nfam; ++g) {
    std::vector<int> cuts;
    for (int j = 0; j < J; ++j)
      if (tb[(size_t)g * J + j] > ta[(size_t)g * J + j]) {
        cuts.push_back(ta[(size_t)g * J + j]);
        cuts.push_back(tb[(size_t)g * J + j]);
        ++nband[g];
      }
    std::sort(cuts.begin(), cuts.end());
    cuts.erase(std::unique(cuts.begin(), cuts.end()), cuts.end());
    boff[g] = (int32_t)bend.size();
    if (cuts.size() < 2) continue;
    nblk_g[g] = (int32_t)cuts.size() - 1;
    for (size_t i = 1; i < cuts.size(); ++i) bend.push_back(cuts[i]);
    fam_taps += 32LL * (cuts.back() - cuts.front());
    bool all_a0 = true, all_b_end = true;
    for (int j = 0; j < J; ++j) {
      const int a0 = ta[(size_t)g * J + j], b0 = tb[(size_t)g * J + j];
      if (b0 <= a0) continue;
      ra[(size_t)g * J + j] = (int32_t)(std::lower_bound(cuts.begin(), cuts.end(), a0) - cuts.begin());
      rb[(size_t)g * J + j] = (int32_t)(std::lower_bound(cuts.begin(), cuts.end(), b0) - cuts.begin());
      if (a0 != cuts.front()) all_a0 = false;
      if (b0 != cuts.back()) all_b_end = false;
    }
    mode[g] = all_a0 ? kCombPrefix : (all_b_end ? kCombSuffix : kCombRange);
    // warp tasks: runs of whole blocks of about chunk_steps steps
    int b0 = 0;
    while (b0 < nblk_g[g]) {
      int b1 = b0 + 1;
      while (b1 < nblk_g[g] && cuts[b1] - cuts[b0] < chunk_steps) ++b1;
      if (nblk_g[g] - b1 > 0 && cuts[nblk_g[g]] - cuts[b1] < chunk_steps / 2) b1 = nblk_g[g];  // no tiny last task
      task_fam.push_back(g);
      task_blk0.push_back(boff[g] + b0);
      task_blk1.push_back(boff[g] + b1);
      task_step0.push_back(cuts[b0]);
      b0 = b1;
    }
  }
  const int nblk_total = (int)bend.size();
  if (nblk_total == 0) return 0;
  // worthwhile only if the shared work plus the direct remainder is well below the per-band total
  if ((fam_taps + direct_taps) * 2 > p->g.support_taps) return 0;
  {  // neighbouring warps should share rows of c: order the tasks by their first step
    std::vector<int> idx(task_fam.size());
    for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
    std::stable_sort(idx.begin(), idx.end(), [&](int x, int y) { return task_step0[x] < task_step0[y]; });
    auto permute = [&](std::vector<int32_t>& v) { std::vector<int32_t> o(v.size()); for (size_t i = 0; i < idx.size(); ++i) o[i] = v[idx[i]]; v.swap(o); };
    permute(task_fam); permute(task_blk0); permute(task_blk1); permute(task_step0);
  }
  for (int g = 0; g < kCombMaxFam; ++g) {
    p->comb_mode[g] = mode[g]; p->comb_nband[g] = nband[g]; p->comb_boff[g] = boff[g]; p->comb_nblk[g] = nblk_g[g];
  }

  std::vector<double> tpad((size_t)nfam * p->ld);
  for (int g = 0; g < nfam; ++g)
    for (int k = 0; k < p->ld; ++k) tpad[(size_t)g * p->ld + k] = tm[(size_t)g * N + (k % N)];
  int rc;
  if ((rc = upload(&p->d_tmpl, tpad)) || (rc = upload(&p->d_wsq, wsq)) || (rc = upload(&p->d_bend, bend)) ||
      (rc = upload(&p->d_ra, ra)) || (rc = upload(&p->d_rb, rb)) || (rc = upload(&p->d_nrng, nrng)) ||
      (rc = upload(&p->d_rng, rng)) || (rc = upload(&p->d_task_fam, task_fam)) ||
      (rc = upload(&p->d_task_blk0, task_blk0)) || (rc = upload(&p->d_task_blk1, task_blk1)) ||
      (rc = upload(&p->d_task_step0, task_step0)))
    return rc;
  if (!p->d_next) FDLP_CUDA(cudaMalloc((void**)&p->d_next, 2 * sizeof(unsigned long long)));
  {
    cudaDeviceProp prop;
    FDLP_CUDA(cudaGetDeviceProperties(&prop, p->device));
    p->sms = prop.multiProcessorCount;
  }
  p->nfam = nfam; p->nblk = nblk_total; p->fam_ntask = (int)task_fam.size();
  p->fam_smem = (size_t)kAcWarps * fold_smem_doubles(p->fam_na) * sizeof(double);
  p->g.family_taps = fam_taps;
  p->g.direct_taps = direct_taps;
  p->g.nfam = nfam;
  p->g.block_len = (int32_t)(fam_taps / nblk_total);  // average block length in samples
  p->fam_ok = true;
  return 0;
}

extern "C" const char* fdlp_last_error(void) { return g_err; }
extern "C" int fdlp_version(void) { return FDLP_ABI_VERSION; }
extern "C" int64_t fdlp_sizeof_struct(int32_t which) {
  switch (which) {
    case 0: return sizeof(fdlp_params);
    case 1: return sizeof(fdlp_geometry);
    case 2: return sizeof(fdlp_batch);
    case 3: return sizeof(mel_params);
    case 4: return sizeof(mel_batch);
    case 5: return sizeof(fdlp_reverb_batch);
    case 6: return sizeof(mfcc_params);
    case 7: return sizeof(mfcc_batch);
    default: return -1;
  }
}
// every params / batch struct leads with the caller's sizeof(): a binding made from another header version is refused
#define FDLP_CHECK_SIZE(ptr, type)                                                                        \
  if ((ptr)->struct_size != sizeof(type))                                                                 \
    return set_err(FDLP_E_ARG, "%s.struct_size = %ld, this library expects %ld (include/fdlp_b200.h, ABI 200)",   \
                   #type, (long)(ptr)->struct_size, (long)sizeof(type))

extern "C" int fdlp_plan_create(const fdlp_params* prm, fdlp_plan** out) {
  if (!prm || !out) return set_err(FDLP_E_ARG, "null argument%s", "");
  *out = nullptr;
  FDLP_CHECK_SIZE(prm, fdlp_params);
  if (prm->nfilters < 1 || prm->order < 1 || prm->coeff_num < 2 || prm->frate < 1 || prm->srate < 1 ||
      !(prm->fduration > 0) || !(prm->overlap_fraction >= 0 && prm->overlap_fraction < 1) || !prm->weights)
    return set_err(FDLP_E_ARG, "invalid fdlp_params%s", "");
  fdlp_plan* p = new (std::nothrow) fdlp_plan();
  if (!p) return set_err(FDLP_E_NOMEM, "out of host memory%s", "");
  p->prm = *prm;
  p->env_f32 = prm->env_f64 == 0;
  cudaGetDevice(&p->device);


  // ---- geometry, with the reference's exact float expressions ----
  const double of = 1 - prm->overlap_fraction;               // computeFDLPSpectrogram.py:104
  const double lfr = 1 / (of * prm->fduration);              // :173
  fdlp_geometry& g = p->g;
  g.N = (int)(prm->srate * prm->fduration);                  // features.py:134
  g.S = (int)(prm->srate / lfr);                             // features.py:135
  if (prm->seg_shift > 0) g.S = prm->seg_shift;             // computeModulationSpectrum.py:143-144: int(srate / frate)
  g.M = 2 * (int)(prm->fduration * prm->frate);              // :201
  g.kk = (int)np_round_half_even(prm->fduration * prm->frate);       // :203
  g.kkb2 = (int)np_round_half_even(prm->fduration * prm->frate / 2); // :204
  g.kks = (int)np_round_half_even(prm->fduration * prm->frate * of); // :220
  g.L = prm->cepstrum_mode ? prm->coeff_num : std::min(prm->coeff_num, g.M);
  g.nlags = prm->order + 2;
  const int N = g.N, J = prm->nfilters;
  if (N < 64 || N % 32 != 0 || g.S < 1 || g.M < 2 || g.kk < 1 || g.kk > g.M) {
    plan_free(p);
    return set_err(FDLP_E_UNSUPPORTED, "segment length %s%ld must be a multiple of 32 (and kk <= M)", "", N);
  }
  // lag blocking for the DMMA autocorrelation: blocks of 8 lags, NA blocks per warp
  const int kmax = g.nlags - 1;
  const int na_total = (kmax + 7) / 8 + 1;
  int groups = 1;
  while ((na_total + groups - 1) / groups > 20) groups *= 2;
  if (groups > 4 || na_total + 8 > N / 8) {
    plan_free(p);
    return set_err(FDLP_E_UNSUPPORTED, "LPC order %s%ld too large for this segment length", "", prm->order);
  }
  p->a_groups = groups;
  p->na = (((na_total + groups - 1) / groups) + 3) & ~3;
  int ext_rows = p->na * groups;  // rows past N the fragment streams may read
  if (groups > 1) ext_rows = std::max(ext_rows, 19 * ((g.nlags + 151) / 152 - 1) + 20);  // family lag groups (build_families)
  p->pad = 8 * (ext_rows + 12);
  p->ld = N + p->pad;
  if (p->pad > N / 2) {
    plan_free(p);
    return set_err(FDLP_E_UNSUPPORTED, "LPC order %s%ld too large for this segment length", "", prm->order);
  }

  int rc = build_fft_tables(N / 2, &p->fft);
  if (rc) { plan_free(p); return rc; }

  // ---- tables ----
  {
    std::vector<double> ham = np_hamming(N);
    std::vector<double2> B(N);
    const long double sc = 2.0L / sqrtl(2.0L * N);
    for (int k = 0; k < N; ++k) {
      const long double ang = -M_PIl * k / (2.0L * N);
      B[k] = make_double2((double)(cosl(ang) * sc), (double)(sinl(ang) * sc));
    }
    if ((rc = upload(&p->d_ham, ham)) || (rc = upload(&p->d_postB, B))) { plan_free(p); return rc; }
    // split tables for the unpack step (dct.cuh): k = 64 a + b
    const int n = N / 2;
    p->unp_na = n / 64 + 1;
    std::vector<double2> U;
    for (int a = 0; a < p->unp_na; ++a) { const long double ang = -2.0L * M_PIl * (64.0L * a) / N; U.push_back(make_double2((double)cosl(ang), (double)sinl(ang))); }
    for (int b = 0; b < 64; ++b) { const long double ang = -2.0L * M_PIl * b / N; U.push_back(make_double2((double)cosl(ang), (double)sinl(ang))); }
    for (int a = 0; a < p->unp_na; ++a) { const long double ang = -M_PIl * (64.0L * a) / (2.0L * N); U.push_back(make_double2((double)cosl(ang), (double)sinl(ang))); }
    for (int b = 0; b < 64; ++b) { const long double ang = -M_PIl * b / (2.0L * N); U.push_back(make_double2((double)(cosl(ang) * sc), (double)(sinl(ang) * sc))); }
    if ((rc = upload(&p->d_unp, U))) { plan_free(p); return rc; }
    // split Hamming table (dct.cuh): i = 64 a + b
    p->ham_na = (N + 63) / 64 + 1;
    std::vector<double2> H;
    if (prm->window_kind < FDLP_WINDOW_HAMMING || prm->window_kind > FDLP_WINDOW_SQUARE) {
      plan_free(p); return set_err(FDLP_E_ARG, "invalid window kind%s", "");
    }
    const long double theta = N > 1 ? 2.0L * M_PIl / (N - 1) : 0.0L;
    for (int a = 0; a < p->ham_na; ++a) H.push_back(make_double2((double)cosl(theta * 64.0L * a), (double)sinl(theta * 64.0L * a)));
    for (int b = 0; b < 64; ++b) H.push_back(make_double2((double)cosl(theta * b), (double)sinl(theta * b)));
    if ((rc = upload(&p->d_hamt, H))) { plan_free(p); return rc; }
  }
  {
    std::vector<double> full;
    if (prm->fbank_kind == FDLP_FBANK_MEL) full = make_mel_fbank(J, 2 * N, prm->srate, prm->fbank_args[0]);
    else if (prm->fbank_kind == FDLP_FBANK_COCHLEAR)
      full = make_cochlear_fbank(J, 2 * N, prm->srate, prm->fbank_args[0], prm->fbank_args[1], (int)prm->fbank_args[2],
                                 prm->fbank_args[3], prm->fbank_args[4]);
    else if (prm->fbank_kind == FDLP_FBANK_CUSTOM && prm->custom_fbank)
      full.assign(prm->custom_fbank, prm->custom_fbank + (size_t)J * (N + 1));
    else { plan_free(p); return set_err(FDLP_E_ARG, "invalid filter bank kind%s", ""); }
    // fbank[j, 0:-1]  (computeFDLPSpectrogram.py:190)
    p->h_fbank.resize((size_t)J * N);
    std::vector<int32_t> lo(J), hi(J);
    const int rows = N / 8;
    // Row range of the autocorrelation per band: taps that are exactly zero (mel triangles) are skipped, and
    // so are tails below tail_eps * max|F_j| (cochlear skirts decay as 10^(-2.5 bark)).  A dropped tap changes
    // r[k] by at most tail_eps^2 * |c|^2; with the default 1e-12 that is below the float64 rounding of r[0]
    // for any input whose full-band / in-band energy ratio is under 1e8.
    const double eps = prm->tail_eps > 0 ? prm->tail_eps : 0.0;
    int64_t support = 0;
    for (int j = 0; j < J; ++j) {
      double vmax = 0.0;
      for (int k = 0; k < N; ++k) {
        const double v = full[(size_t)j * (N + 1) + k];
        p->h_fbank[(size_t)j * N + k] = v;
        vmax = std::max(vmax, fabs(v));
      }
      const double thr = eps * vmax;
      int first = N, last = -1;
      for (int k = 0; k < N; ++k) {
        const double v = fabs(p->h_fbank[(size_t)j * N + k]);
        if (v != 0.0 && v >= thr) { if (k < first) first = k; last = k; }
      }
      if (last < 0) { lo[j] = 0; hi[j] = 0; }
      else {
        lo[j] = (first / 8) & ~3;
        hi[j] = std::min(rows, ((last / 8 + 1) + 3) & ~3);
      }
      support += (int64_t)(hi[j] - lo[j]) * 8;
    }
    g.support_taps = support;
    std::vector<double> fpad((size_t)J * p->ld);
    for (int j = 0; j < J; ++j)
      for (int k = 0; k < p->ld; ++k) fpad[(size_t)j * p->ld + k] = p->h_fbank[(size_t)j * N + (k % N)];
    if ((rc = upload(&p->d_fbank, fpad)) || (rc = upload(&p->d_row_lo, lo)) || (rc = upload(&p->d_row_hi, hi))) {
      plan_free(p); return rc;
    }
    if ((rc = build_families(p, lo, hi))) { plan_free(p); return rc; }
  }
  {
    std::vector<double> w(prm->weights, prm->weights + g.L);
    p->Lp = (g.L + 3) & ~3;
    p->env_ny = (g.kk + 79) / 80;          // envelope kernel: 10 blocks of 8 frames per warp pass
    p->ldb = p->env_ny * 80;
    std::vector<double> inv_n(g.L, 0.0), wr(g.kk), basis((size_t)p->Lp * p->ldb, 0.0);
    for (int n = 1; n < g.L; ++n) inv_n[n] = 1.0 / n;
    // cosine basis of fft(ms, M): cos(2 pi n t / M), argument reduced exactly in integers
    for (int n = 0; n < g.L; ++n)
      for (int t = 0; t < g.kk; ++t)
        basis[(size_t)n * p->ldb + t] = (double)cosl(2.0L * M_PIl * (long double)(((long long)n * t) % g.M) / g.M);
    std::vector<double> hn = np_hanning(g.kk), hm = np_hamming(g.kk);
    for (int t = 0; t < g.kk; ++t) wr[t] = hn[t] / hm[t];
    if ((rc = upload(&p->d_weights, w)) || (rc = upload(&p->d_inv_n, inv_n)) || (rc = upload(&p->d_basis, basis)) ||
        (rc = upload(&p->d_wratio, wr))) {
      plan_free(p); return rc;
    }
  }

  // ---- kernel attributes ----
  int max_smem = 0;
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, p->device);
  if (max_smem <= 0) max_smem = 227 * 1024;  // building / probing without a device
  // split twiddle and unpack tables staged behind the data
  const size_t tws_bytes = ((size_t)p->fft.tws_n + 2 * (size_t)(p->unp_na + 64) + (size_t)(p->ham_na + 64)) * sizeof(double2);
  p->dct_fp64 = !prm->dct_fp32 && (size_t)N * 8 + tws_bytes <= (size_t)max_smem;
  p->dct_smem = (size_t)N * (p->dct_fp64 ? 8 : 4) + tws_bytes;
  if (!p->dct_fp64 && !prm->dct_fp32 && N % 4 == 0) {
    // float64 does not fit one CTA: split the first radix-2 stage over two CTAs (dct.cuh dct_half_kernel)
    int radix[kMaxRadixPasses], npass = 0;
    if (factor_radices(N / 4, radix, &npass) == 0) {
      if ((rc = build_fft_tables(N / 4, &p->fft_half))) { plan_free(p); return rc; }
      const int n2 = N / 4;
      p->tw1_na = (n2 + 63) / 64;
      std::vector<double2> t1;
      for (int a = 0; a < p->tw1_na; ++a) { const long double ang = -2.0L * M_PIl * (64.0L * a) / (N / 2); t1.push_back(make_double2((double)cosl(ang), (double)sinl(ang))); }
      for (int b = 0; b < 64; ++b) { const long double ang = -2.0L * M_PIl * b / (N / 2); t1.push_back(make_double2((double)cosl(ang), (double)sinl(ang))); }
      if ((rc = upload(&p->d_tw1, t1))) { plan_free(p); return rc; }
      const size_t half_bytes = (size_t)n2 * 16 + ((size_t)p->fft_half.tws_n + 2 * (size_t)(p->unp_na + 64) +
                                                  (size_t)(p->ham_na + 64) + (size_t)(p->tw1_na + 64)) * sizeof(double2);
      if (half_bytes <= (size_t)max_smem) {
        p->dct_half = true;
        p->dct_fp64 = true;
        p->dct_smem = half_bytes;
      }
    }
  }
  g.dct_is_fp64 = p->dct_fp64 ? 1 : 0;
  if (p->dct_smem > (size_t)max_smem) {
    plan_free(p);
    return set_err(FDLP_E_UNSUPPORTED, "segment length %s%ld does not fit in shared memory", "", N);
  }
  p->ac_smem = (size_t)(kAcWarps / p->a_groups) * (p->na * p->a_groups) * 64 * sizeof(double);
  p->lpc_threads = kLpcMaxThreads;
  while (p->lpc_threads > 32 && (size_t)(p->lpc_threads / kLpcGroup) * lpc_task_stride(prm->order, g.nlags, g.L) * sizeof(double) > (size_t)max_smem / 2)
    p->lpc_threads /= 2;
  p->lpc_smem = (size_t)(p->lpc_threads / kLpcGroup) * lpc_task_stride(prm->order, g.nlags, g.L) * sizeof(double);
  p->ola_smem = (size_t)J * (kOlaFT + 1) * sizeof(double);
  if (p->lpc_smem > (size_t)max_smem || p->ola_smem > (size_t)max_smem || p->ac_smem > (size_t)max_smem) {
    plan_free(p);
    return set_err(FDLP_E_UNSUPPORTED, "configuration needs more shared memory than available%s", "");
  }
#define SET_SMEM(kern, bytes)                                                                              \
  do {                                                                                                     \
    cudaError_t e = raise_smem_limit(kern, (size_t)(bytes));                                               \
    if (e != cudaSuccess) { plan_free(p); return set_err(FDLP_E_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e)); } \
  } while (0)
  if (p->dct_half) {
    SET_SMEM((dct_half_kernel<int16_t>), p->dct_smem);
    SET_SMEM((dct_half_kernel<float>), p->dct_smem);
    SET_SMEM((dct_half_kernel<double>), p->dct_smem);
  } else if (p->dct_fp64) {
    SET_SMEM((dct_segments_kernel<double, int16_t>), p->dct_smem);
    SET_SMEM((dct_segments_kernel<double, float>), p->dct_smem);
    SET_SMEM((dct_segments_kernel<double, double>), p->dct_smem);
  } else {
    SET_SMEM((dct_segments_kernel<float, int16_t>), p->dct_smem);
    SET_SMEM((dct_segments_kernel<float, float>), p->dct_smem);
    SET_SMEM((dct_segments_kernel<float, double>), p->dct_smem);
  }
  SET_SMEM(lpc_cepstrum_kernel, p->lpc_smem);
  SET_SMEM(ola_log_kernel<float>, p->ola_smem);
  SET_SMEM(ola_log_kernel<double>, p->ola_smem);
#define SET_AC(NA_)                                                                     \
  case NA_:                                                                             \
    if (p->a_groups > 1) SET_SMEM((autocorr_dmma_kernel<NA_, true>), p->ac_smem);       \
    else SET_SMEM((autocorr_dmma_kernel<NA_, false>), p->ac_smem);                      \
    break;
  switch (p->na) {
    SET_AC(4) SET_AC(8) SET_AC(12) SET_AC(16) SET_AC(20)
    default: plan_free(p); return set_err(FDLP_E_UNSUPPORTED, "unsupported lag blocking%s", "");
  }
#undef SET_AC
  if (p->fam_ok) {
#define SET_FAM(NA_)                                                        \
  case NA_:                                                                 \
    SET_SMEM((family_block_autocorr_kernel<NA_, false>), p->fam_smem);      \
    SET_SMEM((autocorr_ranges_kernel<NA_, false>), p->fam_smem);            \
    break;
    if (p->fam_groups > 1) {
      SET_SMEM((family_block_autocorr_kernel<20, true>), p->fam_smem);
      SET_SMEM((autocorr_ranges_kernel<20, true>), p->fam_smem);
    } else {
      switch (p->fam_na) {
        SET_FAM(4) SET_FAM(8) SET_FAM(12) SET_FAM(16) SET_FAM(20)
        default: p->fam_ok = false;
      }
    }
#undef SET_FAM
  }
#undef SET_SMEM
  {
    int lo_prio = 0, hi_prio = 0;
    cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio);
    cudaError_t e = cudaStreamCreateWithPriority(&p->side, cudaStreamNonBlocking, hi_prio);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_start, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&p->ev_side, cudaEventDisableTiming);
    for (int i = 0; i < kMaxChunks && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&p->ev_chunk[i], cudaEventDisableTiming);
    if (e != cudaSuccess) { plan_free(p); return set_err(FDLP_E_CUDA, "stream/event creation: %s", cudaGetErrorString(e)); }
  }
  {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, p->device);
    if (sms > 0) p->overlap_cap = sms;
    if (const char* e = getenv("FDLP_OVERLAP_CAP")) p->overlap_cap = atoi(e);
    if (const char* e = getenv("FDLP_OVERLAP_CHUNKS")) p->overlap_chunks = std::max(1, std::min(kMaxChunks, atoi(e)));
  }
  *out = p;
  return FDLP_OK;
}

extern "C" int fdlp_plan_destroy(fdlp_plan* p) {
  plan_free(p);
  return FDLP_OK;
}

extern "C" int fdlp_plan_get_geometry(const fdlp_plan* p, fdlp_geometry* out) {
  if (!p || !out) return set_err(FDLP_E_ARG, "null argument%s", "");
  *out = p->g;
  return FDLP_OK;
}

extern "C" int fdlp_plan_get_fbank(const fdlp_plan* p, double* host_out) {
  if (!p || !host_out) return set_err(FDLP_E_ARG, "null argument%s", "");
  memcpy(host_out, p->h_fbank.data(), p->h_fbank.size() * sizeof(double));
  return FDLP_OK;
}

extern "C" int fdlp_make_fbank(int32_t fbank_kind, const double* fbank_args, int32_t nfilters, int32_t nfft, int32_t srate,
                               double* host_out) {
  if (!fbank_args || !host_out || nfilters < 1 || nfft < 2 || srate < 1) return set_err(FDLP_E_ARG, "bad argument%s", "");
  std::vector<double> f;
  if (fbank_kind == FDLP_FBANK_MEL) f = make_mel_fbank(nfilters, nfft, srate, fbank_args[0]);
  else if (fbank_kind == FDLP_FBANK_COCHLEAR)
    f = make_cochlear_fbank(nfilters, nfft, srate, fbank_args[0], fbank_args[1], (int)fbank_args[2], fbank_args[3], fbank_args[4]);
  else return set_err(FDLP_E_ARG, "fbank_kind must be FDLP_FBANK_MEL or FDLP_FBANK_COCHLEAR%s", "");
  memcpy(host_out, f.data(), f.size() * sizeof(double));
  return FDLP_OK;
}

extern "C" int64_t fdlp_plan_scratch_bytes(const fdlp_plan* p, int32_t n_seg) {
  if (!p) return -1;
  const int64_t J = p->prm.nfilters;
  const int64_t fam = (p->fam_ok && p->ac_mode == 0) ? (int64_t)p->nblk * p->g.nlags : 0;
  return (int64_t)n_seg * 8 * ((int64_t)p->ld + J * p->g.nlags + J * p->g.kk + J * p->Lp + fam);
}

// ---------------------------------------------------------------- launches
static int launch_dct(fdlp_plan* p, const fdlp_batch* b, double* out, int out_ld, int out_pad, cudaStream_t st,
                      int seg0 = 0, int nseg = -1) {
  if (nseg < 0) nseg = b->n_seg;
  DctArgs a;
  memset(&a, 0, sizeof(a));
  a.wav = b->wav; a.utt_wav_off = b->utt_wav_off; a.utt_seg_off = b->utt_seg_off; a.seg_utt = b->seg_utt;
  if (b->noise_mode != FDLP_NOISE_NONE) {
    a.noise = (const int16_t*)b->noise; a.utt_noise_off = b->utt_noise_off; a.utt_alpha = p->d_alpha;
  }
  a.N = p->g.N; a.S = p->g.S; a.n = p->g.N / 2; a.npass = p->fft.npass;
  for (int i = 0; i < p->fft.npass; ++i) a.radix[i] = p->fft.radix[i];
  a.ham = p->d_ham; a.tw = p->fft.d_tw; a.perm = p->fft.d_perm; a.unpackA = p->fft.d_unpackA; a.postB = p->d_postB;
  a.out = out; a.out_ld = out_ld; a.out_pad = out_pad; a.seg0 = seg0;
  a.tws = p->fft.d_tws; a.tws_n = p->fft.tws_n; a.unp = p->d_unp; a.unp_na = p->unp_na; a.hamt = p->d_hamt; a.ham_na = p->ham_na;
  a.win_a0 = p->prm.window_kind == FDLP_WINDOW_HANNING ? 0.5 : p->prm.window_kind == FDLP_WINDOW_SQUARE ? 1.0 : 0.54;
  a.win_a1 = p->prm.window_kind == FDLP_WINDOW_HANNING ? 0.5 : p->prm.window_kind == FDLP_WINDOW_SQUARE ? 0.0 : 0.46;
  for (int i = 0; i < p->fft.npass; ++i) { a.tw_off[i] = p->fft.tw_off[i]; a.tw_na[i] = p->fft.tw_na[i]; a.magic[i] = p->fft.magic[i]; a.fuse[i] = p->fft.fuse[i]; }
  a.vec_ok = (((uintptr_t)b->wav) % 16 == 0) ? 1 : 0;
  static const int dct_threads_env = getenv("FDLP_DCT_THREADS") ? atoi(getenv("FDLP_DCT_THREADS")) : 0;
  const dim3 grid(nseg), block(dct_threads_env >= 64 && dct_threads_env <= 512 ? (dct_threads_env & ~31) : p->dct_threads);
  if (p->dct_half) {
    const FftTables& f = p->fft_half;
    a.npass = f.npass;
    for (int i = 0; i < f.npass; ++i) { a.radix[i] = f.radix[i]; a.tw_off[i] = f.tw_off[i]; a.tw_na[i] = f.tw_na[i]; a.magic[i] = f.magic[i]; a.fuse[i] = f.fuse[i]; }
    a.perm = f.d_perm; a.tws = f.d_tws; a.tws_n = f.tws_n; a.tw1 = p->d_tw1; a.tw1_na = p->tw1_na;
    const dim3 grid2(2 * (unsigned)nseg);
    if (b->wav_dtype == FDLP_WAV_INT16) dct_half_kernel<int16_t><<<grid2, block, p->dct_smem, st>>>(a);
    else if (b->wav_dtype == FDLP_WAV_FLOAT64) dct_half_kernel<double><<<grid2, block, p->dct_smem, st>>>(a);
    else dct_half_kernel<float><<<grid2, block, p->dct_smem, st>>>(a);
    FDLP_CUDA(cudaGetLastError());
    return 0;
  }
  if (b->wav_dtype == FDLP_WAV_INT16) {
    if (p->dct_fp64) dct_segments_kernel<double, int16_t><<<grid, block, p->dct_smem, st>>>(a);
    else dct_segments_kernel<float, int16_t><<<grid, block, p->dct_smem, st>>>(a);
  } else if (b->wav_dtype == FDLP_WAV_FLOAT64) {
    if (p->dct_fp64) dct_segments_kernel<double, double><<<grid, block, p->dct_smem, st>>>(a);
    else dct_segments_kernel<float, double><<<grid, block, p->dct_smem, st>>>(a);
  } else {
    if (p->dct_fp64) dct_segments_kernel<double, float><<<grid, block, p->dct_smem, st>>>(a);
    else dct_segments_kernel<float, float><<<grid, block, p->dct_smem, st>>>(a);
  }
  FDLP_CUDA(cudaGetLastError());
  return 0;
}

static int launch_autocorr_direct(fdlp_plan* p, const double* dct, int n_seg, double* r, cudaStream_t st) {
  AcArgs a;
  a.dct = dct; a.fbank = p->d_fbank; a.row_lo = p->d_row_lo; a.row_hi = p->d_row_hi; a.r = r;
  a.N = p->g.N; a.J = p->prm.nfilters; a.nlags = p->g.nlags; a.rows = p->g.N / 8; a.n_seg = n_seg; a.ld = p->ld;
  a.a_groups = p->a_groups; a.na_total = p->na * p->a_groups;
  const int bpc = kAcWarps / p->a_groups;
  const int nbg = (a.J + bpc - 1) / bpc;
  const dim3 grid((unsigned)((size_t)n_seg * nbg)), block(kAcWarps * 32);
#define RUN_AC(NA_)                                                                                   \
  case NA_:                                                                                           \
    if (p->a_groups > 1) autocorr_dmma_kernel<NA_, true><<<grid, block, p->ac_smem, st>>>(a);         \
    else autocorr_dmma_kernel<NA_, false><<<grid, block, p->ac_smem, st>>>(a);                        \
    break;
  switch (p->na) {
    RUN_AC(4) RUN_AC(8) RUN_AC(12) RUN_AC(16) RUN_AC(20)
    default: return set_err(FDLP_E_UNSUPPORTED, "unsupported lag blocking%s", "");
  }
#undef RUN_AC
  FDLP_CUDA(cudaGetLastError());
  return 0;
}

static int ensure_family_scratch(fdlp_plan* p, int n_seg) {
  if (n_seg > p->cap_seg_p) {
    cudaFree(p->d_p); p->d_p = nullptr; p->cap_seg_p = 0;
    FDLP_CUDA(cudaMalloc((void**)&p->d_p, (size_t)n_seg * p->nblk * p->g.nlags * sizeof(double)));
    p->cap_seg_p = n_seg;
  }
  return 0;
}

// family path: shared block sums -> per-band combination -> direct remainder (autocorr.cuh)
static int launch_autocorr_family(fdlp_plan* p, const double* dct, int n_seg, double* r, cudaStream_t st, bool prof) {
  int rc = ensure_family_scratch(p, n_seg);
  if (rc) return rc;
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev_ac[0], st));
  FamArgs f;
  f.dct = dct; f.tmpl = p->d_tmpl; f.p = p->d_p; f.bend = p->d_bend; f.task_fam = p->d_task_fam;
  f.task_blk0 = p->d_task_blk0; f.task_blk1 = p->d_task_blk1; f.task_step0 = p->d_task_step0;
  f.n_seg = n_seg; f.ld = p->ld; f.nlags = p->g.nlags; f.nblk_total = p->nblk; f.ntask = p->fam_ntask;
  f.lag_groups = p->fam_groups;
  // persistent grids: every resident warp draws work items from a counter until none are left
  FDLP_CUDA(cudaMemsetAsync(p->d_next, 0, 2 * sizeof(unsigned long long), st));
  f.next = p->d_next;
  static const int ac_ctas_env = getenv("FDLP_AC_CTAS") ? atoi(getenv("FDLP_AC_CTAS")) : 0;  // A/B runs: CTAs per SM
  const int resident = p->sms * (ac_ctas_env > 0 ? ac_ctas_env : (p->fam_groups > 1 ? 2 : 3));
  const long long tasks = (long long)n_seg * p->fam_ntask * p->fam_groups;
  const dim3 gridA((unsigned)std::min<long long>((tasks + kAcWarps - 1) / kAcWarps, resident)), block(kAcWarps * 32);
  AcRangeArgs d;
  d.dct = dct; d.fbank = p->d_fbank; d.nrng = p->d_nrng; d.rng = p->d_rng; d.r = r;
  d.J = p->prm.nfilters; d.nlags = p->g.nlags; d.n_seg = n_seg; d.ld = p->ld;
  d.lag_groups = p->fam_groups;
  d.next = p->d_next + 1;
  const long long items = (long long)n_seg * d.J * p->fam_groups;
  const dim3 gridD((unsigned)std::min<long long>((items + kAcWarps - 1) / kAcWarps, resident));
  d.p = p->d_p; d.ra = p->d_ra; d.rb = p->d_rb; d.wsq = p->d_wsq; d.nfam = p->nfam; d.nblk_total = p->nblk;
  ScanArgs c;
  c.p = p->d_p; c.n_scan = 0; c.n_seg = n_seg; c.nblk_total = p->nblk; c.nlags = p->g.nlags;
  for (int g = 0; g < kCombMaxFam; ++g) {
    c.mode[g] = 0; c.boff[g] = 0; c.nblk[g] = 0;
    d.mode[g] = p->comb_mode[g]; d.boff[g] = p->comb_boff[g];
  }
  for (int g = 0; g < p->nfam; ++g)
    if (p->comb_mode[g] != kCombRange && p->comb_nband[g] > 0) {
      c.mode[c.n_scan] = p->comb_mode[g]; c.boff[c.n_scan] = p->comb_boff[g]; c.nblk[c.n_scan] = p->comb_nblk[g];
      ++c.n_scan;
    }
  const dim3 gridC((unsigned)((size_t)n_seg * std::max(c.n_scan, 1))), blockC((unsigned)((p->g.nlags + 31) & ~31));
#define RUN_FAM_A(NA_) case NA_: family_block_autocorr_kernel<NA_, false><<<gridA, block, p->fam_smem, st>>>(f); break;
#define RUN_FAM_D(NA_) case NA_: autocorr_ranges_kernel<NA_, false><<<gridD, block, p->fam_smem, st>>>(d); break;
  if (p->fam_groups > 1) {
    family_block_autocorr_kernel<20, true><<<gridA, block, p->fam_smem, st>>>(f);
  } else {
    switch (p->fam_na) {
      RUN_FAM_A(4) RUN_FAM_A(8) RUN_FAM_A(12) RUN_FAM_A(16) RUN_FAM_A(20)
      default: return set_err(FDLP_E_UNSUPPORTED, "unsupported lag blocking%s", "");
    }
  }
  FDLP_CUDA(cudaGetLastError());
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev_ac[1], st));
  if (c.n_scan > 0) family_scan_kernel<<<gridC, blockC, 0, st>>>(c);
  FDLP_CUDA(cudaGetLastError());
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev_ac[2], st));
  if (p->fam_groups > 1) {
    autocorr_ranges_kernel<20, true><<<gridD, block, p->fam_smem, st>>>(d);
  } else {
    switch (p->fam_na) {
      RUN_FAM_D(4) RUN_FAM_D(8) RUN_FAM_D(12) RUN_FAM_D(16) RUN_FAM_D(20)
      default: return set_err(FDLP_E_UNSUPPORTED, "unsupported lag blocking%s", "");
    }
  }
#undef RUN_FAM_A
#undef RUN_FAM_D
  FDLP_CUDA(cudaGetLastError());
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev_ac[3], st));
  return 0;
}

static int launch_autocorr(fdlp_plan* p, const double* dct, int n_seg, double* r, cudaStream_t st, bool prof = false) {
  if (p->fam_ok && p->ac_mode == 0) return launch_autocorr_family(p, dct, n_seg, r, st, prof);
  return launch_autocorr_direct(p, dct, n_seg, r, st);
}

static int launch_lpc(fdlp_plan* p, const double* r, int n_seg, double* lpc_out, double* gain_out, double* cep_out,
                      void* env, int32_t* status, const int32_t* seg_utt, cudaStream_t st, size_t task0 = 0,
                      int grid_cap = 0, bool env_f32 = false) {
  LpcArgs a;
  a.r = r; a.n_tasks = n_seg * p->prm.nfilters; a.J = p->prm.nfilters;
  a.order = p->prm.order; a.nlags = p->g.nlags; a.L = p->g.L; a.Lp = p->Lp;
  a.weights = p->d_weights; a.inv_n = p->d_inv_n;
  a.lpc_out = lpc_out; a.gain_out = gain_out; a.cep_out = cep_out;
  // every chunk starts on a multiple of 8 tasks (chunk boundaries are whole segments, J tasks each, padded below)
  a.ms = p->d_ms + (task0 + 7) / 8 * 8 * p->Lp;
  a.status = status; a.seg_utt = seg_utt;
  // register-resident kernel when the order and the cepstrum length fit 8 lanes x NS slots, else shared memory
  const int need = std::max(p->prm.order + 1, p->g.L);
  static const bool force_smem = getenv("FDLP_LPC_SMEM") != nullptr;
  int lpc_grid = (a.n_tasks + 15) / 16;
  if (grid_cap > 0) lpc_grid = std::min(lpc_grid, grid_cap);
  static const bool force_rot = getenv("FDLP_LPC_ROT") != nullptr;  // the rotation-per-order register kernel (A/B runs)
  static const bool force_l16 = getenv("FDLP_LPC_L16") != nullptr;  // A/B: 16 lanes per solve for orders below 160
  if (need <= 80 && !force_smem && !force_rot) {
    lpc_cepstrum_blk_kernel<2, 8><<<lpc_grid, 128, 0, st>>>(a);
  } else if (p->prm.order + 1 <= 160 && force_l16 && !force_smem && !force_rot) {
    int grid16 = (a.n_tasks + 7) / 8;
    if (grid_cap > 0) grid16 = std::min(grid16, grid_cap);
    lpc_cepstrum_blk_kernel<2, 16><<<grid16, 128, 0, st>>>(a);
  } else if (p->prm.order + 1 <= 160 && !force_smem && !force_rot) {
    // any cepstrum length: beyond 160 coefficients the kernel's vectors are circular (REVERB recipe: L = 300)
    lpc_cepstrum_blk_kernel<4, 8><<<lpc_grid, 128, 0, st>>>(a);
  } else if (p->prm.order + 1 <= 320 && !force_smem && !force_rot) {
    // orders 160..319: 16 lanes per solve, 8 solves per CTA
    int grid16 = (a.n_tasks + 7) / 8;
    if (grid_cap > 0) grid16 = std::min(grid16, grid_cap);
    lpc_cepstrum_blk_kernel<4, 16><<<grid16, 128, 0, st>>>(a);
  } else if (need <= 64 && !force_smem) {
    lpc_cepstrum_reg_kernel<8><<<lpc_grid, 128, 0, st>>>(a);
  } else if (need <= 160 && !force_smem) {
    lpc_cepstrum_reg_kernel<20><<<lpc_grid, 128, 0, st>>>(a);
  } else {
    const int tpc = p->lpc_threads / kLpcGroup;
    lpc_cepstrum_kernel<<<(a.n_tasks + tpc - 1) / tpc, p->lpc_threads, p->lpc_smem, st>>>(a);
  }
  FDLP_CUDA(cudaGetLastError());
  if (p->profiling && p->ev_lpc) FDLP_CUDA(cudaEventRecord(p->ev_lpc, st));
  if (!env) return 0;  // cepstra only (fdlp_modspec_forward)
  EnvArgs e;
  e.ms = a.ms; e.basis = p->d_basis; e.wratio = p->d_wratio; e.env = env;
  e.n_tasks = a.n_tasks; e.Lp = p->Lp; e.ldb = p->ldb; e.kk = p->g.kk;
  e.status = env_f32 ? status : nullptr; e.seg_utt = seg_utt; e.J = p->prm.nfilters;
  const dim3 grid((a.n_tasks + 8 * kEnvMA * kEnvWarps - 1) / (8 * kEnvMA * kEnvWarps), p->env_ny);
  if (env_f32) envelope_dmma_kernel<10, float><<<grid, kEnvWarps * 32, 0, st>>>(e);
  else envelope_dmma_kernel<10, double><<<grid, kEnvWarps * 32, 0, st>>>(e);
  FDLP_CUDA(cudaGetLastError());
  return 0;
}

static int check_batch(const fdlp_plan* p, const fdlp_batch* b) {
  if (!p || !b) return set_err(FDLP_E_ARG, "null argument%s", "");
  FDLP_CHECK_SIZE(b, fdlp_batch);
  if (b->n_utt < 0 || b->n_seg < 0 || b->n_frames < 0) return set_err(FDLP_E_ARG, "negative batch sizes%s", "");
  if (b->n_seg > 0 && (!b->wav || !b->utt_wav_off || !b->utt_seg_off || !b->seg_utt))
    return set_err(FDLP_E_ARG, "null batch pointer%s", "");
  if (b->wav_dtype != FDLP_WAV_INT16 && b->wav_dtype != FDLP_WAV_FLOAT32 && b->wav_dtype != FDLP_WAV_FLOAT64)
    return set_err(FDLP_E_ARG, "bad wav dtype%s", "");
  if (b->noise_mode != FDLP_NOISE_NONE && (!b->noise || !b->utt_noise_off || !b->utt_snr_db))
    return set_err(FDLP_E_ARG, "noise mixing requested without noise arrays%s", "");
  if (b->noise_mode != FDLP_NOISE_NONE && b->wav_dtype == FDLP_WAV_FLOAT64)
    return set_err(FDLP_E_ARG, "float64 samples are the output of fdlp_reverb_forward, which mixes the noise itself%s", "");
  return 0;
}

static int launch_noise(fdlp_plan* p, const fdlp_batch* b, cudaStream_t st) {
  if (b->noise_mode == FDLP_NOISE_NONE) return 0;
  NoiseArgs a;
  a.wav = b->wav; a.wav_dtype = b->wav_dtype; a.utt_wav_off = b->utt_wav_off; a.noise = (const int16_t*)b->noise;
  a.utt_noise_off = b->utt_noise_off; a.utt_snr_db = b->utt_snr_db; a.utt_alpha = p->d_alpha; a.mode = b->noise_mode;
  noise_alpha_kernel<<<b->n_utt, kNoiseThreads, 0, st>>>(a);
  FDLP_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int fdlp_forward(fdlp_plan* p, const fdlp_batch* b, void* stream) {
  int rc = check_batch(p, b);
  if (rc) return rc;
  if (b->n_seg == 0 || b->n_utt == 0) return FDLP_OK;
  if (!b->feats || !b->utt_frame_off || !b->seg_ptr) return set_err(FDLP_E_ARG, "null output / frame pointers%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  if ((rc = ensure_scratch(p, b->n_seg, b->n_utt))) return rc;
  const bool prof = p->profiling;
  const int J = p->prm.nfilters;
  int nchunk = 1;
  if (p->overlap && !prof && p->side) nchunk = std::min(p->overlap_chunks, b->n_seg / 16);
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev[0], st));
  if ((rc = launch_noise(p, b, st))) return rc;
  if (nchunk <= 1) {
    if ((rc = launch_dct(p, b, p->d_dct, p->ld, p->pad, st))) return rc;
    if (prof) FDLP_CUDA(cudaEventRecord(p->ev[1], st));
    if ((rc = launch_autocorr(p, p->d_dct, b->n_seg, p->d_r, st, prof))) return rc;
    if (prof) FDLP_CUDA(cudaEventRecord(p->ev[2], st));
    if ((rc = launch_lpc(p, p->d_r, b->n_seg, nullptr, nullptr, nullptr, p->d_env, b->status, b->seg_utt, st, 0, 0,
                         p->env_f32))) return rc;
    if (prof) FDLP_CUDA(cudaEventRecord(p->ev[3], st));
  } else {
    // The serial recursions of K4 are latency bound and leave the FP64 pipe idle; K3 saturates it.  Chunks of
    // segments flow through two streams so that K4/K5 of chunk c share the SMs with K2/K3 of chunk c+1.
    FDLP_CUDA(cudaEventRecord(p->ev_start, st));
    FDLP_CUDA(cudaStreamWaitEvent(p->side, p->ev_start, 0));
    for (int c = 0; c < nchunk; ++c) {
      const int s0 = (int)((int64_t)b->n_seg * c / nchunk), s1 = (int)((int64_t)b->n_seg * (c + 1) / nchunk);
      const int ns = s1 - s0;
      double* dct_c = p->d_dct + (size_t)s0 * p->ld;
      double* r_c = p->d_r + (size_t)s0 * J * p->g.nlags;
      if ((rc = launch_dct(p, b, dct_c, p->ld, p->pad, st, s0, ns))) return rc;
      if ((rc = launch_autocorr(p, dct_c, ns, r_c, st))) return rc;
      FDLP_CUDA(cudaEventRecord(p->ev_chunk[c], st));
      FDLP_CUDA(cudaStreamWaitEvent(p->side, p->ev_chunk[c], 0));
      void* env_c = p->env_f32 ? (void*)((float*)p->d_env + (size_t)s0 * J * p->g.kk)
                               : (void*)(p->d_env + (size_t)s0 * J * p->g.kk);
      if ((rc = launch_lpc(p, r_c, ns, nullptr, nullptr, nullptr, env_c, b->status,
                           b->seg_utt + s0, p->side, (size_t)s0 * J, p->overlap_cap, p->env_f32))) return rc;
    }
    FDLP_CUDA(cudaEventRecord(p->ev_side, p->side));
    FDLP_CUDA(cudaStreamWaitEvent(st, p->ev_side, 0));
  }
  OlaArgs o;
  o.env = p->d_env; o.utt_seg_off = b->utt_seg_off; o.utt_frame_off = b->utt_frame_off; o.seg_ptr = b->seg_ptr;
  o.feats = b->feats; o.cmvn_part = nullptr; o.J = p->prm.nfilters; o.kk = p->g.kk; o.kkb2 = p->g.kkb2;
  const int64_t cmvn_slots = cmvn_slot_count(b->n_frames, b->n_utt, kOlaFT);
  const int cmvn_w = 2 * (p->prm.nfilters + 1);
  if (b->cmvn_acc) {
    if (cmvn_slots > p->cap_cmvn_slots) {
      cudaFree(p->d_cmvn_part); p->d_cmvn_part = nullptr; p->cap_cmvn_slots = 0;
      FDLP_CUDA(cudaMalloc((void**)&p->d_cmvn_part, (size_t)cmvn_slots * cmvn_w * sizeof(double)));
      p->cap_cmvn_slots = cmvn_slots;
    }
    FDLP_CUDA(cudaMemsetAsync(p->d_cmvn_part, 0, (size_t)cmvn_slots * cmvn_w * sizeof(double), st));
    o.cmvn_part = p->d_cmvn_part;
  }
  // grid.y covers the longest utterance; the host passes it through n_frames / n_utt bounds
  int64_t max_tiles = (b->max_utt_frames + kOlaFT - 1) / kOlaFT;
  if (max_tiles < 1) max_tiles = 1;
  if (max_tiles > 65535) return set_err(FDLP_E_ARG, "utterance too long%s", "");
  const dim3 grid(b->n_utt, (unsigned)max_tiles), block(256);
  if (p->env_f32) ola_log_kernel<float><<<grid, block, p->ola_smem, st>>>(o);
  else ola_log_kernel<double><<<grid, block, p->ola_smem, st>>>(o);
  FDLP_CUDA(cudaGetLastError());
  if (b->cmvn_acc) {  // fixed-order sum of the per-CTA partials: bit-reproducible stats (no floating-point atomics)
    cmvn_reduce_kernel<<<cmvn_w, 256, 0, st>>>(p->d_cmvn_part, cmvn_slots, cmvn_w, b->cmvn_acc);
    FDLP_CUDA(cudaGetLastError());
  }
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev[4], st));
  return FDLP_OK;
}

extern "C" int fdlp_plan_set_overlap(fdlp_plan* p, int32_t enable) {
  if (!p) return set_err(FDLP_E_ARG, "null argument%s", "");
  p->overlap = enable != 0;
  return FDLP_OK;
}

extern "C" int fdlp_plan_set_profiling(fdlp_plan* p, int32_t enable) {
  if (!p) return set_err(FDLP_E_ARG, "null argument%s", "");
  if (enable && !p->ev[0]) {
    for (int i = 0; i < 5; ++i) FDLP_CUDA(cudaEventCreate(&p->ev[i]));
    for (int i = 0; i < 4; ++i) FDLP_CUDA(cudaEventCreate(&p->ev_ac[i]));
    FDLP_CUDA(cudaEventCreate(&p->ev_lpc));
  }
  p->profiling = enable != 0;
  return FDLP_OK;
}

extern "C" int fdlp_plan_get_stage_ms(fdlp_plan* p, float* out4) {
  if (!p || !out4 || !p->ev[0]) return set_err(FDLP_E_ARG, "profiling was not enabled%s", "");
  FDLP_CUDA(cudaEventSynchronize(p->ev[4]));
  for (int i = 0; i < 4; ++i) FDLP_CUDA(cudaEventElapsedTime(&out4[i], p->ev[i], p->ev[i + 1]));
  return FDLP_OK;
}

extern "C" int fdlp_plan_set_autocorr_mode(fdlp_plan* p, int32_t mode) {
  if (!p || mode < 0 || mode > 1) return set_err(FDLP_E_ARG, "bad argument%s", "");
  p->ac_mode = mode;
  return FDLP_OK;
}

extern "C" int fdlp_plan_get_autocorr_ms(fdlp_plan* p, float* out3) {
  if (!p || !out3 || !p->ev[0]) return set_err(FDLP_E_ARG, "profiling was not enabled%s", "");
  out3[0] = out3[1] = out3[2] = 0.f;
  FDLP_CUDA(cudaEventSynchronize(p->ev[4]));
  if (p->fam_ok && p->ac_mode == 0) {
    for (int i = 0; i < 3; ++i) FDLP_CUDA(cudaEventElapsedTime(&out3[i], p->ev_ac[i], p->ev_ac[i + 1]));
  } else {
    FDLP_CUDA(cudaEventElapsedTime(&out3[2], p->ev[1], p->ev[2]));
  }
  return FDLP_OK;
}

extern "C" int fdlp_plan_get_kernel_ms(fdlp_plan* p, float* out7) {
  if (!p || !out7 || !p->ev[0]) return set_err(FDLP_E_ARG, "profiling was not enabled%s", "");
  for (int i = 0; i < 7; ++i) out7[i] = 0.f;
  FDLP_CUDA(cudaEventSynchronize(p->ev[4]));
  FDLP_CUDA(cudaEventElapsedTime(&out7[0], p->ev[0], p->ev[1]));
  if (p->fam_ok && p->ac_mode == 0) {
    for (int i = 0; i < 3; ++i) FDLP_CUDA(cudaEventElapsedTime(&out7[1 + i], p->ev_ac[i], p->ev_ac[i + 1]));
  } else {
    FDLP_CUDA(cudaEventElapsedTime(&out7[3], p->ev[1], p->ev[2]));
  }
  FDLP_CUDA(cudaEventElapsedTime(&out7[4], p->ev[2], p->ev_lpc));
  FDLP_CUDA(cudaEventElapsedTime(&out7[5], p->ev_lpc, p->ev[3]));
  FDLP_CUDA(cudaEventElapsedTime(&out7[6], p->ev[3], p->ev[4]));
  return FDLP_OK;
}

extern "C" int fdlp_stage_dct(fdlp_plan* p, const fdlp_batch* b, double* dct_out, void* stream) {
  int rc = check_batch(p, b);
  if (rc) return rc;
  if (!dct_out) return set_err(FDLP_E_ARG, "null output%s", "");
  if (b->n_seg == 0) return FDLP_OK;
  if ((rc = ensure_scratch(p, 1, b->n_utt))) return rc;
  if ((rc = launch_noise(p, b, (cudaStream_t)stream))) return rc;
  return launch_dct(p, b, dct_out, p->g.N, 0, (cudaStream_t)stream);
}

extern "C" int fdlp_stage_autocorr(fdlp_plan* p, const double* dct, int32_t n_seg, double* r_out, void* stream) {
  if (!p || !dct || !r_out || n_seg < 0) return set_err(FDLP_E_ARG, "bad argument%s", "");
  if (n_seg == 0) return FDLP_OK;
  int rc = ensure_scratch(p, n_seg, 1);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t N8 = (size_t)p->g.N * sizeof(double), ld8 = (size_t)p->ld * sizeof(double);
  FDLP_CUDA(cudaMemcpy2DAsync(p->d_dct, ld8, dct, N8, N8, n_seg, cudaMemcpyDeviceToDevice, st));
  FDLP_CUDA(cudaMemcpy2DAsync(p->d_dct + p->g.N, ld8, dct, N8, (size_t)p->pad * sizeof(double), n_seg,
                              cudaMemcpyDeviceToDevice, st));
  return launch_autocorr(p, p->d_dct, n_seg, r_out, st);
}

extern "C" int fdlp_stage_lpc_envelope(fdlp_plan* p, const double* r, int32_t n_seg, double* lpc_out, double* gain_out,
                                       double* cep_out, double* env_out, int32_t* seg_status, void* stream) {
  if (!p || !r || !env_out || n_seg < 0) return set_err(FDLP_E_ARG, "bad argument%s", "");
  if (n_seg == 0) return FDLP_OK;
  int rc = ensure_scratch(p, n_seg, 1);
  if (rc) return rc;
  return launch_lpc(p, r, n_seg, lpc_out, gain_out, cep_out, env_out, seg_status, nullptr, (cudaStream_t)stream);
}

// ================================================================ mel plan
struct mel_plan {
  mel_params prm;
  int flen = 0, shift = 0, n = 0;
  FftTables fft;
  std::vector<double> h_fbank;  // [J, n+1]
  double* d_ham = nullptr;
  float* d_fbank = nullptr;
  int32_t *d_lo = nullptr, *d_hi = nullptr;
  int32_t* d_frame_utt = nullptr;
  int64_t cap_frames = 0;
  size_t smem = 0;
  // warp-per-frame kernel (nfft/2 = 32 * R2, R2 in {16, 32})
  int warp_r2 = 0, sms = 148;
  size_t warp_smem = 0;
  int warp_m1 = 0, wts_smem = 0;
  float *d_ham32 = nullptr, *d_wts = nullptr;
  float2 *d_tw4 = nullptr, *d_unpackA32 = nullptr;
  int4* d_band = nullptr;
  double* d_cmvn_part = nullptr;  // [kMelCmvnCtas, 2, J+1]
  // second-generation kernel (mel_mma.cuh): nfft = 2048, magnitudes in registers, filterbank on the tensor cores
  bool mma_ok = false;
  MelMmaArgs mma;             // tile tables filled at plan creation, pointers per call
  float2* d_lane_tw = nullptr;
  float4* d_bfrag = nullptr;
};

constexpr int kMelCmvnCtas = 296;

// CMVN partial sums of the finished feature matrix, slots added in a fixed order (bit-reproducible)
static int mel_cmvn(mel_plan* p, const mel_batch* b, cudaStream_t st) {
  if (!b->cmvn_acc) return 0;
  const int J = p->prm.nfilters, w = 2 * (J + 1);
  if (!p->d_cmvn_part) FDLP_CUDA(cudaMalloc((void**)&p->d_cmvn_part, (size_t)kMelCmvnCtas * w * sizeof(double)));
  CmvnSumArgs c;
  c.feats = b->feats; c.part = p->d_cmvn_part; c.n_frames = b->n_frames; c.J = J;
  c.chunk = (b->n_frames + kMelCmvnCtas - 1) / kMelCmvnCtas;
  const int rows_par = std::max(1, 256 / J);
  cmvn_sums_kernel<<<kMelCmvnCtas, 256, (size_t)rows_par * 2 * std::min(J, 256) * sizeof(double), st>>>(c);
  FDLP_CUDA(cudaGetLastError());
  cmvn_reduce_kernel<<<w, 256, 0, st>>>(p->d_cmvn_part, kMelCmvnCtas, w, b->cmvn_acc);
  FDLP_CUDA(cudaGetLastError());
  return 0;
}

template <int R2, int M1, typename S>
static cudaError_t mel_warp_launch(const MelWarpArgs& w, dim3 grid, size_t smem, cudaStream_t st) {
  cudaError_t e = raise_smem_limit(mel_frames_warp_kernel<R2, M1, S>, smem);
  if (e != cudaSuccess) return e;
  mel_frames_warp_kernel<R2, M1, S><<<grid, kMelWarps * 32, smem, st>>>(w);
  return cudaGetLastError();
}
template <int R2, typename S>
static cudaError_t mel_warp_launch_m1(int m1, const MelWarpArgs& w, dim3 grid, size_t smem, cudaStream_t st) {
  if (m1 == 5) return mel_warp_launch<R2, 5, S>(w, grid, smem, st);
  if (m1 == 8 && R2 > 16) return mel_warp_launch<R2, 8, S>(w, grid, smem, st);
  return mel_warp_launch<R2, 0, S>(w, grid, smem, st);
}

static void mel_free(mel_plan* p) {
  if (!p) return;
  free_fft_tables(&p->fft);
  cudaFree(p->d_ham); cudaFree(p->d_fbank); cudaFree(p->d_lo); cudaFree(p->d_hi); cudaFree(p->d_frame_utt);
  cudaFree(p->d_ham32); cudaFree(p->d_tw4); cudaFree(p->d_unpackA32); cudaFree(p->d_wts); cudaFree(p->d_band);
  cudaFree(p->d_cmvn_part); cudaFree(p->d_lane_tw); cudaFree(p->d_bfrag);
  delete p;
}

extern "C" int mel_plan_create(const mel_params* prm, mel_plan** out) {
  if (!prm || !out) return set_err(FDLP_E_ARG, "null argument%s", "");
  *out = nullptr;
  FDLP_CHECK_SIZE(prm, mel_params);
  if (prm->nfilters < 1 || prm->nfft < 8 || (prm->nfft & 1) || prm->frate < 1 || !(prm->fduration > 0))
    return set_err(FDLP_E_ARG, "invalid mel_params%s", "");
  mel_plan* p = new (std::nothrow) mel_plan();
  if (!p) return set_err(FDLP_E_NOMEM, "out of host memory%s", "");
  p->prm = *prm;
  p->flen = (int)(prm->srate * prm->fduration);   // features.py:134
  p->shift = (int)(prm->srate / (double)prm->frate);  // features.py:135
  p->n = prm->nfft / 2;
  if (p->flen < 1 || p->shift < 1) { mel_free(p); return set_err(FDLP_E_ARG, "empty frames%s", ""); }
  int rc = build_fft_tables(p->n, &p->fft);
  if (rc) { mel_free(p); return rc; }
  const int J = prm->nfilters, nb = p->n + 1;
  if (prm->fbank_kind == FDLP_FBANK_MEL) p->h_fbank = make_mel_fbank(J, prm->nfft, prm->srate, prm->fbank_args[0]);
  else if (prm->fbank_kind == FDLP_FBANK_COCHLEAR)
    p->h_fbank = make_cochlear_fbank(J, prm->nfft, prm->srate, prm->fbank_args[0], prm->fbank_args[1],
                                     (int)prm->fbank_args[2], prm->fbank_args[3], prm->fbank_args[4]);
  else if (prm->fbank_kind == FDLP_FBANK_CUSTOM && prm->custom_fbank)
    p->h_fbank.assign(prm->custom_fbank, prm->custom_fbank + (size_t)J * nb);
  else { mel_free(p); return set_err(FDLP_E_ARG, "invalid filter bank kind%s", ""); }
  std::vector<float> f32(p->h_fbank.size());
  std::vector<int32_t> lo(J), hi(J);
  for (int j = 0; j < J; ++j) {
    int first = nb, last = -1;
    for (int k = 0; k < nb; ++k) {
      const double v = p->h_fbank[(size_t)j * nb + k];
      f32[(size_t)j * nb + k] = (float)v;
      if (v != 0.0) { if (k < first) first = k; last = k; }
    }
    lo[j] = last < 0 ? 0 : first;
    hi[j] = last < 0 ? 0 : last + 1;
  }
  std::vector<double> ham = np_hamming(p->flen);
  if ((rc = upload(&p->d_ham, ham)) || (rc = upload(&p->d_fbank, f32)) || (rc = upload(&p->d_lo, lo)) ||
      (rc = upload(&p->d_hi, hi))) { mel_free(p); return rc; }
  if ((p->n == 512 || p->n == 1024) && J <= 32 * kMelMaxBandsPerLane && !getenv("FDLP_MEL_CTA")) {
    const int R2 = p->n / 32, n = p->n;
    const int used = std::min(p->flen, prm->nfft);
    std::vector<float> ham32((size_t)prm->nfft, 0.f);  // zero beyond the frame: the kernel loads whole 64-sample rows
    for (int i = 0; i < used; ++i) ham32[i] = (float)ham[i];
    std::vector<float2> tw4((size_t)R2 * 32), A32(n + 1);
    for (int k1 = 0; k1 < R2; ++k1)
      for (int l = 0; l < 32; ++l) {
        const long double ang = -2.0L * M_PIl * (long double)((k1 * l) % n) / n;
        tw4[(size_t)k1 * 32 + l] = make_float2((float)cosl(ang), (float)sinl(ang));
      }
    for (int k = 0; k <= n; ++k) {
      const long double ang = -2.0L * M_PIl * k / (2.0L * n);
      A32[k] = make_float2((float)cosl(ang), (float)sinl(ang));
    }
    // band supports packed back to back: weight of (band j, bin b) at wts[band[j].z + b - band[j].x]
    std::vector<float> wts;
    std::vector<int4> band(J);
    for (int j = 0; j < J; ++j) {
      band[j] = make_int4(lo[j], hi[j] - lo[j], (int)wts.size(), 0);
      for (int k = lo[j]; k < hi[j]; ++k) wts.push_back(f32[(size_t)j * nb + k]);
    }
    if (wts.empty()) wts.push_back(0.f);
    if ((rc = upload(&p->d_ham32, ham32)) || (rc = upload(&p->d_tw4, tw4)) || (rc = upload(&p->d_unpackA32, A32)) ||
        (rc = upload(&p->d_wts, wts)) || (rc = upload(&p->d_band, band))) {
      mel_free(p); return rc;
    }
    p->warp_r2 = R2;
    const int rows = (used + 63) / 64;
    p->warp_m1 = rows <= 5 ? 5 : (rows <= 8 && R2 > 16) ? 8 : 0;
    // two CTAs per SM: stage the packed weights in shared memory when they fit beside the tiles
    p->wts_smem = mel_warp_smem_bytes(R2, J, (int)wts.size()) <= 113 * 1024 ? (int)wts.size() : 0;
    p->warp_smem = mel_warp_smem_bytes(R2, J, p->wts_smem);
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms > 0) p->sms = sms;
    // mel_mma.cuh (magnitudes in registers, filterbank as a banded 3xTF32 GEMM) is an evidenced experiment, not the
    // default: it takes the shared-memory pipe from 79 % to 29 % busy but is issue-bound at 2700 instructions per frame
    // (3.9 ms against 3.7 ms for mel_frames_warp_kernel on the bench batch); FDLP_MEL_MMA=1 selects it
    if (n == 1024 && J <= 8 * kMmaMaxTiles && getenv("FDLP_MEL_MMA")) {
      // ---- tables of mel_frames_mma_kernel ----
      std::vector<float2> ltw(5 * 32);
      for (int l = 0; l < 32; ++l) {
        const int mult[4] = {1, 8, 16, 24};
        for (int q = 0; q < 4; ++q) {
          const long double ang = -2.0L * M_PIl * (long double)((mult[q] * l) % n) / n;
          ltw[q * 32 + l] = make_float2((float)cosl(ang), (float)sinl(ang));
        }
        const long double ang = -2.0L * M_PIl * l / (2.0L * n);
        ltw[4 * 32 + l] = make_float2((float)cosl(ang), (float)sinl(ang));
      }
      auto rna_tf32 = [](float x) {  // cvt.rna.tf32.f32: 10 mantissa bits, nearest, ties away from zero
        uint32_t b;
        memcpy(&b, &x, 4);
        b = (b + 0x1000u) & ~0x1FFFu;
        float r;
        memcpy(&r, &b, 4);
        return r;
      };
      MelMmaArgs& m = p->mma;
      memset(&m, 0, sizeof(m));
      const int ntile = (J + 7) / 8;
      std::vector<int> t_k0(ntile, 0), t_nk(ntile, 0), order(ntile);
      for (int ti = 0; ti < ntile; ++ti) {
        int klo = nb, khi = 0;
        for (int j = 8 * ti; j < std::min(J, 8 * ti + 8); ++j)
          if (hi[j] > lo[j]) { klo = std::min(klo, (int)lo[j]); khi = std::max(khi, (int)hi[j]); }
        if (khi > klo) { t_k0[ti] = klo / 8 * 8; t_nk[ti] = (khi - t_k0[ti] + 7) / 8; }
        order[ti] = ti;
      }
      // longest tiles first, each to the least loaded warp
      std::sort(order.begin(), order.end(), [&](int x, int y) { return t_nk[x] != t_nk[y] ? t_nk[x] > t_nk[y] : x < y; });
      std::vector<std::vector<int>> per_warp(kMmaWarps);
      std::vector<int> load(kMmaWarps, 0);
      for (int ti : order) {
        int w = 0;
        for (int x = 1; x < kMmaWarps; ++x) if (load[x] < load[w]) w = x;
        per_warp[w].push_back(ti);
        load[w] += t_nk[ti] + 2;
      }
      std::vector<float4> bfrag;
      int slot = 0;
      for (int w = 0; w < kMmaWarps; ++w) {
        m.warp_tile0[w] = slot;
        for (int ti : per_warp[w]) {
          m.tile_n0[slot] = 8 * ti; m.tile_k0[slot] = t_k0[ti]; m.tile_nk[slot] = t_nk[ti];
          m.tile_b0[slot] = (int)(bfrag.size() / 32);
          for (int s_ = 0; s_ < t_nk[ti]; ++s_)
            for (int l = 0; l < 32; ++l) {
              const int j = 8 * ti + (l >> 2), kk = t_k0[ti] + 8 * s_ + (l & 3);
              // (x 0.5: the kernel's magnitudes carry the factor 2 of the unscaled real-transform unpack)
              auto wv = [&](int k) { return (j < J && k < nb) ? 0.5f * f32[(size_t)j * nb + k] : 0.f; };
              const float b0 = wv(kk), b1 = wv(kk + 4);
              const float h0 = rna_tf32(b0), h1 = rna_tf32(b1);
              bfrag.push_back(make_float4(h0, h1, rna_tf32(b0 - h0), rna_tf32(b1 - h1)));
            }
          ++slot;
        }
      }
      m.warp_tile0[kMmaWarps] = slot;
      if (bfrag.empty()) bfrag.push_back(make_float4(0.f, 0.f, 0.f, 0.f));
      if ((rc = upload(&p->d_lane_tw, ltw)) || (rc = upload(&p->d_bfrag, bfrag))) { mel_free(p); return rc; }
      cudaError_t ea = cudaSuccess;
#define MMA_SMEM(M1_)                                                                              \
      if (ea == cudaSuccess) ea = raise_smem_limit(mel_frames_mma_kernel<M1_, int16_t>, mel_mma_smem_bytes()); \
      if (ea == cudaSuccess) ea = raise_smem_limit(mel_frames_mma_kernel<M1_, float>, mel_mma_smem_bytes());
      MMA_SMEM(5) MMA_SMEM(8) MMA_SMEM(0)
#undef MMA_SMEM
      if (ea != cudaSuccess) { mel_free(p); return set_err(FDLP_E_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(ea)); }
      p->mma_ok = true;
    }
  }
  p->smem = (size_t)prm->nfft * sizeof(float) + (size_t)(p->n + 1) * sizeof(float);
  cudaError_t e = raise_smem_limit(mel_frames_kernel<int16_t>, p->smem);
  if (e == cudaSuccess) e = raise_smem_limit(mel_frames_kernel<float>, p->smem);
  if (e != cudaSuccess) { mel_free(p); return set_err(FDLP_E_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e)); }
  *out = p;
  return FDLP_OK;
}

extern "C" int mel_plan_destroy(mel_plan* p) { mel_free(p); return FDLP_OK; }

extern "C" int mel_plan_get_fbank(const mel_plan* p, double* host_out) {
  if (!p || !host_out) return set_err(FDLP_E_ARG, "null argument%s", "");
  memcpy(host_out, p->h_fbank.data(), p->h_fbank.size() * sizeof(double));
  return FDLP_OK;
}

extern "C" int mel_forward(mel_plan* p, const mel_batch* b, void* stream) {
  if (!p || !b) return set_err(FDLP_E_ARG, "null argument%s", "");
  FDLP_CHECK_SIZE(b, mel_batch);
  if (b->n_utt <= 0 || b->n_frames <= 0) return FDLP_OK;
  if (!b->wav || !b->utt_wav_off || !b->utt_frame_off || !b->feats) return set_err(FDLP_E_ARG, "null batch pointer%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  if (b->n_frames > p->cap_frames) {
    cudaFree(p->d_frame_utt); p->d_frame_utt = nullptr; p->cap_frames = 0;
    FDLP_CUDA(cudaMalloc((void**)&p->d_frame_utt, (size_t)b->n_frames * sizeof(int32_t)));
    p->cap_frames = b->n_frames;
  }
  mel_frame_map_kernel<<<b->n_utt, 128, 0, st>>>(b->utt_frame_off, b->n_utt, p->d_frame_utt);
  FDLP_CUDA(cudaGetLastError());
  if (b->wav_dtype != FDLP_WAV_INT16 && b->wav_dtype != FDLP_WAV_FLOAT32) return set_err(FDLP_E_ARG, "bad wav dtype%s", "");
  if (p->mma_ok) {
    MelMmaArgs m = p->mma;
    m.wav = b->wav; m.utt_wav_off = b->utt_wav_off; m.utt_frame_off = b->utt_frame_off; m.frame_utt = p->d_frame_utt;
    m.n_frames = b->n_frames; m.flen = p->flen; m.shift = p->shift; m.nfft = p->prm.nfft;
    m.ham = p->d_ham32; m.lane_tw = p->d_lane_tw; m.bfrag = p->d_bfrag; m.feats = b->feats;
    m.J = p->prm.nfilters; m.power = p->prm.power;
    const long long rounds = (b->n_frames + kMmaFrames - 1) / kMmaFrames;
    const dim3 grid((unsigned)std::min<long long>(rounds, (long long)p->sms * 2)), block(kMmaWarps * 32);
    const bool i16 = b->wav_dtype == FDLP_WAV_INT16;
    const size_t sm = mel_mma_smem_bytes();
#define MMA_RUN(M1_)                                                                      \
    if (i16) mel_frames_mma_kernel<M1_, int16_t><<<grid, block, sm, st>>>(m);             \
    else mel_frames_mma_kernel<M1_, float><<<grid, block, sm, st>>>(m);
    if (p->warp_m1 == 5) { MMA_RUN(5) } else if (p->warp_m1 == 8) { MMA_RUN(8) } else { MMA_RUN(0) }
#undef MMA_RUN
    FDLP_CUDA(cudaGetLastError());
    return mel_cmvn(p, b, st);
  }
  if (p->warp_r2 > 0) {
    MelWarpArgs w;
    memset(&w, 0, sizeof(w));
    w.wav = b->wav; w.utt_wav_off = b->utt_wav_off; w.utt_frame_off = b->utt_frame_off; w.frame_utt = p->d_frame_utt;
    w.n_frames = b->n_frames; w.flen = p->flen; w.shift = p->shift; w.nfft = p->prm.nfft; w.n = p->n;
    w.ham = p->d_ham32; w.tw4 = p->d_tw4; w.unpackA = p->d_unpackA32; w.wts = p->d_wts; w.band = p->d_band;
    w.wts_smem = p->wts_smem; w.feats = b->feats; w.cmvn = b->cmvn_acc;
    w.J = p->prm.nfilters; w.power = p->prm.power;
    const long long need = (b->n_frames + kMelWarps - 1) / kMelWarps;
    const dim3 grid((unsigned)std::min<long long>(need, (long long)p->sms * 2));
    const bool i16 = b->wav_dtype == FDLP_WAV_INT16;
    cudaError_t le;
    if (p->warp_r2 == 32)
      le = i16 ? mel_warp_launch_m1<32, int16_t>(p->warp_m1, w, grid, p->warp_smem, st)
               : mel_warp_launch_m1<32, float>(p->warp_m1, w, grid, p->warp_smem, st);
    else
      le = i16 ? mel_warp_launch_m1<16, int16_t>(p->warp_m1, w, grid, p->warp_smem, st)
               : mel_warp_launch_m1<16, float>(p->warp_m1, w, grid, p->warp_smem, st);
    FDLP_CUDA(le);
    FDLP_CUDA(cudaGetLastError());
    return mel_cmvn(p, b, st);
  }
  MelArgs a;
  memset(&a, 0, sizeof(a));
  a.wav = b->wav; a.utt_wav_off = b->utt_wav_off; a.utt_frame_off = b->utt_frame_off; a.frame_utt = p->d_frame_utt;
  a.flen = p->flen; a.shift = p->shift; a.nfft = p->prm.nfft; a.n = p->n; a.npass = p->fft.npass;
  for (int i = 0; i < p->fft.npass; ++i) a.radix[i] = p->fft.radix[i];
  a.ham = p->d_ham; a.tw = p->fft.d_tw; a.perm = p->fft.d_perm; a.unpackA = p->fft.d_unpackA;
  a.fbank = p->d_fbank; a.band_lo = p->d_lo; a.band_hi = p->d_hi; a.feats = b->feats; a.cmvn = b->cmvn_acc;
  a.J = p->prm.nfilters; a.power = p->prm.power;
  if (b->n_frames > 2147483647LL) return set_err(FDLP_E_ARG, "too many frames in one batch%s", "");
  const dim3 grid((unsigned)b->n_frames), block(128);
  if (b->wav_dtype == FDLP_WAV_INT16) mel_frames_kernel<int16_t><<<grid, block, p->smem, st>>>(a);
  else if (b->wav_dtype == FDLP_WAV_FLOAT32) mel_frames_kernel<float><<<grid, block, p->smem, st>>>(a);
  else return set_err(FDLP_E_ARG, "bad wav dtype%s", "");
  FDLP_CUDA(cudaGetLastError());
  return mel_cmvn(p, b, st);
}

// ================================================================ MFCC
struct mfcc_plan {
  mfcc_params prm;
  int flen = 0, shift = 0, P = 0, nb = 0, nbp = 0, used = 0;
  float* d_ham = nullptr;
  float2* d_tw = nullptr;
  float* d_weff = nullptr;
  double* d_dctb = nullptr;
  int32_t* d_frame_utt = nullptr;
  float* d_tmp = nullptr;      // unspliced features when a context is requested
  int64_t cap_frames = 0, cap_tmp = 0;
  double* d_alpha = nullptr;
  int cap_utt = 0;
  size_t smem = 0;
  // tensor-core path (csrc/mfcc_tc.cuh): nfft = 1024 (513-point transform), 320-sample frames, <= 32 bands
  bool tc_ok = false;   // the configuration fits the tensor-core kernel and its tables are built
  bool tc_on = true;    // mfcc_plan_set_tensor_cores / FDLP_MFCC_TC=0
  float *d_bh = nullptr, *d_bl = nullptr, *d_wblk = nullptr, *d_w0 = nullptr;
  int32_t* d_blk_off = nullptr;
  uint8_t* d_blk = nullptr;
  int tc_nblk = 0;
  double* d_dctb32 = nullptr;
  float *d_ah = nullptr, *d_al = nullptr, *d_dc = nullptr;
  int64_t cap_tc = 0;
  CUtensorMap tm_bh, tm_bl;
  int sms = 148;
};

static void mfcc_free(mfcc_plan* p) {
  if (!p) return;
  cudaFree(p->d_ham); cudaFree(p->d_tw); cudaFree(p->d_weff); cudaFree(p->d_dctb); cudaFree(p->d_frame_utt);
  cudaFree(p->d_tmp); cudaFree(p->d_alpha);
  cudaFree(p->d_bh); cudaFree(p->d_bl); cudaFree(p->d_wblk); cudaFree(p->d_blk); cudaFree(p->d_blk_off); cudaFree(p->d_w0); cudaFree(p->d_dctb32);
  cudaFree(p->d_ah); cudaFree(p->d_al); cudaFree(p->d_dc);
  delete p;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn tensor_map_encoder() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    return (EncodeTiledFn)f;
  }();
  return fn;
}
// float32 matrix [rows, 320] row-major -> boxes of {32 floats, box_rows rows} with the 128-byte swizzle
static bool mfcc_tmap(CUtensorMap* tm, const float* ptr, uint64_t rows, uint32_t box_rows) {
  EncodeTiledFn fn = tensor_map_encoder();
  if (!fn) return false;
  cuuint64_t dims[2] = {(cuuint64_t)tc::kK, rows};
  cuuint64_t strides[1] = {(cuuint64_t)tc::kK * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)tc::kBK, box_rows};
  cuuint32_t es[2] = {1, 1};
  return fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, tc::kBK == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// tables of the tensor-core path: B = cos | -sin of bins 1..256 cut into a TF32 part and the remainder, the folded
// filterbank transposed (bin-major, 32 bands), the DCT basis padded to 32 bands
static int mfcc_tc_setup(mfcc_plan* p, const std::vector<float>& weff, const std::vector<double>& dctb) {
  const int J = p->prm.nfilters;
  if (p->P != 2 * tc::kUnits * tc::kBinsUnit + 1 || p->used != tc::kK || J > tc::kJP) return 0;
  const int ncol = tc::kUnits * tc::kNUnit;
  std::vector<float> bh((size_t)ncol * tc::kK), bl((size_t)ncol * tc::kK);
  for (int col = 0; col < ncol; ++col) {
    const int u = col / tc::kNUnit, part = (col % tc::kNUnit) / tc::kBinsUnit, i = col % tc::kBinsUnit;
    const int k = 1 + u * tc::kBinsUnit + i;
    for (int n = 0; n < tc::kK; ++n) {
      const double ang = 2.0 * M_PI * (double)(((long long)n * k) % p->P) / (double)p->P;
      const float v = (float)(part == 0 ? cos(ang) : -sin(ang));
      uint32_t bits;
      memcpy(&bits, &v, 4);
      bits &= 0xffffe000u;
      float hi;
      memcpy(&hi, &bits, 4);
      bh[(size_t)col * tc::kK + n] = hi;
      bl[(size_t)col * tc::kK + n] = v - hi;
    }
  }
  // folded filterbank in 32-bin chunks: the (chunk, band) blocks with a non-zero weight, and which bands touch a chunk
  if (p->prm.ncep > tc::kMaxCep) return 0;
  std::vector<float> w0(tc::kJP, 0.f), wblk;
  std::vector<uint8_t> blk_band;
  std::vector<int32_t> blk_off(tc::kChunks + 1, 0);
  for (int j = 0; j < J; ++j) w0[j] = weff[(size_t)j * p->nbp];
  for (int ch = 0; ch < tc::kChunks; ++ch) {
    blk_off[ch] = (int32_t)blk_band.size();
    for (int j = 0; j < J; ++j) {
      float w[32];
      bool any = false;
      for (int i = 0; i < 32; ++i) {
        w[i] = weff[(size_t)j * p->nbp + (1 + ch * 32 + i)];
        any = any || w[i] != 0.f;
      }
      if (!any) continue;
      if ((int)blk_band.size() >= tc::kMaxBlk) return 0;  // a denser filterbank than the table holds: FP32-pipe kernel
      blk_band.push_back((uint8_t)j);
      wblk.insert(wblk.end(), w, w + 32);
    }
  }
  blk_off[tc::kChunks] = (int32_t)blk_band.size();
  p->tc_nblk = (int)blk_band.size();
  if (wblk.empty()) { wblk.assign(32, 0.f); blk_band.push_back(0); }
  std::vector<double> d32((size_t)p->prm.ncep * tc::kJP, 0.0);
  for (int c = 0; c < p->prm.ncep; ++c)
    for (int j = 0; j < J; ++j) d32[(size_t)c * tc::kJP + j] = dctb[(size_t)c * J + j];
  int rc;
  if ((rc = upload(&p->d_bh, bh)) || (rc = upload(&p->d_bl, bl)) || (rc = upload(&p->d_wblk, wblk)) ||
      (rc = upload(&p->d_blk, blk_band)) || (rc = upload(&p->d_blk_off, blk_off)) || (rc = upload(&p->d_w0, w0)) ||
      (rc = upload(&p->d_dctb32, d32)))
    return rc;
  if (!mfcc_tmap(&p->tm_bh, p->d_bh, ncol, tc::kNUnit / tc::kCluster) || !mfcc_tmap(&p->tm_bl, p->d_bl, ncol, tc::kNUnit / tc::kCluster))
    return 0;
  if (cudaFuncSetAttribute(tc::mfcc_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::kSmemBytes) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  p->tc_ok = true;
  return 0;
}

extern "C" int mfcc_plan_create(const mfcc_params* prm, mfcc_plan** out) {
  if (!prm || !out) return set_err(FDLP_E_ARG, "null argument%s", "");
  *out = nullptr;
  FDLP_CHECK_SIZE(prm, mfcc_params);
  if (prm->nfilters < 1 || prm->nfft < 8 || (prm->nfft & 1) || prm->frate < 1 || !(prm->fduration > 0) ||
      prm->ncep < 1 || prm->ncep > prm->nfilters)
    return set_err(FDLP_E_ARG, "invalid mfcc_params%s", "");
  mfcc_plan* p = new (std::nothrow) mfcc_plan();
  if (!p) return set_err(FDLP_E_NOMEM, "out of host memory%s", "");
  p->prm = *prm;
  p->flen = (int)(prm->srate * prm->fduration);       // features.py:134
  p->shift = (int)(prm->srate / (double)prm->frate);  // features.py:135
  if (p->flen < 1 || p->shift < 1) { mfcc_free(p); return set_err(FDLP_E_ARG, "empty frames%s", ""); }
  const int J = prm->nfilters;
  p->P = prm->nfft / 2 + 1;                 // computeMfccFeatures.py:125  fft(time_frames, int(nfft / 2 + 1))
  p->nb = p->P / 2 + 1;                     // |X[P-k]| = |X[k]|
  p->nbp = p->nb | 1;
  p->used = std::min(p->flen, p->P);
  const std::vector<double> fb = make_mel_fbank(J, prm->nfft, prm->srate, 1.0);  // [J, P]  (:72)
  std::vector<float> weff((size_t)J * p->nbp, 0.f);
  for (int j = 0; j < J; ++j)
    for (int k = 0; k < p->nb; ++k) {
      double w = fb[(size_t)j * p->P + k];
      if (k >= 1 && p->P - k != k) w += fb[(size_t)j * p->P + (p->P - k)];
      weff[(size_t)j * p->nbp + k] = (float)w;
    }
  const std::vector<double> hw = np_hamming(p->flen);
  std::vector<float> ham(p->used);
  for (int i = 0; i < p->used; ++i) ham[i] = (float)(hw[i] / 32768.0);   // signal / np.power(2, 15)  (:113)
  std::vector<float2> tw(p->P);
  for (int m = 0; m < p->P; ++m) {
    const double ang = -2.0 * M_PI * (double)m / (double)p->P;
    tw[m] = make_float2((float)cos(ang), (float)sin(ang));
  }
  std::vector<double> dctb((size_t)prm->ncep * J);   // scipy.fftpack.dct type II: y[c] = 2 sum_j x[j] cos(pi c (2j+1) / (2J))
  for (int c = 0; c < prm->ncep; ++c)
    for (int j = 0; j < J; ++j) dctb[(size_t)c * J + j] = 2.0 * cos(M_PI * c * (2.0 * j + 1.0) / (2.0 * J));
  int rc;
  if ((rc = upload(&p->d_ham, ham)) || (rc = upload(&p->d_tw, tw)) || (rc = upload(&p->d_weff, weff)) ||
      (rc = upload(&p->d_dctb, dctb))) {
    mfcc_free(p);
    return rc;
  }
  p->smem = mfcc_smem_bytes(p->used, p->P, p->nbp, J);
  int dev = 0, max_smem = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (max_smem <= 0) max_smem = 227 * 1024;
  if (p->smem + 2048 > (size_t)max_smem) { mfcc_free(p); return set_err(FDLP_E_UNSUPPORTED, "nfft too large for the MFCC kernel%s", ""); }
  cudaError_t e = raise_smem_limit(mfcc_frames_kernel<int16_t>, p->smem);
  if (e == cudaSuccess) e = raise_smem_limit(mfcc_frames_kernel<float>, p->smem);
  if (e != cudaSuccess) { mfcc_free(p); return set_err(FDLP_E_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e)); }
  cudaDeviceGetAttribute(&p->sms, cudaDevAttrMultiProcessorCount, dev);
  if ((rc = mfcc_tc_setup(p, weff, dctb))) { mfcc_free(p); return rc; }
  *out = p;
  return FDLP_OK;
}

extern "C" int mfcc_plan_destroy(mfcc_plan* p) { mfcc_free(p); return FDLP_OK; }

extern "C" int mfcc_plan_set_tensor_cores(mfcc_plan* p, int32_t enable) {
  if (!p) return set_err(FDLP_E_ARG, "null plan%s", "");
  p->tc_on = enable != 0;
  return (p->tc_ok && p->tc_on) ? 1 : 0;
}

extern "C" int mfcc_forward(mfcc_plan* p, const mfcc_batch* b, void* stream) {
  if (!p || !b) return set_err(FDLP_E_ARG, "null argument%s", "");
  FDLP_CHECK_SIZE(b, mfcc_batch);
  if (b->n_utt <= 0 || b->n_frames <= 0) return FDLP_OK;
  if (!b->wav || !b->utt_wav_off || !b->utt_frame_off || !b->feats) return set_err(FDLP_E_ARG, "null batch pointer%s", "");
  if (b->wav_dtype != FDLP_WAV_INT16 && b->wav_dtype != FDLP_WAV_FLOAT32) return set_err(FDLP_E_ARG, "bad wav dtype%s", "");
  if (b->context < 0) return set_err(FDLP_E_ARG, "negative context%s", "");
  if (b->noise && (!b->utt_noise_off || !b->utt_snr_db)) return set_err(FDLP_E_ARG, "noise mixing needs offsets and SNRs%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  if (b->n_frames > p->cap_frames) {
    cudaFree(p->d_frame_utt); p->d_frame_utt = nullptr; p->cap_frames = 0;
    FDLP_CUDA(cudaMalloc((void**)&p->d_frame_utt, (size_t)b->n_frames * sizeof(int32_t)));
    p->cap_frames = b->n_frames;
  }
  if (b->context > 0 && b->n_frames > p->cap_tmp) {
    cudaFree(p->d_tmp); p->d_tmp = nullptr; p->cap_tmp = 0;
    FDLP_CUDA(cudaMalloc((void**)&p->d_tmp, (size_t)b->n_frames * p->prm.ncep * sizeof(float)));
    p->cap_tmp = b->n_frames;
  }
  if (b->noise && b->n_utt > p->cap_utt) {
    cudaFree(p->d_alpha); p->d_alpha = nullptr; p->cap_utt = 0;
    FDLP_CUDA(cudaMalloc((void**)&p->d_alpha, (size_t)b->n_utt * sizeof(double)));
    p->cap_utt = b->n_utt;
  }
  mel_frame_map_kernel<<<b->n_utt, 128, 0, st>>>(b->utt_frame_off, b->n_utt, p->d_frame_utt);
  FDLP_CUDA(cudaGetLastError());
  if (b->noise) {  // both signals are divided by 2^15 in the reference, which leaves alpha unchanged: true energies
    NoiseArgs n;
    n.wav = b->wav; n.wav_dtype = b->wav_dtype; n.utt_wav_off = b->utt_wav_off; n.noise = b->noise;
    n.utt_noise_off = b->utt_noise_off; n.utt_snr_db = b->utt_snr_db; n.utt_alpha = p->d_alpha; n.mode = FDLP_NOISE_FLOAT;
    noise_alpha_kernel<<<b->n_utt, kNoiseThreads, 0, st>>>(n);
    FDLP_CUDA(cudaGetLastError());
  }
  MfccArgs a;
  memset(&a, 0, sizeof(a));
  a.wav = b->wav; a.wav_dtype = b->wav_dtype; a.utt_wav_off = b->utt_wav_off; a.utt_frame_off = b->utt_frame_off;
  a.frame_utt = p->d_frame_utt; a.n_frames = b->n_frames;
  a.flen = p->flen; a.shift = p->shift; a.P = p->P; a.nb = p->nb; a.nbp = p->nbp; a.used = p->used;
  a.ham = p->d_ham; a.tw = p->d_tw; a.weff = p->d_weff; a.dctb = p->d_dctb;
  a.feats = (b->context > 0) ? p->d_tmp : b->feats;
  a.J = p->prm.nfilters; a.ncep = p->prm.ncep;
  a.noise = b->noise; a.utt_noise_off = b->utt_noise_off; a.utt_alpha = p->d_alpha;
  const long long tiles = (b->n_frames + kMfccFT - 1) / kMfccFT;
  if (tiles > 2147483647LL) return set_err(FDLP_E_ARG, "too many frames in one batch%s", "");
  static const bool env_tc = !(getenv("FDLP_MFCC_TC") && atoi(getenv("FDLP_MFCC_TC")) == 0);  // FDLP_MFCC_TC=0: FFMA kernel
  if (p->tc_ok && p->tc_on && env_tc) {
    // tensor-core path: frames -> TF32 planes, then the GEMM kernel (csrc/mfcc_tc.cuh)
    if (b->n_frames > p->cap_tc) {
      cudaFree(p->d_ah); cudaFree(p->d_al); cudaFree(p->d_dc);
      p->d_ah = p->d_al = p->d_dc = nullptr; p->cap_tc = 0;
      FDLP_CUDA(cudaMalloc((void**)&p->d_ah, (size_t)b->n_frames * tc::kK * sizeof(float)));
      FDLP_CUDA(cudaMalloc((void**)&p->d_al, (size_t)b->n_frames * tc::kK * sizeof(float)));
      FDLP_CUDA(cudaMalloc((void**)&p->d_dc, (size_t)b->n_frames * sizeof(float)));
      p->cap_tc = b->n_frames;
    }
    tc::PrepArgs pa;
    pa.wav = b->wav; pa.utt_wav_off = b->utt_wav_off; pa.utt_frame_off = b->utt_frame_off; pa.frame_utt = p->d_frame_utt;
    pa.n_frames = b->n_frames; pa.flen = p->flen; pa.shift = p->shift; pa.ham = p->d_ham;
    pa.ah = p->d_ah; pa.al = p->d_al; pa.dc = p->d_dc;
    pa.noise = b->noise; pa.utt_noise_off = b->utt_noise_off; pa.utt_alpha = p->d_alpha;
    const unsigned pblocks = (unsigned)((b->n_frames + 7) / 8);
    if (b->wav_dtype == FDLP_WAV_INT16) tc::mfcc_tc_prep_kernel<int16_t><<<pblocks, 256, 0, st>>>(pa);
    else tc::mfcc_tc_prep_kernel<float><<<pblocks, 256, 0, st>>>(pa);
    FDLP_CUDA(cudaGetLastError());
    CUtensorMap tm_ah, tm_al;
    if (!mfcc_tmap(&tm_ah, p->d_ah, (uint64_t)b->n_frames, tc::kM) || !mfcc_tmap(&tm_al, p->d_al, (uint64_t)b->n_frames, tc::kM))
      return set_err(FDLP_E_CUDA, "cuTensorMapEncodeTiled failed%s", "");
    tc::Args ta;
    ta.dc = p->d_dc; ta.wblk = p->d_wblk; ta.blk_band = p->d_blk; ta.blk_off = p->d_blk_off; ta.n_blk = p->tc_nblk;
    ta.w0 = p->d_w0; ta.dctb = p->d_dctb32; ta.feats = a.feats;
    ta.n_frames = b->n_frames; ta.J = a.J; ta.ncep = a.ncep;
    ta.n_tiles = (int)((b->n_frames + tc::kM - 1) / tc::kM);
    const int grid = std::min((ta.n_tiles + tc::kCluster - 1) / tc::kCluster * tc::kCluster, p->sms / tc::kCluster * tc::kCluster);
    tc::mfcc_tc_kernel<<<grid, tc::kThreads, tc::kSmemBytes, st>>>(tm_ah, tm_al, p->tm_bh, p->tm_bl, ta);
    FDLP_CUDA(cudaGetLastError());
#ifdef FDLP_X_TC_PROF
    {
      long long h[8];
      cudaStreamSynchronize(st);
      cudaMemcpyFromSymbol(h, tc::g_tc_prof, sizeof(h));
      fprintf(stderr, "tc prof (cycles summed over %d CTAs): producer wait empty %lld, mma wait tmem_empty %lld, mma wait full %lld, "
              "epilogue wait tmem_full %lld, cta total %lld; epilogue: dc + log10 %lld, mags + filterbank %lld, dc + log10 + dct + store %lld\n", grid, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
      memset(h, 0, sizeof(h));
      cudaMemcpyToSymbol(tc::g_tc_prof, h, sizeof(h));
    }
#endif
  } else {
    const int threads = std::min(288, (p->nb + 31) / 32 * 32);
    if (b->wav_dtype == FDLP_WAV_INT16) mfcc_frames_kernel<int16_t><<<(unsigned)tiles, threads, p->smem, st>>>(a);
    else mfcc_frames_kernel<float><<<(unsigned)tiles, threads, p->smem, st>>>(a);
    FDLP_CUDA(cudaGetLastError());
  }
  if (b->context > 0) {
    SpliceArgs s;
    s.in = p->d_tmp; s.out = b->feats; s.utt_frame_off = b->utt_frame_off; s.frame_utt = p->d_frame_utt;
    s.n_frames = b->n_frames; s.D = p->prm.ncep; s.context = b->context;
    splice_kernel<<<(unsigned)std::min<long long>((b->n_frames + 31) / 32, 148LL * 32), 256, 0, st>>>(s);
    FDLP_CUDA(cudaGetLastError());
  }
  return FDLP_OK;
}

// ================================================================ modulation spectrum
extern "C" int fdlp_modspec_forward(fdlp_plan* p, const fdlp_batch* b, const int32_t* sel_idx, const double* sel_scale,
                                    int32_t feat_len, int32_t absolute, float* out, void* stream) {
  if (!p || !b || feat_len < 0) return set_err(FDLP_E_ARG, "bad argument%s", "");
  if (!p->prm.cepstrum_mode) return set_err(FDLP_E_ARG, "the plan was not created with cepstrum_mode = 1%s", "");
  if (b->n_utt <= 0 || b->n_seg <= 0 || feat_len == 0) return FDLP_OK;
  if (!b->wav || !b->utt_wav_off || !b->utt_seg_off || !b->seg_utt || !sel_idx || !sel_scale || !out)
    return set_err(FDLP_E_ARG, "null batch pointer%s", "");
  if (b->noise_mode != FDLP_NOISE_NONE) return set_err(FDLP_E_ARG, "noise mixing is not part of this path%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  int rc = ensure_scratch(p, b->n_seg, b->n_utt);
  if (rc) return rc;
  const size_t tasks = (size_t)b->n_seg * p->prm.nfilters;
  if (tasks > p->cap_cep) {
    cudaFree(p->d_cep); p->d_cep = nullptr; p->cap_cep = 0;
    FDLP_CUDA(cudaMalloc((void**)&p->d_cep, tasks * p->g.L * sizeof(double)));
    p->cap_cep = tasks;
  }
  if ((rc = launch_dct(p, b, p->d_dct, p->ld, p->pad, st))) return rc;
  if ((rc = launch_autocorr(p, p->d_dct, b->n_seg, p->d_r, st))) return rc;
  if ((rc = launch_lpc(p, p->d_r, b->n_seg, nullptr, nullptr, p->d_cep, nullptr, b->status, b->seg_utt, st))) return rc;
  ModspecGatherArgs g;
  g.cep = p->d_cep; g.sel_idx = sel_idx; g.sel_scale = sel_scale; g.out = out;
  g.n_tasks = (int64_t)tasks; g.L = p->g.L; g.feat_len = feat_len; g.absolute = absolute;
  const long long blocks = std::min<long long>(((long long)tasks * feat_len + 255) / 256, 148LL * 16);
  modspec_gather_kernel<<<(unsigned)blocks, 256, 0, st>>>(g);
  FDLP_CUDA(cudaGetLastError());
  return FDLP_OK;
}

// ---- complex-modulation path (computeModulationSpectrum.py --complex_modulation) ----
extern "C" int fdlp_stage_frames(fdlp_plan* p, const fdlp_batch* b, double* frames_out, void* stream) {
  int rc = check_batch(p, b);
  if (rc) return rc;
  if (!frames_out) return set_err(FDLP_E_ARG, "null output%s", "");
  if (b->noise_mode != FDLP_NOISE_NONE) return set_err(FDLP_E_ARG, "noise mixing is not part of this path%s", "");
  if (b->n_seg == 0) return FDLP_OK;
  FramesArgs a;
  a.wav = b->wav; a.utt_wav_off = b->utt_wav_off; a.utt_seg_off = b->utt_seg_off; a.seg_utt = b->seg_utt;
  a.N = p->g.N; a.S = p->g.S; a.out = frames_out; a.n_seg = b->n_seg;
  a.a0 = p->prm.window_kind == FDLP_WINDOW_HANNING ? 0.5 : p->prm.window_kind == FDLP_WINDOW_SQUARE ? 1.0 : 0.54;
  a.a1 = p->prm.window_kind == FDLP_WINDOW_HANNING ? 0.5 : p->prm.window_kind == FDLP_WINDOW_SQUARE ? 0.0 : 0.46;
  const unsigned blocks = (unsigned)std::min<long long>(b->n_seg, 148LL * 16);
  cudaStream_t st = (cudaStream_t)stream;
  if (b->wav_dtype == FDLP_WAV_INT16) window_frames_kernel<int16_t><<<blocks, 256, 0, st>>>(a);
  else if (b->wav_dtype == FDLP_WAV_FLOAT32) window_frames_kernel<float><<<blocks, 256, 0, st>>>(a);
  else window_frames_kernel<double><<<blocks, 256, 0, st>>>(a);
  FDLP_CUDA(cudaGetLastError());
  return FDLP_OK;
}

extern "C" int fdlp_modspec_complex_forward(const double* spec, int64_t spec_ld, int64_t n_seg, const int32_t* seg_utt,
                                       int32_t M, const double* fbank, const int32_t* row_lo, const int32_t* row_hi,
                                       int32_t max_width, int32_t nfilters, int32_t order, int32_t coeff_n,
                                       const int32_t* sel_idx, const double* sel_scale, int32_t n_sel,
                                       int32_t absolute, float* out, int32_t* status, void* stream) {
  if (n_seg < 0 || M < 4 || spec_ld < M || nfilters < 1 || n_sel < 0 || max_width < 1 || max_width > M)
    return set_err(FDLP_E_ARG, "bad size%s", "");
  if (order < 1 || order + 2 > M || order > 4096 || coeff_n < 2 || coeff_n > 4096)
    return set_err(FDLP_E_ARG, "need 1 <= order <= M - 2 and 2 <= coeff_n (order %s%ld)", "", (long)order);
  if (n_seg == 0 || n_sel == 0) return FDLP_OK;
  if (!spec || !seg_utt || !fbank || !row_lo || !row_hi || !sel_idx || !sel_scale || !out)
    return set_err(FDLP_E_ARG, "null pointer%s", "");
  CplxModspecArgs a;
  a.spec = reinterpret_cast<const double2*>(spec); a.ld = spec_ld; a.fbank = fbank; a.row_lo = row_lo; a.row_hi = row_hi;
  a.M = M; a.J = nfilters; a.order = order; a.lim = coeff_n; a.n_sel = n_sel; a.absolute = absolute ? 1 : 0;
  a.max_width = max_width; a.sel_idx = sel_idx; a.sel_scale = sel_scale; a.out = out; a.status = status;
  a.seg_utt = seg_utt; a.n_tasks = n_seg * (int64_t)nfilters;
  const size_t smem = sizeof(double2) * ((size_t)max_width + (order + 2) + (std::max(order, coeff_n) + 2) + coeff_n);
  if (smem > 227 * 1024) return set_err(FDLP_E_UNSUPPORTED, "band of %s%ld taps does not fit in shared memory", "", (long)max_width);
  static std::mutex mu;
  static size_t attr_smem[64] = {0};
  int dev = 0;
  FDLP_CUDA(cudaGetDevice(&dev));
  if (smem > 48 * 1024) {
    std::lock_guard<std::mutex> lk(mu);
    if (dev < 64 && attr_smem[dev] < smem) {
      FDLP_CUDA(cudaFuncSetAttribute(modspec_complex_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(227 * 1024)));
      attr_smem[dev] = 227 * 1024;
    }
  }
  const int threads = std::max(64, (order + 2 + 31) / 32 * 32);
  if (threads > 1024) return set_err(FDLP_E_UNSUPPORTED, "order %s%ld needs more than 1024 lags per CTA", "", (long)order);
  const unsigned blocks = (unsigned)std::min<long long>(a.n_tasks, 148LL * 32);
  modspec_complex_kernel<<<blocks, threads, smem, (cudaStream_t)stream>>>(a);
  FDLP_CUDA(cudaGetLastError());
  return FDLP_OK;
}

// ================================================================ host-side WAV ingest (CLI pipeline)
extern "C" int fdlp_wav_probe(const char* const* paths, int32_t n, int32_t threads, int64_t* info) {
  if (n < 0 || (n > 0 && (!paths || !info))) return set_err(FDLP_E_ARG, "bad argument%s", "");
  run_indexed(n, threads, [&](int i) { wav_probe_one(paths[i], info + 4 * (size_t)i); });
  return FDLP_OK;
}

extern "C" int fdlp_wav_read(const char* const* paths, const int64_t* data_off, const int64_t* nbytes, const int64_t* dst_off,
                             void* dst_base, int32_t n, int32_t threads, int32_t* status) {
  if (n < 0 || (n > 0 && (!paths || !data_off || !nbytes || !dst_off || !dst_base || !status)))
    return set_err(FDLP_E_ARG, "bad argument%s", "");
  run_indexed(n, threads, [&](int i) {
    status[i] = wav_read_one(paths[i], data_off[i], nbytes[i], static_cast<unsigned char*>(dst_base) + dst_off[i]);
  });
  return FDLP_OK;
}

// ================================================================ 'diff' augmentation
extern "C" int fdlp_diff_filter(const void* wav, int32_t wav_dtype, const int64_t* utt_wav_off, int32_t n_utt,
                                int64_t max_len, float* out, void* stream) {
  if (n_utt < 0 || max_len < 0) return set_err(FDLP_E_ARG, "bad argument%s", "");
  if (n_utt == 0 || max_len == 0) return FDLP_OK;
  if (!wav || !utt_wav_off || !out) return set_err(FDLP_E_ARG, "null argument%s", "");
  if (wav_dtype != FDLP_WAV_INT16 && wav_dtype != FDLP_WAV_FLOAT32) return set_err(FDLP_E_ARG, "bad wav dtype%s", "");
  if (n_utt > 65535) return set_err(FDLP_E_ARG, "too many utterances in one call%s", "");
  DiffArgs a;
  a.wav = wav; a.wav_dtype = wav_dtype; a.utt_wav_off = utt_wav_off; a.n_utt = n_utt; a.out = out;
  const long long bx = std::min<long long>((max_len + 255) / 256, 1024);
  diff_fir_kernel<<<dim3((unsigned)bx, (unsigned)n_utt), 256, 0, (cudaStream_t)stream>>>(a);
  FDLP_CUDA(cudaGetLastError());
  return FDLP_OK;
}

extern "C" int fdlp_noise_mix(const void* wav, int32_t wav_dtype, const int64_t* utt_wav_off, int32_t n_utt,
                              int64_t max_len, const int16_t* noise, const int64_t* utt_noise_off,
                              const double* utt_snr_db, int32_t noise_mode, double* utt_alpha, float* out, void* stream) {
  if (n_utt < 0 || max_len < 0) return set_err(FDLP_E_ARG, "bad argument%s", "");
  if (n_utt == 0 || max_len == 0) return FDLP_OK;
  if (!wav || !utt_wav_off || !noise || !utt_noise_off || !utt_snr_db || !utt_alpha || !out)
    return set_err(FDLP_E_ARG, "null argument%s", "");
  if (wav_dtype != FDLP_WAV_INT16 && wav_dtype != FDLP_WAV_FLOAT32) return set_err(FDLP_E_ARG, "bad wav dtype%s", "");
  if (noise_mode != FDLP_NOISE_INT16_WRAP && noise_mode != FDLP_NOISE_FLOAT) return set_err(FDLP_E_ARG, "bad noise mode%s", "");
  if (n_utt > 65535) return set_err(FDLP_E_ARG, "too many utterances in one call%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  NoiseArgs a;
  a.wav = wav; a.wav_dtype = wav_dtype; a.utt_wav_off = utt_wav_off; a.noise = noise; a.utt_noise_off = utt_noise_off;
  a.utt_snr_db = utt_snr_db; a.utt_alpha = utt_alpha; a.mode = noise_mode;
  noise_alpha_kernel<<<n_utt, kNoiseThreads, 0, st>>>(a);
  FDLP_CUDA(cudaGetLastError());
  NoiseMixArgs m;
  m.wav = wav; m.wav_dtype = wav_dtype; m.utt_wav_off = utt_wav_off; m.noise = noise; m.utt_noise_off = utt_noise_off;
  m.utt_alpha = utt_alpha; m.out = out;
  const long long bx = std::min<long long>((max_len + 255) / 256, 1024);
  noise_mix_kernel<<<dim3((unsigned)bx, (unsigned)n_utt), 256, 0, st>>>(m);
  FDLP_CUDA(cudaGetLastError());
  return FDLP_OK;
}

// ================================================================ --add_reverb (features.py:110-115)
struct fdlp_reverb_plan {
  int R = 0, B = 0, F = 0, logF = 0, P = 0;
  int conv_m = 1;  // block pairs per convolution CTA (2 halves the spectrum reads but holds accumulators in
                   // registers: one CTA per SM instead of three, slower at F = 4096)
  size_t smem = 0;
  double2* d_tw = nullptr;
  double2* d_H = nullptr;
  // scratch, grown on demand
  int cap_blk = 0, cap_utt = 0;
  double2* d_X = nullptr;
  double2* d_O = nullptr;
  double* d_full = nullptr;
  double2* d_C = nullptr;
  double* d_pmax = nullptr;
  int32_t* d_parg = nullptr;
  double* d_alpha = nullptr;
  bool profiling = false;
  cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
};

static void reverb_free(fdlp_reverb_plan* p) {
  if (!p) return;
  cudaFree(p->d_tw); cudaFree(p->d_H); cudaFree(p->d_X); cudaFree(p->d_O); cudaFree(p->d_full); cudaFree(p->d_C);
  cudaFree(p->d_pmax); cudaFree(p->d_parg); cudaFree(p->d_alpha);
  for (auto& e : p->ev) if (e) cudaEventDestroy(e);
  delete p;
}

static void reverb_fill(const fdlp_reverb_plan* p, RvArgs& a) {
  memset(&a, 0, sizeof(a));
  a.R = p->R; a.B = p->B; a.F = p->F; a.logF = p->logF; a.P = p->P; a.tw = p->d_tw; a.H = p->d_H;
  a.X = p->d_X; a.O = p->d_O; a.C = p->d_C; a.full = p->d_full; a.part_max = p->d_pmax; a.part_arg = p->d_parg;
}

extern "C" int fdlp_reverb_plan_create(const double* rir, int32_t R, int32_t block, fdlp_reverb_plan** out) {
  if (!rir || !out || R < 1) return set_err(FDLP_E_ARG, "invalid impulse response%s", "");
  if (block != 0 && (block < 64 || block > 4096 || (block & (block - 1))))
    return set_err(FDLP_E_ARG, "block must be 0 (auto) or a power of two in 64..4096%s", "");
  fdlp_reverb_plan* p = new (std::nothrow) fdlp_reverb_plan();
  if (!p) return set_err(FDLP_E_NOMEM, "out of host memory%s", "");
  p->R = R;
  int B = block;
  if (B == 0) { B = 64; while (8 * B < R && B < 2048) B <<= 1; }  // about 8 partitions; 2048 measured best at R = 16000
  p->B = B; p->F = 2 * B; p->P = (R + B - 1) / B;
  p->logF = 0;
  while ((1 << p->logF) < p->F) ++p->logF;
  const int na = p->F >> 6;
  std::vector<double2> tw(na + 64);
  for (int a = 0; a < na; ++a) {
    const double th = -2.0 * M_PI * (64.0 * a) / p->F;
    tw[a] = make_double2(cos(th), sin(th));
  }
  for (int b = 0; b < 64; ++b) {
    const double th = -2.0 * M_PI * b / p->F;
    tw[na + b] = make_double2(cos(th), sin(th));
  }
  p->smem = (size_t)(p->F + (p->F >> 6)) * sizeof(double2) + tw.size() * sizeof(double2);  // rows of 64, pitch 65
  int rc = upload(&p->d_tw, tw);
  if (rc) { reverb_free(p); return rc; }
  cudaError_t e = raise_smem_limit(rv_fwd_kernel<int16_t>, p->smem);
  if (e == cudaSuccess) e = raise_smem_limit(rv_fwd_kernel<float>, p->smem);
  if (const char* m = getenv("FDLP_RV_CONVM")) p->conv_m = atoi(m) == 2 ? 2 : 1;
  if (e == cudaSuccess) e = raise_smem_limit(rv_conv_kernel<1, 1>, p->smem);
#define RV_CONV_SMEM(BINS) if (e == cudaSuccess) e = raise_smem_limit(rv_conv_kernel<BINS, 2>, p->smem);
  RV_CONV_SMEM(1) RV_CONV_SMEM(2) RV_CONV_SMEM(4) RV_CONV_SMEM(8) RV_CONV_SMEM(16)
#undef RV_CONV_SMEM
  if (e == cudaSuccess) e = raise_smem_limit(rv_corr_kernel, p->smem);
  if (e == cudaSuccess) e = raise_smem_limit(rv_rir_kernel, p->smem);
  if (e != cudaSuccess) { reverb_free(p); return set_err(FDLP_E_CUDA, "cudaFuncSetAttribute: %s", cudaGetErrorString(e)); }
  double* d_rir = nullptr;
  std::vector<double> h(rir, rir + R);
  if ((rc = upload(&d_rir, h))) { reverb_free(p); return rc; }
  if (cudaMalloc((void**)&p->d_H, (size_t)p->P * p->F * sizeof(double2)) != cudaSuccess) {
    cudaFree(d_rir); reverb_free(p); return set_err(FDLP_E_NOMEM, "out of device memory%s", "");
  }
  RvArgs a;
  reverb_fill(p, a);
  rv_rir_kernel<<<p->P, kRvThreads, p->smem>>>(a, d_rir, p->d_H);
  e = cudaDeviceSynchronize();
  cudaFree(d_rir);
  if (e != cudaSuccess) { reverb_free(p); return set_err(FDLP_E_CUDA, "impulse-response spectra: %s", cudaGetErrorString(e)); }
  *out = p;
  return FDLP_OK;
}

extern "C" int fdlp_reverb_plan_destroy(fdlp_reverb_plan* p) {
  reverb_free(p);
  return FDLP_OK;
}

extern "C" int fdlp_reverb_plan_get_block(const fdlp_reverb_plan* p, int32_t* block, int32_t* partitions,
                                          int32_t* pairs_per_cta) {
  if (!p || !block) return set_err(FDLP_E_ARG, "null argument%s", "");
  *block = p->B;
  if (partitions) *partitions = p->P;
  if (pairs_per_cta) *pairs_per_cta = p->conv_m;
  return FDLP_OK;
}

extern "C" int fdlp_reverb_plan_set_profiling(fdlp_reverb_plan* p, int32_t enable) {
  if (!p) return set_err(FDLP_E_ARG, "null argument%s", "");
  if (enable && !p->ev[0])
    for (auto& e : p->ev) FDLP_CUDA(cudaEventCreate(&e));
  p->profiling = enable != 0;
  return FDLP_OK;
}

extern "C" int fdlp_reverb_plan_get_kernel_ms(fdlp_reverb_plan* p, float* out4) {
  if (!p || !out4 || !p->ev[0]) return set_err(FDLP_E_ARG, "profiling was not enabled%s", "");
  FDLP_CUDA(cudaEventSynchronize(p->ev[4]));
  for (int i = 0; i < 4; ++i) FDLP_CUDA(cudaEventElapsedTime(&out4[i], p->ev[i], p->ev[i + 1]));
  return FDLP_OK;
}

extern "C" int64_t fdlp_reverb_plan_scratch_bytes(const fdlp_reverb_plan* p, int32_t n_pair, int32_t n_utt) {
  if (!p) return 0;
  return ((int64_t)2 * n_pair + (int64_t)n_utt * (p->P - 1)) * p->F * (int64_t)sizeof(double2) +
         (int64_t)2 * n_pair * p->B * (int64_t)sizeof(double) + (int64_t)n_utt * p->P * p->F * (int64_t)sizeof(double2);
}

extern "C" int fdlp_reverb_forward(fdlp_reverb_plan* p, const fdlp_reverb_batch* b, void* stream) {
  if (!p || !b) return set_err(FDLP_E_ARG, "null argument%s", "");
  FDLP_CHECK_SIZE(b, fdlp_reverb_batch);
  if (b->n_utt < 0 || b->n_pair < 0 || b->n_conv < 0) return set_err(FDLP_E_ARG, "negative batch sizes%s", "");
  if (b->n_utt == 0 || b->n_pair == 0) return FDLP_OK;
  if (!b->wav || !b->utt_wav_off || !b->utt_pair_off || !b->pair_utt || !b->utt_conv_off || !b->conv_utt || !b->out)
    return set_err(FDLP_E_ARG, "null batch pointer%s", "");
  if (b->wav_dtype != FDLP_WAV_INT16 && b->wav_dtype != FDLP_WAV_FLOAT32) return set_err(FDLP_E_ARG, "bad wav dtype%s", "");
  if (b->out_dtype != FDLP_WAV_FLOAT32 && b->out_dtype != FDLP_WAV_FLOAT64) return set_err(FDLP_E_ARG, "bad output dtype%s", "");
  if (b->noise_mode != FDLP_NOISE_NONE && (!b->noise || !b->utt_noise_off || !b->utt_snr_db))
    return set_err(FDLP_E_ARG, "noise mixing requested without noise arrays%s", "");
  if (b->n_utt > 65535) return set_err(FDLP_E_ARG, "too many utterances in one call%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n_rows = (size_t)b->n_pair + (size_t)b->n_utt * (p->P - 1);
  if (b->n_pair > p->cap_blk || b->n_utt > p->cap_utt) {
    FDLP_CUDA(cudaStreamSynchronize(st));
    cudaFree(p->d_X); cudaFree(p->d_O); cudaFree(p->d_full); cudaFree(p->d_pmax); cudaFree(p->d_parg); cudaFree(p->d_alpha);
    cudaFree(p->d_C);
    p->d_X = p->d_O = p->d_C = nullptr; p->d_full = nullptr; p->d_pmax = nullptr; p->d_parg = nullptr; p->d_alpha = nullptr;
    const int np = std::max(b->n_pair, p->cap_blk), nu = std::max(b->n_utt, p->cap_utt);  // capacities only grow
    p->cap_blk = p->cap_utt = 0;
    FDLP_CUDA(cudaMalloc((void**)&p->d_X, ((size_t)np + (size_t)nu * (p->P - 1)) * p->F * sizeof(double2)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_O, (size_t)np * p->F * sizeof(double2)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_full, (size_t)2 * np * p->B * sizeof(double)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_C, (size_t)nu * p->P * p->F * sizeof(double2)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_pmax, (size_t)nu * p->P * sizeof(double)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_parg, (size_t)nu * p->P * sizeof(int32_t)));
    FDLP_CUDA(cudaMalloc((void**)&p->d_alpha, (size_t)nu * sizeof(double)));
    p->cap_blk = np; p->cap_utt = nu;
  }
  RvArgs a;
  reverb_fill(p, a);
  a.wav = b->wav; a.wav_dtype = b->wav_dtype; a.utt_wav_off = b->utt_wav_off;
  a.utt_pair_off = b->utt_pair_off; a.pair_utt = b->pair_utt; a.n_pair = b->n_pair;
  a.utt_conv_off = b->utt_conv_off; a.conv_utt = b->conv_utt;
  a.out = b->out; a.out_dtype = b->out_dtype; a.ind = b->ind; a.out_len = b->out_len;
  const bool prof = p->profiling;
  if (b->noise_mode != FDLP_NOISE_NONE) {
    NoiseArgs n;
    n.wav = b->wav; n.wav_dtype = b->wav_dtype; n.utt_wav_off = b->utt_wav_off; n.noise = (const int16_t*)b->noise;
    n.utt_noise_off = b->utt_noise_off; n.utt_snr_db = b->utt_snr_db; n.utt_alpha = p->d_alpha; n.mode = b->noise_mode;
    noise_alpha_kernel<<<b->n_utt, kNoiseThreads, 0, st>>>(n);
    FDLP_CUDA(cudaGetLastError());
    a.noise = (const int16_t*)b->noise; a.utt_noise_off = b->utt_noise_off; a.utt_alpha = p->d_alpha;
  }
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev[0], st));
  if (b->wav_dtype == FDLP_WAV_INT16) rv_fwd_kernel<int16_t><<<(unsigned)n_rows, kRvThreads, p->smem, st>>>(a);
  else rv_fwd_kernel<float><<<(unsigned)n_rows, kRvThreads, p->smem, st>>>(a);
  FDLP_CUDA(cudaGetLastError());
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev[1], st));
  if (p->conv_m == 1) {
    // radix-16 transforms (128 registers): 256 threads so that two CTAs share an SM; 512 at F = 8192 (one CTA fits)
    rv_conv_kernel<1, 1><<<b->n_conv, p->F <= 4096 ? 256 : kRvThreads, p->smem, st>>>(a);
  } else {
    const int bins = std::max(1, p->F / kRvThreads);
#define RV_CONV_LAUNCH(BINS) case BINS: rv_conv_kernel<BINS, 2><<<b->n_conv, kRvThreads, p->smem, st>>>(a); break;
    switch (bins) {
      RV_CONV_LAUNCH(1) RV_CONV_LAUNCH(2) RV_CONV_LAUNCH(4) RV_CONV_LAUNCH(8) RV_CONV_LAUNCH(16)
      default: return set_err(FDLP_E_UNSUPPORTED, "unsupported transform length%s", "");
    }
#undef RV_CONV_LAUNCH
  }
  FDLP_CUDA(cudaGetLastError());
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev[2], st));
  rv_corr_mac_kernel<<<dim3((unsigned)((p->F + kRvThreads - 1) / kRvThreads), (unsigned)((p->P + kRvQ - 1) / kRvQ),
                            (unsigned)b->n_utt), kRvThreads, 0, st>>>(a);
  FDLP_CUDA(cudaGetLastError());
  rv_corr_kernel<<<dim3((unsigned)p->P, (unsigned)b->n_utt), kRvThreads, p->smem, st>>>(a);
  FDLP_CUDA(cudaGetLastError());
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev[3], st));
  const long long bx = std::max<long long>(1, std::min<long long>((b->max_len + 1023) / 1024, 64));
  rv_align_kernel<<<dim3((unsigned)bx, (unsigned)b->n_utt), 256, 0, st>>>(a);
  FDLP_CUDA(cudaGetLastError());
  if (prof) FDLP_CUDA(cudaEventRecord(p->ev[4], st));
  return FDLP_OK;
}

// ================================================================ CMVN apply
extern "C" int fdlp_synth_white(int16_t* out, const int64_t* utt_wav_off, const int64_t* utt_id, int32_t n_utt,
                                int64_t max_len, uint64_t seed, int32_t scale_q16, void* stream) {
  if (n_utt <= 0) return FDLP_OK;
  if (!out || !utt_wav_off || !utt_id || max_len < 1 || n_utt > 65535) return set_err(FDLP_E_ARG, "invalid synth arguments%s", "");
  SynthArgs a;
  a.out = out; a.utt_wav_off = utt_wav_off; a.utt_id = utt_id; a.n_utt = n_utt; a.seed = seed; a.scale_q16 = scale_q16;
  const dim3 grid((unsigned)std::min<int64_t>((max_len + 1023) / 1024, 64), (unsigned)n_utt);
  synth_white_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  FDLP_CUDA(cudaGetLastError());
  return FDLP_OK;
}

extern "C" int cmvn_apply(float* feats, int64_t n_frames, int32_t dim, const double* stats, int32_t norm_vars, void* stream) {
  if (!feats || !stats || n_frames < 0 || dim < 1) return set_err(FDLP_E_ARG, "bad argument%s", "");
  if (n_frames == 0) return FDLP_OK;
  const int64_t total = n_frames * dim;
  const int64_t blocks = (total + 255) / 256;
  if (blocks > 2147483647LL) return set_err(FDLP_E_ARG, "too many elements%s", "");
  cmvn_apply_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(feats, n_frames, dim, stats, norm_vars);
  FDLP_CUDA(cudaGetLastError());
  return FDLP_OK;
}

// ================================================================ allocation-free CMVN accumulation
// cmvn_accumulate_noalloc: the CMVN partial sums of a finished feature matrix (the same two kernels mel_forward runs
// for its `cmvn_acc` argument, cmvn_sums_kernel + cmvn_reduce_kernel) for callers that cannot call cudaMalloc: CUDA graph
// capture, tight serving loops, harnesses that forbid allocation.  The per-CTA partial-sum slots live in a file-scope
// device array instead of plan scratch.  Every GPU that loads the library holds these kNoAllocBytes until exit.
constexpr int kNoAllocMaxDim = FDLP_NOALLOC_MAX_DIM;
constexpr size_t kNoAllocDoubles = (size_t)kMelCmvnCtas * 2 * (kNoAllocMaxDim + 1);  // 296 x 2 x 256 doubles = 1.2 MB
static __device__ __align__(256) double g_noalloc_cmvn_part[kNoAllocDoubles];

// In host code `g_noalloc_cmvn_part` is only a stand-in for the symbol, not a device address: the address is taken once
// per device, with that device current, through cudaGetSymbolAddress, and that pointer is what the kernels receive.
constexpr int kNoAllocMaxDevices = 64;
static double* g_noalloc_ptr[kNoAllocMaxDevices] = {};
static std::mutex g_noalloc_mutex;

static int noalloc_scratch(double** out) {
  int dev = -1;
  FDLP_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return set_err(FDLP_E_ARG, "device index beyond %s", "64");
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[dev]) {
    double* ptr = nullptr;
    FDLP_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_cmvn_part));
    if (!ptr) return set_err(FDLP_E_CUDA, "cudaGetSymbolAddress returned a null address%s", "");
    g_noalloc_ptr[dev] = ptr;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}

extern "C" int cmvn_accumulate_noalloc(const float* feats, int64_t n_frames, int32_t dim, double* stats, void* stream) {
  if (!feats || !stats || n_frames < 0 || dim < 1) return set_err(FDLP_E_ARG, "bad argument%s", "");
  if (dim > kNoAllocMaxDim)
    return set_err(FDLP_E_ARG,
                   "cmvn_accumulate_noalloc: dim = %s%ld exceeds FDLP_NOALLOC_MAX_DIM = %ld (the static scratch array); use the "
                   "cmvn_acc argument of mel_forward / fdlp_forward instead", "", (long)dim, (long)kNoAllocMaxDim);
  if (n_frames == 0) return FDLP_OK;
  double* part = nullptr;
  int rc = noalloc_scratch(&part);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int w = 2 * (dim + 1);
  CmvnSumArgs c;
  c.feats = feats; c.part = part; c.n_frames = n_frames; c.J = dim;
  c.chunk = (n_frames + kMelCmvnCtas - 1) / kMelCmvnCtas;
  const int rows_par = std::max(1, 256 / dim);
  // every one of the kMelCmvnCtas slots, all w entries of it, is written by this launch before the reduction reads it:
  // nothing of the previous call's contents survives into the result, so the array needs no clearing
  cmvn_sums_kernel<<<kMelCmvnCtas, 256, (size_t)rows_par * 2 * std::min(dim, 256) * sizeof(double), st>>>(c);
  FDLP_CUDA(cudaGetLastError());
  cmvn_reduce_kernel<<<w, 256, 0, st>>>(part, kMelCmvnCtas, w, stats);
  FDLP_CUDA(cudaGetLastError());
  return FDLP_OK;
}

// ================================================================ FP64 peak probe
namespace fdlp {
__global__ void __launch_bounds__(256) peak_dfma_kernel(double* out, double a, double b, int iters) {
  double acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = fma(acc[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) peak_dmma_kernel(double* out, double a, double b, int iters) {
  double c[16][2];
#pragma unroll
  for (int i = 0; i < 16; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) dmma884(c[i][0], c[i][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace fdlp

extern "C" int fdlp_measure_fp64_peak(double* dfma_tflops, double* dmma_tflops) {
  int dev = 0, sms = 0;
  FDLP_CUDA(cudaGetDevice(&dev));
  FDLP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int blocks = sms * 8, threads = 256, iters = 4096;
  double* out = nullptr;
  FDLP_CUDA(cudaMalloc((void**)&out, sizeof(double) * blocks * threads));
  cudaEvent_t e0, e1;
  FDLP_CUDA(cudaEventCreate(&e0));
  FDLP_CUDA(cudaEventCreate(&e1));
  double best[2] = {1e30, 1e30};
  for (int which = 0; which < 2; ++which) {
    for (int rep = 0; rep < 6; ++rep) {
      FDLP_CUDA(cudaEventRecord(e0));
      if (which == 0) peak_dfma_kernel<<<blocks, threads>>>(out, 1.0000001, 1e-9, iters);
      else peak_dmma_kernel<<<blocks, threads>>>(out, 1.0000001, 1e-9, iters);
      FDLP_CUDA(cudaEventRecord(e1));
      FDLP_CUDA(cudaEventSynchronize(e1));
      float ms = 0;
      FDLP_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      if (rep >= 2 && ms < best[which]) best[which] = ms;
    }
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
  const double threads_total = (double)blocks * threads;
  if (dfma_tflops) *dfma_tflops = threads_total * iters * 16 * 2 / (best[0] * 1e-3) / 1e12;
  if (dmma_tflops) *dmma_tflops = (threads_total / 32) * iters * 16 * 512.0 / (best[1] * 1e-3) / 1e12;
  return FDLP_OK;
}
