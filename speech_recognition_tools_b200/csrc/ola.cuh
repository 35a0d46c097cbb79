// K6: overlap-add of the per-segment envelopes into 100 Hz frames, floor + natural log, FP32 store,
// and the per-dimension CMVN partial sums.
//
// Reference: OLA        src/featgen/computeFDLPSpectrogram.py:182,207-225
//            log / clip src/featgen/computeFDLPSpectrogram.py:227  (np.log(np.clip(feats.T, a_min=1e-14)))
//            CMVN stats Kaldi compute-cmvn-stats, e2e/wsj/run_fdlp_e1.sh:280 (un-vendored: layout restated)
//
// Gather form: every output frame pulls its contributors in ascending segment order (the order the
// reference adds them), so the result is deterministic and needs no atomics.  The segment pointers
// (with the randrange(2) jitter of line 225) are computed by the host and passed in.
// A block owns one (utterance, tile of FT frames): phase 1 reads envelopes with the frame index fastest
// (coalesced along t), phase 2 writes [frame, band] rows with the band index fastest (coalesced).
#pragma once
#include "common.cuh"

namespace fdlp {

struct OlaArgs {
  const void* env;               // [n_seg, J, kk] float32 (fdlp_forward) or float64
  const int32_t* utt_seg_off;    // [n_utt+1]
  const int64_t* utt_frame_off;  // [n_utt+1]
  const int32_t* seg_ptr;        // [n_seg]
  float* feats;                  // [n_frames, J]
  double* cmvn_part;             // [cmvn_slots, 2, J+1] per-CTA partial sums (zeroed by the launcher), or null
  int J, kk, kkb2;
};

// CMVN partial sums are bit-reproducible: every CTA writes its own slot (no floating-point atomics), and
// cmvn_reduce_kernel adds the slots in a fixed order.  Slot of (utterance u, tile y): foff_u / FT + u + y, unique
// because floor(a) + ceil(b) <= floor(a + b) + 1; at most n_frames / FT + n_utt + 1 slots.
__host__ __device__ inline int64_t cmvn_slot_count(int64_t n_frames, int n_utt, int ft) { return n_frames / ft + n_utt + 1; }

constexpr int kOlaFT = 64;  // frames per block

template <typename ET>
__global__ void __launch_bounds__(256) ola_log_kernel(const OlaArgs a) {
  const ET* __restrict__ env = reinterpret_cast<const ET*>(a.env);
  extern __shared__ __align__(16) double tile[];  // [J][FT+1]
  const int u = blockIdx.x;
  const int64_t foff = a.utt_frame_off[u];
  const int Fr = (int)(a.utt_frame_off[u + 1] - foff);
  const int f0 = blockIdx.y * kOlaFT;
  if (f0 >= Fr) return;
  const int nf = min(kOlaFT, Fr - f0);
  const int s0 = a.utt_seg_off[u], nseg = a.utt_seg_off[u + 1] - s0;
  const int J = a.J, kk = a.kk, kkb2 = a.kkb2;
  constexpr int LD = kOlaFT + 1;

  // Segments that can reach this tile: the pointers increase with the segment index, so the contributors of
  // frames [f0, f0 + nf) are one contiguous run (2 or 3 segments); every thread finds it with the same scan.
  int i_lo = nseg, i_hi = 0;  // later segments (i >= 1) with ptr_i < f0 + nf and ptr_i + kk > f0
  for (int i = 1; i < nseg; ++i) {
    const int ptr = a.seg_ptr[s0 + i];
    if (ptr < f0 + nf && ptr + kk > f0) { i_lo = min(i_lo, i); i_hi = max(i_hi, i + 1); }
  }
  const int d1 = Fr < kkb2 ? Fr : kkb2;  // feats[:, 0:kkb2] += E[kkb2:]  (or feats[:, :] += E[kkb2:kkb2+Fr] when Fr < kkb2)
  for (int e = threadIdx.x; e < J * kOlaFT; e += blockDim.x) {
    const int j = e / kOlaFT, fl = e % kOlaFT;
    const int f = f0 + fl;
    double acc = 0.0;
    if (fl < nf) {
      if (f < d1 && kkb2 + f < kk) acc += (double)env[((size_t)s0 * J + j) * kk + kkb2 + f];  // segment 0 first: the reference's order
      for (int i = i_lo; i < i_hi; ++i) {
        const int t = f - a.seg_ptr[s0 + i];
        if (t >= 0 && t < kk) acc += (double)env[((size_t)(s0 + i) * J + j) * kk + t];
      }
    }
    tile[j * LD + fl] = acc;
  }
  __syncthreads();

  {
    // element e = fl * J + j for e = tid, tid + 256, ...: advance (fl, j) without a division per element
    int fl = (int)threadIdx.x / J, j = (int)threadIdx.x % J;
    const int dfl = (int)blockDim.x / J, dj = (int)blockDim.x % J;
    float* frow = a.feats + (size_t)(foff + f0) * J;
    for (; fl < nf;) {
      const double v = tile[j * LD + fl];
      // np.clip propagates NaN; fmax would not
      const double lg = isnan(v) ? v : log(v < 1e-14 ? 1e-14 : v);
      const float o = (float)lg;
      frow[(size_t)fl * J + j] = o;
      tile[j * LD + fl] = (double)o;  // keep the rounded value for the CMVN sums
      fl += dfl;
      j += dj;
      if (j >= J) { j -= J; ++fl; }
    }
  }
  if (a.cmvn_part) {
    __syncthreads();
    double* part = a.cmvn_part + (size_t)(foff / kOlaFT + u + blockIdx.y) * (2 * (J + 1));
    for (int j = threadIdx.x; j < J; j += blockDim.x) {
      double s1 = 0.0, s2 = 0.0;
      for (int fl = 0; fl < nf; ++fl) {
        const double v = tile[j * LD + fl];
        s1 += v;
        s2 = fma(v, v, s2);
      }
      part[j] = s1;
      part[(J + 1) + j] = s2;
    }
    if (threadIdx.x == 0) part[J] = (double)nf;
  }
}

// Partial CMVN sums of a finished feature matrix (mel path): CTA b owns the frames [b * chunk, (b + 1) * chunk) and
// adds them row by row, in order; slot b of `part` receives [sum x (J), frames; sum x^2 (J), 0].
struct CmvnSumArgs {
  const float* feats;  // [n_frames, J]
  double* part;        // [gridDim.x, 2, J+1]
  int64_t n_frames, chunk;
  int J;
};

__global__ void __launch_bounds__(256) cmvn_sums_kernel(const CmvnSumArgs a) {
  extern __shared__ __align__(16) double sh[];  // [rows_par][2][J]
  const int J = a.J;
  const int rows_par = blockDim.x / J > 0 ? blockDim.x / J : 1;
  const int64_t f0 = (int64_t)blockIdx.x * a.chunk;
  const int64_t f1 = f0 + a.chunk < a.n_frames ? f0 + a.chunk : a.n_frames;
  double* part = a.part + (size_t)blockIdx.x * (2 * (J + 1));
  for (int jb = 0; jb < J; jb += blockDim.x) {  // J > blockDim.x: several passes over the bands
    const int rl = (int)threadIdx.x / J, j = jb + ((J <= (int)blockDim.x) ? (int)threadIdx.x % J : (int)threadIdx.x);
    double s1 = 0.0, s2 = 0.0;
    if (rl < rows_par && j < J)
      for (int64_t f = f0 + rl; f < f1; f += rows_par) {
        const double v = (double)a.feats[f * J + j];
        s1 += v;
        s2 = fma(v, v, s2);
      }
    if (rl < rows_par && j < J) { sh[(rl * 2) * J + j - jb] = s1; sh[(rl * 2 + 1) * J + j - jb] = s2; }
    __syncthreads();
    if (rl == 0 && j < J) {
      for (int q = 1; q < rows_par; ++q) { s1 += sh[(q * 2) * J + j - jb]; s2 += sh[(q * 2 + 1) * J + j - jb]; }
      part[j] = s1;
      part[(J + 1) + j] = s2;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { part[J] = f1 > f0 ? (double)(f1 - f0) : 0.0; part[2 * J + 1] = 0.0; }
}

// cmvn[c] += sum over slots of part[slot][c], slots added in a fixed order: thread t takes slots t, t + 256, ... in
// sequence, then a fixed binary tree over the 256 partial sums.  One CTA per column c of the [2, J+1] layout.
__global__ void __launch_bounds__(256) cmvn_reduce_kernel(const double* part, int64_t n_slots, int width, double* cmvn) {
  __shared__ double sh[256];
  const int c = blockIdx.x;
  double s = 0.0;
  for (int64_t q = threadIdx.x; q < n_slots; q += 256) s += part[q * width + c];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) cmvn[c] += sh[0];
}

// modulation-spectrum output (computeModulationSpectrum.py:176-191): selected cepstral coefficients per band
struct ModspecGatherArgs {
  const double* cep;       // [n_tasks, L]
  const int32_t* sel_idx;  // [feat_len]
  const double* sel_scale; // [feat_len]
  float* out;              // [n_tasks, feat_len]  == [n_seg, J * feat_len]
  int64_t n_tasks;
  int L, feat_len, absolute;
};

__global__ void __launch_bounds__(256) modspec_gather_kernel(const ModspecGatherArgs a) {
  const int64_t total = a.n_tasks * a.feat_len;
  for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t task = e / a.feat_len;
    const int q = (int)(e - task * a.feat_len);
    double v = a.cep[task * a.L + a.sel_idx[q]] * a.sel_scale[q];
    if (a.absolute) v = fabs(v);
    a.out[e] = (float)v;
  }
}

// Kaldi apply-cmvn (--norm-vars) with global stats: x <- (x - mean) / sqrt(var)
__global__ void cmvn_apply_kernel(float* feats, int64_t n, int dim, const double* stats, int norm_vars) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * dim) return;
  const int d = (int)(i % dim);
  const double cnt = stats[dim];
  const double mean = stats[d] / cnt;
  double v = (double)feats[i] - mean;
  if (norm_vars) {
    double var = stats[(dim + 1) + d] / cnt - mean * mean;
    if (var < 1e-20) var = 1e-20;
    v *= rsqrt(var);
  }
  feats[i] = (float)v;
}

}  // namespace fdlp
