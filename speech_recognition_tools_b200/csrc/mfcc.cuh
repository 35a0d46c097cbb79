// Kc: MFCC features (reference: src/featgen/computeMfccFeatures.py:111-131).
//
//   signal / 2^15 [+ alpha * noise / 2^15]  ->  Hamming frames (getFrames, features.py:118-154)
//   spec = abs(fft(frames, nfft/2 + 1))             <- a P = nfft/2 + 1 POINT transform (513 for nfft = 1024): the
//                                                      frame is truncated / zero-padded to P samples and all P bins,
//                                                      the mirrored upper half included, meet the filterbank
//   mel  = log10(spec @ createFbank(nfilters, nfft, srate).T)          [F, nfilters]
//   out  = scipy dct (type II, unnormalised) along the bands, first 13 coefficients
//
// P is odd, so there is no power-of-two FFT to reuse: the transform is evaluated as a direct DFT, 2 * used * nb real
// FMAs per frame (used = min(flen, P) samples, nb = P/2 + 1 distinct magnitudes because |X[P-k]| = |X[k]| for real
// input; the filterbank columns k and P-k are folded on the host).  One CTA owns a tile of kMfccFT frames: thread k
// keeps bin k of all frames in registers, frame samples are broadcast reads of shared memory, the twiddle
// exp(-2 pi i (n k mod P) / P) comes from a P-entry shared table every 16th sample and from a recurrence in between.
// FP32 products and sums in the DFT and the filterbank (relative 1e-6), float64 from the logarithm on.
#pragma once
#include "common.cuh"

namespace fdlp {

constexpr int kMfccFT = 32;  // frames per CTA

struct MfccArgs {
  const void* wav;
  int wav_dtype;
  const int64_t* utt_wav_off;    // [n_utt+1]
  const int64_t* utt_frame_off;  // [n_utt+1]
  const int32_t* frame_utt;      // [n_frames]
  int64_t n_frames;
  int flen, shift, P, nb, nbp, used;
  const float* ham;     // [used]  hamming(flen)[i] / 2^15
  const float2* tw;     // [P]     exp(-2 pi i m / P)
  const float* weff;    // [J, nbp] folded filterbank: fb[j][k] + fb[j][P-k] (k >= 1), fb[j][0]
  const double* dctb;   // [ncep, J] 2 cos(pi c (2 j + 1) / (2 J))
  float* feats;         // [n_frames, ncep]
  int J, ncep;
  const int16_t* noise;          // or null
  const int64_t* utt_noise_off;  // [n_utt]
  const double* utt_alpha;       // [n_utt]
};

__host__ __device__ inline size_t mfcc_smem_bytes(int used, int P, int nbp, int J) {
  size_t b = (size_t)used * kMfccFT * sizeof(float);       // xs[used][FT]
  b += (size_t)P * sizeof(float2);                          // tw
  b += (size_t)kMfccFT * nbp * sizeof(float);               // mag[FT][nbp]
  b = (b + 7) & ~(size_t)7;
  b += (size_t)kMfccFT * J * sizeof(double);                // le[FT][J]
  return b;
}

template <typename S>
__global__ void __launch_bounds__(288) mfcc_frames_kernel(const MfccArgs a) {
  constexpr int FT = kMfccFT;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* xs = reinterpret_cast<float*>(smem_raw);                        // [used][FT]
  float2* tw = reinterpret_cast<float2*>(xs + (size_t)a.used * FT);      // [P]
  float* mag = reinterpret_cast<float*>(tw + a.P);                       // [FT][nbp]
  size_t off = (size_t)a.used * FT * sizeof(float) + (size_t)a.P * sizeof(float2) + (size_t)FT * a.nbp * sizeof(float);
  off = (off + 7) & ~(size_t)7;
  double* le = reinterpret_cast<double*>(smem_raw + off);               // [FT][J]
  __shared__ int64_t s_base[FT];   // sample offset of the utterance in the packed batch
  __shared__ int64_t s_T[FT];      // utterance length (0: frame beyond the batch)
  __shared__ int64_t s_q0[FT];     // first sample of the frame (before reflection)
  __shared__ int64_t s_noff[FT];
  __shared__ float s_alpha[FT];

  const int64_t fr0 = (int64_t)blockIdx.x * FT;
  if (threadIdx.x < FT) {
    const int64_t fr = fr0 + threadIdx.x;
    int64_t T = 0, base = 0, q0 = 0, noff = 0;
    float alpha = 0.f;
    if (fr < a.n_frames) {
      const int u = a.frame_utt[fr];
      const int k = (int)(fr - a.utt_frame_off[u]);
      base = a.utt_wav_off[u];
      T = a.utt_wav_off[u + 1] - base;
      const int ext = (a.flen % 2 == 0) ? a.flen / 2 - 1 : (a.flen - 1) / 2;  // features.py:137-146
      q0 = (int64_t)k * a.shift - ext;
      if (a.noise) { noff = a.utt_noise_off[u]; alpha = (float)a.utt_alpha[u]; }
    }
    s_base[threadIdx.x] = base; s_T[threadIdx.x] = T; s_q0[threadIdx.x] = q0;
    s_noff[threadIdx.x] = noff; s_alpha[threadIdx.x] = alpha;
  }
  for (int m = threadIdx.x; m < a.P; m += blockDim.x) tw[m] = __ldg(a.tw + m);
  __syncthreads();

  // ---- frames -> xs[i][f] (frame index fastest: conflict-free stores) ----
  const S* wav = reinterpret_cast<const S*>(a.wav);
  for (int idx = threadIdx.x; idx < a.used * FT; idx += blockDim.x) {
    const int f = idx % FT, i = idx / FT;
    float v = 0.f;
    const int64_t T = s_T[f];
    if (T > 0) {
      const int64_t q = (T > 1) ? reflect_index(s_q0[f] + i, T) : 0;
      float x = (float)wav[s_base[f] + q];
      if (a.noise) x = fmaf(s_alpha[f], (float)a.noise[s_noff[f] + q], x);
      v = x * __ldg(a.ham + i);
    }
    xs[idx] = v;
  }
  __syncthreads();

  // ---- direct DFT: thread k accumulates bin k of the FT frames ----
  const float4* xs4 = reinterpret_cast<const float4*>(xs);
  for (int k = threadIdx.x; k < a.nb; k += blockDim.x) {
    float re[FT], im[FT];
#pragma unroll
    for (int f = 0; f < FT; ++f) re[f] = im[f] = 0.f;
    // twiddle exp(-2 pi i n k / P): read from the table every 16th sample, advanced by w_k = exp(-2 pi i k / P) in
    // between (a gather per sample kept the shared-memory pipe as busy as the FMA pipe; 15 float32 products drift by
    // 1e-6 at most before the next table value)
    const float2 wk = tw[k];
    const int k16 = (int)((16LL * k) % a.P);
    int m = 0;  // (n * k) mod P at n = 0, 16, 32, ...
    float2 w = make_float2(1.f, 0.f);
    for (int n = 0; n < a.used; ++n) {
      if ((n & 15) == 0) {
        w = tw[m];
        m += k16;
        if (m >= a.P) m -= a.P;
      } else {
        w = make_float2(w.x * wk.x - w.y * wk.y, w.x * wk.y + w.y * wk.x);
      }
#pragma unroll
      for (int f4 = 0; f4 < FT / 4; ++f4) {
        const float4 x = xs4[n * (FT / 4) + f4];
        re[4 * f4 + 0] = fmaf(x.x, w.x, re[4 * f4 + 0]); im[4 * f4 + 0] = fmaf(x.x, w.y, im[4 * f4 + 0]);
        re[4 * f4 + 1] = fmaf(x.y, w.x, re[4 * f4 + 1]); im[4 * f4 + 1] = fmaf(x.y, w.y, im[4 * f4 + 1]);
        re[4 * f4 + 2] = fmaf(x.z, w.x, re[4 * f4 + 2]); im[4 * f4 + 2] = fmaf(x.z, w.y, im[4 * f4 + 2]);
        re[4 * f4 + 3] = fmaf(x.w, w.x, re[4 * f4 + 3]); im[4 * f4 + 3] = fmaf(x.w, w.y, im[4 * f4 + 3]);
      }
    }
#pragma unroll
    for (int f = 0; f < FT; ++f) mag[f * a.nbp + k] = sqrtf(re[f] * re[f] + im[f] * im[f]);
  }
  __syncthreads();

  // ---- folded filterbank, log10 ----
  for (int idx = threadIdx.x; idx < FT * a.J; idx += blockDim.x) {
    const int f = idx % FT, j = idx / FT;  // nbp is odd: the 32 rows of a warp sit in 32 different banks
    const float* mrow = mag + f * a.nbp;
    const float* wrow = a.weff + (size_t)j * a.nbp;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int k = 0;
    for (; k + 3 < a.nb; k += 4) {
      s0 = fmaf(__ldg(wrow + k), mrow[k], s0);
      s1 = fmaf(__ldg(wrow + k + 1), mrow[k + 1], s1);
      s2 = fmaf(__ldg(wrow + k + 2), mrow[k + 2], s2);
      s3 = fmaf(__ldg(wrow + k + 3), mrow[k + 3], s3);
    }
    for (; k < a.nb; ++k) s0 = fmaf(__ldg(wrow + k), mrow[k], s0);
    le[f * a.J + j] = log10((double)((s0 + s1) + (s2 + s3)));
  }
  __syncthreads();

  // ---- DCT-II along the bands, first ncep coefficients ----
  for (int idx = threadIdx.x; idx < FT * a.ncep; idx += blockDim.x) {
    const int c = idx % a.ncep, f = idx / a.ncep;
    if (fr0 + f >= a.n_frames) continue;
    const double* lrow = le + f * a.J;
    const double* brow = a.dctb + (size_t)c * a.J;
    double s = 0.0;
    for (int j = 0; j < a.J; ++j) s = fma(__ldg(brow + j), lrow[j], s);
    a.feats[(fr0 + f) * a.ncep + c] = (float)s;
  }
}

// spliceFeats (features.py:157-169): out[i, w * D + d] = in[i + w - context, d] (zero outside the utterance) for
// i < F - context; the reference's loop leaves the last `context` rows zero.
struct SpliceArgs {
  const float* in;               // [n_frames, D]
  float* out;                    // [n_frames, D * (2 context + 1)]
  const int64_t* utt_frame_off;  // [n_utt+1]
  const int32_t* frame_utt;      // [n_frames]
  int64_t n_frames;
  int D, context;
};

__global__ void __launch_bounds__(256) splice_kernel(const SpliceArgs a) {
  // a CTA owns kSpliceFT consecutive frames: 32-bit index arithmetic inside the tile (the first version divided a 64-bit
  // flat index three times per element and was bound by that: 0.62 ms for 487 MB of output)
  constexpr int kSpliceFT = 32;
  const int W = 2 * a.context + 1;
  const int row = W * a.D;  // output floats per frame
  __shared__ int64_t s_f0[kSpliceFT], s_F[kSpliceFT];
  for (int64_t t0 = (int64_t)blockIdx.x * kSpliceFT; t0 < a.n_frames; t0 += (int64_t)gridDim.x * kSpliceFT) {
    const int nfr = (int)((a.n_frames - t0 < kSpliceFT) ? a.n_frames - t0 : kSpliceFT);
    __syncthreads();
    if ((int)threadIdx.x < nfr) {
      const int u = a.frame_utt[t0 + threadIdx.x];
      s_f0[threadIdx.x] = a.utt_frame_off[u];
      s_F[threadIdx.x] = a.utt_frame_off[u + 1] - a.utt_frame_off[u];
    }
    __syncthreads();
    float* out = a.out + t0 * row;
    for (int li = threadIdx.x; li < nfr * row; li += blockDim.x) {
      const int lf = li / row, r = li - lf * row;
      const int w = r / a.D, d = r - w * a.D;
      const int64_t f0 = s_f0[lf], F = s_F[lf];
      const int64_t i = t0 + lf - f0, src = i + w - a.context;
      float v = 0.f;
      if (i < F - a.context && src >= 0 && src < F) v = __ldg(a.in + (f0 + src) * a.D + d);
      out[li] = v;
    }
  }
}

}  // namespace fdlp
