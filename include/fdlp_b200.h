/*
 * fdlp_b200.h -- C ABI of libfdlp_b200.so: the B200 (sm_100a) FDLP-spectrogram / mel-spectrogram
 * feature-extraction hot path of sadhusamik/speech_recognition_tools, behind plain C entry points.
 *
 * The reference has no FFI: its hot path is the Python function `getFeats(args)`
 * (src/featgen/computeFDLPSpectrogram.py:29-237) calling numpy/scipy.  The entry points below are what a
 * ctypes binding in that file would call instead of the inner loops (INTEGRATION.md shows the stub):
 *
 *   fdlp_plan_create    <- one-time setup of getFeats: filterbank (computeFDLPSpectrogram.py:49-63,
 *                          features.py:172-219), mask/lifter/gamma weights (:94-118) and the derived
 *                          integer geometry (:104,173,201-204,220-221; features.py:134-135)
 *   fdlp_forward        <- the per-utterance body (:159-227): noise mix (features.py:24-31), getFrames
 *                          (features.py:118-154), DCT (:178), the (segment, band) hot loop (:188-218:
 *                          computeLpcFast features.py:222-230, computeModSpecFromLpc features.py:233-246,
 *                          envelope :201-205), overlap-add (:207-225), log (:227) -- for a packed batch
 *   mel_plan_create / mel_forward <- compute_mel_spectrum (src/featgen/computeMelSpectrum.py:40-157)
 *   cmvn partial sums   <- Kaldi `compute-cmvn-stats` (e2e/wsj/run_fdlp_e1.sh:280), accumulated in the
 *                          last kernel; the cross-GPU sum is an NCCL all-reduce done by the host.
 *
 * Conventions
 *   - every function returns an int status: 0 = ok, <0 = error (FDLP_E_*); no C++ exception crosses the ABI;
 *   - "dev" pointers are device pointers owned by the caller (e.g. torch tensors' data_ptr()); the library
 *     only borrows them for the duration of the stream-ordered work it enqueues;
 *   - "host" pointers are read during the call only;
 *   - the plan owns its tables (filterbank, twiddles, windows, cosine basis) and its scratch buffers;
 *     one plan per (device, configuration); a plan is not re-entrant, distinct plans are independent;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); nothing synchronises the device
 *     except plan creation/destruction and scratch growth;
 *   - every params / batch struct starts with `uint32_t struct_size`, which the caller sets to sizeof(the struct)
 *     as ITS binding declares it.  A binding generated from an older or newer header hands over a different size
 *     and the call returns FDLP_E_ARG instead of reading fields that are not there (fdlp_sizeof_struct reports the
 *     library's own sizes; FDLP_ABI_VERSION / fdlp_version() change whenever a struct layout does).
 */
#ifndef FDLP_B200_H
#define FDLP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FDLP_ABI_VERSION 200

#define FDLP_OK 0
#define FDLP_E_ARG (-1)         /* invalid argument / unsupported configuration            */
#define FDLP_E_CUDA (-2)        /* a CUDA runtime call failed (see fdlp_last_error)         */
#define FDLP_E_UNSUPPORTED (-3) /* segment length not factorable into radices 2,3,4,5 etc. */
#define FDLP_E_NOMEM (-4)

/* per-utterance status bits written by fdlp_forward (0 = clean) */
#define FDLP_ST_SINGULAR 1 /* a Levinson step hit a zero/non-finite prediction error (reference: LinAlgError) */
#define FDLP_ST_NONFINITE 2 /* gain <= 0 or non-finite: the band is NaN in the reference as well              */
#define FDLP_ST_ENV_RANGE 4 /* an envelope value left the float32 range of the K5 -> K6 hand-over (finite and non-zero
                               in float64): re-run with fdlp_params.env_f64 = 1                                 */

#define FDLP_FBANK_MEL 0      /* createFbank(nfilters, 2N, srate, warp)                 features.py:172-190 */
#define FDLP_FBANK_COCHLEAR 1 /* createFbankCochlear(nfilters, 2N, srate, om_w, alp, fixed, bet, warp) :197-219 */
#define FDLP_FBANK_CUSTOM 2   /* caller-supplied [nfilters, N+1] float64 rows (host pointer)                 */

#define FDLP_WINDOW_HAMMING 0
#define FDLP_WINDOW_HANNING 1
#define FDLP_WINDOW_SQUARE 2

#define FDLP_WAV_INT16 0
#define FDLP_WAV_FLOAT32 1
#define FDLP_WAV_FLOAT64 2 /* fdlp_forward / fdlp_modspec_forward only: the waveform fdlp_reverb_forward returns */

#define FDLP_NOISE_NONE 0
#define FDLP_NOISE_INT16_WRAP 1 /* energies squared and averaged with int16 wrap-around, features.py:27-28 on int16 input */
#define FDLP_NOISE_FLOAT 2      /* energies in floating point (float32 signal: float32 squares; float64 mean)  */

typedef struct fdlp_plan fdlp_plan;
typedef struct mel_plan mel_plan;

/* Arguments of the reference's argparse surface that shape the computation
 * (computeFDLPSpectrogram.py:241-262).  String options are pre-parsed by the Python host. */
typedef struct {
  uint32_t struct_size;     /* = sizeof(fdlp_params) */
  int32_t srate;            /* 16000 */
  int32_t nfilters;         /* --nfilters */
  int32_t order;            /* --order */
  int32_t coeff_num;        /* --coeff_num */
  int32_t frate;            /* --frate */
  double fduration;         /* --fduration */
  double overlap_fraction;  /* --overlap_fraction as given by the user (the reference flips it to 1-x) */
  int32_t fbank_kind;       /* FDLP_FBANK_* */
  double fbank_args[5];     /* mel: {warp}; cochlear: {om_w, alp, fixed, bet, warp} (--fbank_type) */
  const double* weights;    /* host, [coeff_num]: mask * lifter * gamma (and odd_mod_zero) folded, :94-118,194-200 */
  const double* custom_fbank; /* host, [nfilters, N+1] when fbank_kind == FDLP_FBANK_CUSTOM, else NULL */
  int32_t dct_fp32;         /* 0: float64 shared-memory FFT when it fits (default); 1: force float32 */
  double tail_eps;          /* autocorrelation skips filter taps below tail_eps * max|F_j| (0: only exact zeros) */
  /* modulation-spectrum extractor (computeModulationSpectrum.py): all zero for the FDLP spectrogram */
  int32_t window_kind;      /* FDLP_WINDOW_*: getFrames window (computeFDLPSpectrogram.py:29 hamming,
                               computeModulationSpectrum.py:31 hanning, :88-90 square) */
  int32_t seg_shift;        /* > 0: segment shift in samples, int(srate / frate) (computeModulationSpectrum.py:143-144);
                               0: derived from overlap_fraction */
  int32_t cepstrum_mode;    /* 1: the plan emits cepstra (fdlp_modspec_forward): L = coeff_num, no truncation to M */
  int32_t env_f64;          /* 0 (default): envelopes cross from K5 to K6 as float32 (a frame sums at most two, the output
                               is float32: ~1e-7 in the log features; values that leave the float32 range raise
                               FDLP_ST_ENV_RANGE); 1: float64 hand-over (extreme --lifter_config / --gamma_weight scalings) */
} fdlp_params;

/* Integer geometry derived exactly as the reference does (np.round = half-to-even). */
typedef struct {
  int32_t N;     /* segment length int(srate*fduration) */
  int32_t S;     /* segment shift int(srate / (1/((1-of)*fduration))) */
  int32_t M;     /* envelope FFT length 2*int(fduration*frate) */
  int32_t kk;    /* round(fduration*frate) */
  int32_t kkb2;  /* round(fduration*frate/2) */
  int32_t kks;   /* round(fduration*frate*(1-of)) */
  int32_t L;     /* min(coeff_num, M): cepstral coefficients that reach the envelope */
  int32_t nlags; /* order + 2 */
  int32_t dct_is_fp64;
  int64_t support_taps; /* sum over bands of the taps inside the filters' support (<= nfilters * N) */
  /* family decomposition of the autocorrelation (0 when the plan uses the direct per-band kernel only) */
  int64_t family_taps;  /* taps per segment of the band-independent block sums (nfam * N) */
  int64_t direct_taps;  /* sum over bands of the taps that still go through the direct per-band kernel */
  int32_t nfam;         /* number of templates (pass band + skirts) */
  int32_t block_len;    /* samples per block */
} fdlp_geometry;

/* One packed batch.  All pointers are DEVICE pointers unless noted. */
typedef struct {
  uint32_t struct_size;         /* = sizeof(fdlp_batch) */
  const void* wav;              /* packed samples of all utterances, int16 or float32 */
  int32_t wav_dtype;            /* FDLP_WAV_* */
  int32_t n_utt;
  int32_t n_seg;                /* total segments in the batch = sum_u ceil((T_u-1)/S) */
  int64_t n_frames;             /* total output frames = sum_u ceil(T_u*frate/srate) */
  int32_t max_utt_frames;       /* longest utterance of the batch, in frames */
  const int64_t* utt_wav_off;   /* [n_utt+1] sample offsets into wav */
  const int32_t* utt_seg_off;   /* [n_utt+1] first segment of each utterance */
  const int64_t* utt_frame_off; /* [n_utt+1] first output frame of each utterance */
  const int32_t* seg_utt;       /* [n_seg] owning utterance */
  const int32_t* seg_ptr;       /* [n_seg] overlap-add frame pointer of the segment (jitter included,
                                   computeFDLPSpectrogram.py:220-225), computed by the host */
  /* optional noise mixing (features.py:24-31); noise_mode == FDLP_NOISE_NONE disables it */
  int32_t noise_mode;
  const void* noise;            /* noise samples, same dtype rules as the reference: int16 */
  const int64_t* utt_noise_off; /* [n_utt] start offset into noise for each utterance (rand_num) */
  const double* utt_snr_db;     /* [n_utt] SNR in dB */
  /* outputs */
  float* feats;                 /* [n_frames, nfilters] row-major log features */
  double* cmvn_acc;             /* [2, nfilters+1] running sums (ADDED to; caller zeroes), or NULL */
  int32_t* status;              /* [n_utt] FDLP_ST_* bits (OR-ed into; caller zeroes) */
} fdlp_batch;

const char* fdlp_last_error(void);
int fdlp_version(void); /* FDLP_ABI_VERSION of the loaded library */
/* sizeof() of the library's own struct `which`: 0 fdlp_params, 1 fdlp_geometry, 2 fdlp_batch, 3 mel_params,
 * 4 mel_batch, 5 fdlp_reverb_batch, 6 mfcc_params, 7 mfcc_batch; -1 for an unknown index */
int64_t fdlp_sizeof_struct(int32_t which);

int fdlp_plan_create(const fdlp_params* params, fdlp_plan** out_plan);
int fdlp_plan_destroy(fdlp_plan* plan);
int fdlp_plan_get_geometry(const fdlp_plan* plan, fdlp_geometry* out);
/* copy the plan's filterbank ([nfilters, N], last column already dropped) to a host buffer */
int fdlp_plan_get_fbank(const fdlp_plan* plan, double* host_out);
/* the filterbank builders on their own (host only): createFbank / createFbankCochlear(nfilters, nfft, srate, ...)
 * (features.py:172-219) as float64 [nfilters, floor(nfft/2 + 1)], fbank_args as in fdlp_params */
int fdlp_make_fbank(int32_t fbank_kind, const double* fbank_args /*host [5]*/, int32_t nfilters, int32_t nfft,
                    int32_t srate, double* host_out);
/* bytes of scratch the plan will hold for a batch of n_seg segments */
int64_t fdlp_plan_scratch_bytes(const fdlp_plan* plan, int32_t n_seg);

/* the whole hot path for one packed batch, stream-ordered */
int fdlp_forward(fdlp_plan* plan, const fdlp_batch* batch, void* stream);

/* optional per-stage timing of fdlp_forward with CUDA events recorded on the launching stream:
 * out4 = {framing+DCT (incl. noise coefficient), autocorrelation, LPC+cepstrum+envelope, OLA+log} in ms
 * for the most recent fdlp_forward (get_stage_ms waits for that call to finish) */
int fdlp_plan_set_profiling(fdlp_plan* plan, int32_t enable);
/* fdlp_forward pipelines chunks of segments over two streams (the latency-bound LPC recursions of one chunk
 * overlap the FP64-bound autocorrelation of the next).  Off by default: on B200 the two kernels do not become
 * co-resident (measured), so it brings nothing yet; results are identical either way. */
int fdlp_plan_set_overlap(fdlp_plan* plan, int32_t enable);
int fdlp_plan_get_stage_ms(fdlp_plan* plan, float* out4);
/* autocorrelation algorithm: 0 (default) = family decomposition when the filterbank allows it (cochlear filters:
 * block sums shared by all bands + a direct remainder), 1 = always the direct per-band kernel.  Both evaluate
 * computeLpcFast's circular autocorrelation (features.py:223-225) with float64 products and sums. */
int fdlp_plan_set_autocorr_mode(fdlp_plan* plan, int32_t mode);
/* out3 = {family block sums, per-band combination, direct per-band kernel} in ms for the most recent
 * profiled fdlp_forward */
int fdlp_plan_get_autocorr_ms(fdlp_plan* plan, float* out3);
/* out7 = per-kernel ms of the most recent profiled fdlp_forward: {framing+DCT, family block sums, family scan,
 * direct per-band autocorrelation (+ family gather), LPC+cepstrum, envelope, overlap-add+log} */
int fdlp_plan_get_kernel_ms(fdlp_plan* plan, float* out7);

/* stage-level entry points (used by the parity tests to check each kernel against the oracle;
 * they run the same kernels fdlp_forward launches, on the plan's scratch) */
int fdlp_stage_dct(fdlp_plan* plan, const fdlp_batch* batch, double* dct_out /*[n_seg,N]*/, void* stream);
int fdlp_stage_autocorr(fdlp_plan* plan, const double* dct /*[n_seg,N]*/, int32_t n_seg,
                        double* r_out /*[n_seg,nfilters,nlags]*/, void* stream);
int fdlp_stage_lpc_envelope(fdlp_plan* plan, const double* r /*[n_seg,nfilters,nlags]*/, int32_t n_seg,
                            double* lpc_out /*[n_seg,nfilters,order+1] or NULL*/, double* gain_out /*[n_seg,nfilters] or NULL*/,
                            double* cep_out /*[n_seg,nfilters,L] or NULL*/, double* env_out /*[n_seg,nfilters,kk]*/,
                            int32_t* seg_status /*[n_seg] or NULL*/, void* stream);

/* ---- mel-spectrogram baseline (computeMelSpectrum.py) ---- */
typedef struct {
  uint32_t struct_size;  /* = sizeof(mel_params) */
  int32_t srate;
  int32_t nfilters;
  int32_t nfft;          /* --nfft (power of two) */
  int32_t frate;
  double fduration;      /* --fduration (0.02) */
  int32_t fbank_kind;
  double fbank_args[5];
  const double* custom_fbank; /* host [nfilters, nfft/2+1] */
  int32_t power;         /* 0: log10 (spectrum_type=log), 1: squared (spectrum_type=power) */
} mel_params;

typedef struct {
  uint32_t struct_size;         /* = sizeof(mel_batch) */
  const void* wav;
  int32_t wav_dtype;
  int32_t n_utt;
  int64_t n_frames;             /* sum_u ceil((T_u-1)/shift) */
  const int64_t* utt_wav_off;   /* [n_utt+1] */
  const int64_t* utt_frame_off; /* [n_utt+1] */
  float* feats;                 /* [n_frames, nfilters] */
  double* cmvn_acc;             /* [2, nfilters+1] or NULL */
} mel_batch;

int mel_plan_create(const mel_params* params, mel_plan** out_plan);
int mel_plan_destroy(mel_plan* plan);
int mel_plan_get_fbank(const mel_plan* plan, double* host_out /*[nfilters, nfft/2+1]*/);
int mel_forward(mel_plan* plan, const mel_batch* batch, void* stream);

/* ---- modulation-spectrum features (computeModulationSpectrum.py:143-198, real-modulation path): per segment
 *      (100 Hz shift, hanning window) and band the cepstrum of the sub-band LPC model; out[seg, j*feat_len + q] =
 *      cep[seg, j, sel_idx[q]] * sel_scale[q] (absolute value if `absolute`).  Uses wav / utt_wav_off /
 *      utt_seg_off / seg_utt / n_utt / n_seg / status of the batch; the plan must have cepstrum_mode = 1. ---- */
int fdlp_modspec_forward(fdlp_plan* plan, const fdlp_batch* batch, const int32_t* sel_idx /*dev [feat_len]*/,
                         const double* sel_scale /*dev [feat_len]*/, int32_t feat_len, int32_t absolute,
                         float* out /*dev [n_seg, nfilters*feat_len]*/, void* stream);

/* ---- --complex_modulation of the same extractor (computeModulationSpectrum.py:45-48, :150-152, :164-173;
 *      features.py:222-246 with keepreal=False).  Three steps, the middle one the caller's:
 *      (1) fdlp_stage_frames: the plan's windowed, reflect-padded frames (getFrames) as float64 [n_seg, N];
 *      (2) the caller takes ifft(frames)[:, :N/2] (scipy.fftpack.ifft: 1/N scaling) -- a library FFT;
 *      (3) fdlp_modspec_complex_forward: per (frame, band) the filterbank taps [row_lo, row_hi) of fbank [nfilters, M]
 *          (createFbank(nfilters, N, ...)[:, :-1], see fdlp_make_fbank) times the
 *          spectrum, its circular complex autocorrelation, the Hermitian Toeplitz solve, the complex cepstrum
 *          c[0..coeff_n-1], and out[seg, j*feat_len + q] = |c[sel_idx[q]] * sel_scale[q]| (absolute, feat_len = n_sel)
 *          or the real parts in [0, n_sel) and the imaginary parts in [n_sel, 2 n_sel) (feat_len = 2 n_sel).
 *          max_width = max(row_hi - row_lo); status [n_utt] is OR-ed into through seg_utt (may be NULL).  All
 *          pointers are device pointers; nothing is allocated. ---- */
int fdlp_stage_frames(fdlp_plan* plan, const fdlp_batch* batch, double* frames_out /*dev [n_seg, N]*/, void* stream);
int fdlp_modspec_complex_forward(const double* spec /*dev [n_seg, spec_ld] complex128 (re, im)*/, int64_t spec_ld /*complex elements per row*/,
                            int64_t n_seg, const int32_t* seg_utt /*dev [n_seg]*/, int32_t M, const double* fbank /*dev [nfilters, M]*/,
                            const int32_t* row_lo /*dev [nfilters]*/, const int32_t* row_hi /*dev [nfilters]*/, int32_t max_width,
                            int32_t nfilters, int32_t order, int32_t coeff_n, const int32_t* sel_idx /*dev [n_sel]*/,
                            const double* sel_scale /*dev [n_sel]*/, int32_t n_sel, int32_t absolute,
                            float* out /*dev [n_seg, nfilters*feat_len]*/, int32_t* status /*dev [n_utt] or NULL*/, void* stream);

/* ---- host-side WAV ingest of the CLI shim (no device work; computeFDLPSpectrogram.py:129-152 reads each file with
 *      scipy.io.wavfile.read).  fdlp_wav_probe: info[i] = {sample rate, format, byte offset of the samples, number of
 *      samples}, format 0 = PCM16, 1 = IEEE float32, -1 = not a plain mono RIFF file of those two kinds (or unreadable):
 *      the caller falls back to the reference's reader.  fdlp_wav_read: nbytes[i] bytes at data_off[i] of file i to
 *      dst_base + dst_off[i] (pinned staging memory); status[i] = 0, or 1 when the file could not be read in full.
 *      Both spread the files over `threads` host threads. ---- */
int fdlp_wav_probe(const char* const* paths, int32_t n, int32_t threads, int64_t* info /*host [n, 4]*/);
int fdlp_wav_read(const char* const* paths, const int64_t* data_off /*host [n]*/, const int64_t* nbytes /*host [n]*/,
                  const int64_t* dst_off /*host [n]*/, void* dst_base /*host*/, int32_t n, int32_t threads,
                  int32_t* status /*host [n]*/);

/* ---- --add_noise diff (computeFDLPSpectrogram.py:162-164): signal = convolve(signal, [1,2,3,2,0,-2,-5,-2,0,2,3,2,1],
 *      mode='same') per utterance; the float32 result (exact for int16 input) is then fed to fdlp_forward ---- */
int fdlp_diff_filter(const void* wav /*dev, packed*/, int32_t wav_dtype, const int64_t* utt_wav_off /*dev [n_utt+1]*/,
                     int32_t n_utt, int64_t max_len /*longest utterance, samples*/, float* out /*dev, packed*/,
                     void* stream);

/* ---- --add_noise type,snr as a waveform of its own (add_noise_to_wav, features.py:24-31: alpha from the energies,
 *      then sig + alpha * noise[off : off + T]); computeMelSpectrum.py:134-141 mixes before framing.  fdlp_forward fuses the
 *      same mix into its framing load; this entry point serves the paths that take a waveform (mel).  utt_alpha: device
 *      scratch of n_utt doubles supplied by the caller, left holding the mixing coefficients.  float32 result. ---- */
int fdlp_noise_mix(const void* wav /*dev, packed*/, int32_t wav_dtype, const int64_t* utt_wav_off /*dev [n_utt+1]*/,
                   int32_t n_utt, int64_t max_len, const int16_t* noise /*dev*/, const int64_t* utt_noise_off /*dev [n_utt]*/,
                   const double* utt_snr_db /*dev [n_utt]*/, int32_t noise_mode /*FDLP_NOISE_INT16_WRAP | FDLP_NOISE_FLOAT*/,
                   double* utt_alpha /*dev [n_utt], out*/, float* out /*dev, packed*/, void* stream);

/* ---- --add_reverb (features.py:110-115 addReverb, called at computeFDLPSpectrogram.py:168-170, computeMelSpectrum.py:143-145,
 *      computeMfccFeatures.py:118-120): out = convolve(sig, rir) (full), xxc = correlate(sig, out, 'valid'),
 *      indM = len(xxc) - argmax(xxc), result = out[indM : indM + len(sig)], per utterance, after the optional noise
 *      mixing (same noise arrays as fdlp_batch).  `rir` is the HOST float64 impulse response as the reference
 *      forms it (rir[:, 1] / 2^15, :76-87).  The convolution runs over blocks of `block` samples
 *      (fdlp_reverb_plan_get_block), two blocks to a complex transform: utterance u has
 *      D_u = ceil(ceil((T_u + R - 1) / block) / 2) block pairs and ceil(D_u / pairs_per_cta) convolution CTAs;
 *      utt_pair_off / pair_utt and utt_conv_off / conv_utt are their prefix sums and owner tables, computed by the
 *      host like utt_seg_off / seg_utt.  The result has the packed layout of the
 *      input, float64 (for fdlp_forward with FDLP_WAV_FLOAT64) or float32 (mel / MFCC paths); out_len[u] is T_u, or
 *      T_u - 1 in the reference's corner case indM == R (the slice runs off the end of the convolution). ---- */
typedef struct fdlp_reverb_plan fdlp_reverb_plan;
typedef struct {
  uint32_t struct_size;         /* = sizeof(fdlp_reverb_batch) */
  const void* wav;              /* dev, packed int16 or float32 */
  int32_t wav_dtype;
  int32_t n_utt;
  const int64_t* utt_wav_off;   /* dev [n_utt+1] */
  int64_t max_len;              /* longest utterance, samples */
  int32_t n_pair;               /* total block pairs */
  const int32_t* utt_pair_off;  /* dev [n_utt+1] */
  const int32_t* pair_utt;      /* dev [n_pair] */
  int32_t n_conv;               /* total convolution CTAs */
  const int32_t* utt_conv_off;  /* dev [n_utt+1] */
  const int32_t* conv_utt;      /* dev [n_conv] */
  int32_t noise_mode;           /* FDLP_NOISE_* ; the mix happens before the reverberation (:158-170) */
  const void* noise;
  const int64_t* utt_noise_off;
  const double* utt_snr_db;
  void* out;                    /* dev, packed like wav */
  int32_t out_dtype;            /* FDLP_WAV_FLOAT32 or FDLP_WAV_FLOAT64 */
  int32_t* ind;                 /* dev [n_utt] indM of each utterance, or NULL */
  int64_t* out_len;             /* dev [n_utt] samples produced, or NULL */
} fdlp_reverb_batch;

int fdlp_reverb_plan_create(const double* rir /*host [R]*/, int32_t R, int32_t block /*0: auto*/, fdlp_reverb_plan** out_plan);
int fdlp_reverb_plan_destroy(fdlp_reverb_plan* plan);
int fdlp_reverb_plan_get_block(const fdlp_reverb_plan* plan, int32_t* block, int32_t* partitions, int32_t* pairs_per_cta);
int fdlp_reverb_plan_set_profiling(fdlp_reverb_plan* plan, int32_t enable);
/* CUDA-event durations of the most recent profiled forward: window spectra, convolution, correlation, alignment */
int fdlp_reverb_plan_get_kernel_ms(fdlp_reverb_plan* plan, float* out4);
int64_t fdlp_reverb_plan_scratch_bytes(const fdlp_reverb_plan* plan, int32_t n_pair, int32_t n_utt);
int fdlp_reverb_forward(fdlp_reverb_plan* plan, const fdlp_reverb_batch* batch, void* stream);

/* ---- MFCC features (computeMfccFeatures.py:111-131: signal / 2^15 [+ noise], Hamming frames, the reference's
 *      (nfft/2 + 1)-point DFT magnitude against createFbank(nfilters, nfft, srate), log10, type-II DCT along the bands,
 *      first `ncep` coefficients; optional spliceFeats context, features.py:157-169) ---- */
typedef struct mfcc_plan mfcc_plan;
typedef struct {
  uint32_t struct_size;  /* = sizeof(mfcc_params) */
  int32_t srate;
  int32_t nfilters;      /* --nfilters (30) */
  int32_t nfft;          /* --nfft (1024), even */
  int32_t frate;         /* --frate (100) */
  double fduration;      /* --fduration (0.02) */
  int32_t ncep;          /* coefficients kept (13, computeMfccFeatures.py:127) */
} mfcc_params;

typedef struct {
  uint32_t struct_size;         /* = sizeof(mfcc_batch) */
  const void* wav;              /* dev, packed int16 / float32 as read from the WAV (not normalised) */
  int32_t wav_dtype;
  int32_t n_utt;
  int64_t n_frames;             /* sum_u frames(T_u), as for mel_batch */
  const int64_t* utt_wav_off;   /* dev [n_utt+1] */
  const int64_t* utt_frame_off; /* dev [n_utt+1] */
  float* feats;                 /* dev [n_frames, ncep * (2 context + 1)] */
  int32_t context;              /* --context (0: no splicing) */
  const int16_t* noise;         /* dev noise recording or NULL (--add_noise type,snr: computeMfccFeatures.py:45-52) */
  const int64_t* utt_noise_off; /* dev [n_utt] floor(rand * (len(noise) - T_u)) */
  const double* utt_snr_db;     /* dev [n_utt] */
} mfcc_batch;

int mfcc_plan_create(const mfcc_params* params, mfcc_plan** out_plan);
int mfcc_plan_destroy(mfcc_plan* plan);
/* The 513-point transform of the default configuration (nfft 1024, 320-sample frames, <= 32 bands) runs as a 3xTF32 GEMM
 * on the tensor cores (tcgen05 / TMA / tensor memory, csrc/mfcc_tc.cuh); other configurations, enable = 0 and
 * FDLP_MFCC_TC=0 use the FP32-pipe kernel.  Returns 1 if the plan uses the tensor-core kernel after the call, else 0. */
int mfcc_plan_set_tensor_cores(mfcc_plan* plan, int32_t enable);
int mfcc_forward(mfcc_plan* plan, const mfcc_batch* batch, void* stream);

/* ---- CMVN apply (ESPnet dump.sh / Kaldi apply-cmvn, e2e/wsj/run_fdlp_e1.sh:293-303) ---- */
int cmvn_apply(float* feats /*dev [n_frames, dim], in place*/, int64_t n_frames, int32_t dim,
               const double* stats /*dev [2, dim+1]*/, int32_t norm_vars, void* stream);

/* ---- allocation-free CMVN accumulation.  NO COUNTERPART IN THE REFERENCE: added
 *      for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 *      allocation).  stats[2, dim+1] += [sum x (dim), n_frames; sum x^2 (dim), 0] over feats[n_frames, dim] -- the same
 *      kernels and summation order as the `cmvn_acc` argument of mel_forward (bit-identical results) -- with the per-CTA
 *      partial sums in a file-scope device array of 296 x 2 x (FDLP_NOALLOC_MAX_DIM + 1) doubles = 1.2 MB that every GPU
 *      loading the library holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library
 *      routine); dim > FDLP_NOALLOC_MAX_DIM returns FDLP_E_ARG.  All calls on a device share the array: each call
 *      overwrites every slot it later reads, so it needs no clearing between calls, but two calls that OVERLAP on the GPU
 *      (different streams, no ordering between them) race -- order them on one stream or with events. ---- */
#define FDLP_NOALLOC_MAX_DIM 255
int cmvn_accumulate_noalloc(const float* feats /*dev [n_frames, dim]*/, int64_t n_frames, int32_t dim,
                            double* stats /*dev [2, dim+1], ADDED to (caller zeroes)*/, void* stream);

/* Synthetic white utterances written straight into a packed device batch (measurement input, SURVEY.md 8d: the
 * 1000-hour workload is never stored).  Bit-identical to speech_recognition_tools_b200/synth.py white_int16(seed,
 * utt_id[u], T_u, scale_q16).  No reference counterpart. */
int fdlp_synth_white(int16_t* out /*dev, packed*/, const int64_t* utt_wav_off /*dev [n_utt+1]*/,
                     const int64_t* utt_id /*dev [n_utt]*/, int32_t n_utt, int64_t max_len, uint64_t seed,
                     int32_t scale_q16 /*5197: sigma ~ 3000*/, void* stream);

/* FP64 pipe probe: runs a register-resident DFMA / DMMA loop and returns TFLOP/s (roofline denominator) */
int fdlp_measure_fp64_peak(double* dfma_tflops, double* dmma_tflops);

#ifdef __cplusplus
}
#endif
#endif /* FDLP_B200_H */
