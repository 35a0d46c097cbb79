"""numpy/scipy restatement of the reference FDLP-spectrogram extractor (TEST INFRASTRUCTURE, see
oracle/__init__.py).  Every function cites the reference lines it follows
(paths relative to /root/reference/src/featgen).

The arithmetic is the reference's: float64 throughout, FFT-based circular autocorrelation,
scipy Levinson (`solve_toeplitz`), the LPC->cepstrum recursion, `abs(exp(fft(ms, M)))`.  The only
liberty taken is batching the per-band work of one segment into array operations (the reference
loops `for j in range(nfilters)` around identical 1-D calls).
"""
from __future__ import annotations

import random as _random

import numpy as np
import scipy.fftpack
import scipy.linalg
import scipy.signal
import scipy.stats

LOG_FLOOR = 0.00000000000001  # computeFDLPSpectrogram.py:227


# --------------------------------------------------------------------------- framing
def get_frames(signal, srate, frate, flength, window=np.hamming):
    """features.py:118-154 (getFrames).  Returns an [nframes, N] float64 array.

    `frate` is a frame RATE in Hz (the FDLP caller passes lfr = 1/(of*fduration)), and the shift is
    `int(srate / frate)` with the float division kept exactly as written in the reference."""
    flength_samples = int(srate * flength)
    frate_samples = int(srate / frate)
    if flength_samples % 2 == 0:
        sp_b = int(flength_samples / 2) - 1
        sp_f = int(flength_samples / 2)
        extend = int(flength_samples / 2) - 1
    else:
        sp_b = sp_f = extend = int((flength_samples - 1) / 2)
    sig_padded = np.pad(signal, extend, "reflect")
    win = window(flength_samples)
    frames = []
    idx = sp_b
    while (idx + sp_f) < len(sig_padded):
        frames.append(sig_padded[idx - sp_b: idx + sp_f + 1] * win)
        idx += frate_samples
    if not frames:
        return np.zeros((0, flength_samples))
    return np.array(frames)


# --------------------------------------------------------------------------- filterbanks
def create_fbank(nfilters, nfft, srate, warp_fact=1.0):
    """features.py:172-190 (createFbank): triangular mel filters on floor((nfft+1)*hz/srate) bins."""
    mel_max = 2595 * np.log10(1 + (srate / warp_fact) / 1400)
    fwarped = np.linspace(0, mel_max, nfilters + 2)
    nbins = int(np.floor(nfft / 2 + 1))
    filts = np.zeros((nfilters, nbins))
    hz_points = warp_fact * (700 * (10 ** (fwarped / 2595) - 1))
    bins = np.floor((nfft + 1) * hz_points / srate)
    for m in range(1, nfilters + 1):
        lo, ce, hi = int(bins[m - 1]), int(bins[m]), int(bins[m + 1])
        k = np.arange(lo, ce)
        filts[m - 1, lo:ce] = (k - bins[m - 1]) / (bins[m] - bins[m - 1]) if ce > lo else 0
        k = np.arange(ce, hi)
        filts[m - 1, ce:hi] = (bins[m + 1] - k) / (bins[m + 1] - bins[m]) if hi > ce else 0
    return filts


def warp_func_bark(x, warp_fact=1.0):
    """features.py:193-194."""
    return 6 * np.arcsinh((x / warp_fact) / 600)


def create_fbank_cochlear(nfilters, nfft, srate, om_w=0.2, alp=2.5, fixed=1, bet=2.5, warp_fact=1.0):
    """features.py:197-219 (createFbankCochlear), vectorised over the bin axis."""
    f_max = srate / 2
    warped_max = warp_func_bark(f_max, warp_fact)
    fwarped_cf = np.linspace(0, warped_max, nfilters)
    nbins = int(np.floor(nfft / 2 + 1))
    f_linear = np.linspace(0, f_max, nbins)
    f_warped = warp_func_bark(f_linear, warp_fact)
    filts = np.zeros((nfilters, nbins))
    alp_c = alp
    for i in range(nfilters):
        fc = fwarped_cf[i]
        a = alp_c if fixed == 1 else alp_c * np.exp(-0.1 * fc)
        d = f_warped - fc
        below = d <= -om_w / 2
        inside = (d > -om_w / 2) & (d < om_w / 2)
        row = np.power(10.0, -bet * (d - om_w / 2))
        row[below] = np.power(10.0, a * (d[below] + om_w / 2))
        row[inside] = 1.0
        filts[i] = row
    return filts


def make_fbank(fbank_type, nfilters, nfft, srate):
    """Filterbank selection, computeFDLPSpectrogram.py:49-63 / computeMelSpectrum.py:52-65."""
    tok = fbank_type.strip().split(",")
    if tok[0] == "mel":
        if len(tok) < 2:
            raise ValueError("Mel filter bank not configured properly....")
        return create_fbank(nfilters, nfft, srate, warp_fact=float(tok[1]))
    if tok[0] == "cochlear":
        if len(tok) < 6:
            raise ValueError("Cochlear filter bank not configured properly....")
        return create_fbank_cochlear(nfilters, nfft, srate, om_w=float(tok[1]), alp=float(tok[2]),
                                     fixed=int(tok[3]), bet=float(tok[4]), warp_fact=float(tok[5]))
    raise ValueError("Invalid type of filter bank, use mel or cochlear with proper configuration")


# --------------------------------------------------------------------------- LPC / cepstrum
def autocorr_circular(x):
    """features.py:223-225: real(ifft(fft(x) * conj(fft(x)))) along the last axis."""
    X = np.fft.fft(x, x.shape[-1], axis=-1)
    return np.real(np.fft.ifft(X * np.conj(X), axis=-1))


def compute_lpc_fast(signal, order):
    """features.py:222-230 (computeLpcFast) for one band: returns (xlpc[order+1], gg)."""
    y = autocorr_circular(signal)
    xlpc = scipy.linalg.solve_toeplitz(y[0:order], -y[1:order + 1])
    xlpc = np.append(1, xlpc)
    gg = y[0] + np.sum(xlpc * y[1:order + 2])
    return xlpc, gg


def compute_modspec_from_lpc(gg, xlpc, lim):
    """features.py:233-246 (computeModSpecFromLpc) for one band (does not mutate its input)."""
    x = np.array(xlpc, dtype=np.float64)
    x[1:] = -x[1:]
    cep = np.zeros(lim, dtype=x.dtype)
    with np.errstate(invalid="ignore", divide="ignore"):
        cep[0] = np.log(np.sqrt(gg))
    cep[1] = x[1]
    if x.shape[0] < lim:
        x = np.append(x, np.zeros(int(lim - x.shape[0] + 1)))
    for n in range(2, lim):
        aa = np.arange(1, n) / n
        bb = np.flipud(x[1:n])
        cc = cep[1:n]
        cep[n] = np.sum(aa * bb * cc) + x[n]
    return cep


def modspec_batch(gg, xlpc, lim):
    """Same recursion as compute_modspec_from_lpc for a [B, p+1] batch (gg: [B])."""
    x = np.array(xlpc, dtype=np.float64)
    x[:, 1:] = -x[:, 1:]
    B = x.shape[0]
    cep = np.zeros((B, lim))
    with np.errstate(invalid="ignore", divide="ignore"):
        cep[:, 0] = np.log(np.sqrt(gg))
    cep[:, 1] = x[:, 1]
    if x.shape[1] < lim:
        x = np.concatenate([x, np.zeros((B, int(lim - x.shape[1] + 1)))], axis=1)
    for n in range(2, lim):
        aa = np.arange(1, n) / n
        cep[:, n] = np.sum(aa * x[:, n - 1:0:-1] * cep[:, 1:n], axis=1) + x[:, n]
    return cep


# --------------------------------------------------------------------------- weights
def build_weights(coeff_num, coeff_range, order, fduration, gamma_weight="None", lifter=None):
    """Mask (computeFDLPSpectrogram.py:94-103), lifter (:43-46,195-196) and gamma weights (:107-118)
    folded into one length-`coeff_num` vector, in the order the reference multiplies them (:194-198)."""
    lo, hi = (int(v) for v in coeff_range.split(","))
    w = np.array([1.0 if (lo <= i <= hi) else 0.0 for i in range(coeff_num)])
    if lifter is not None:
        w = w * np.asarray(lifter, dtype=np.float64)
    g = gamma_weight.strip().split(",") if isinstance(gamma_weight, str) else list(gamma_weight)
    if g[0] != "None":
        x = np.linspace(0, order - 1, order)
        scale, shape, pk_required = float(g[0]), float(g[1]), float(g[2])
        pk_required = pk_required * (2 * fduration)
        pk = (shape - 1) * scale
        loc = -pk + pk_required
        w = w * (scipy.stats.gamma.pdf(x, a=shape, loc=loc, scale=scale) * 3 * scale)
    return w


def derived_sizes(fduration, frate, overlap_fraction, srate=16000):
    """The integer geometry the reference derives inline (computeFDLPSpectrogram.py:104,173,201-204,220-221;
    features.py:134-135).  `overlap_fraction` is the user flag (0.25); the reference flips it to 1-of."""
    of = 1 - overlap_fraction
    lfr = 1 / (of * fduration)
    N = int(srate * fduration)
    S = int(srate / lfr)
    M = 2 * int(fduration * frate)
    kk = int(np.round(fduration * frate))
    kkb2 = int(np.round(fduration * frate / 2))
    kks = int(np.round(fduration * frate * of))
    return dict(N=N, S=S, M=M, kk=kk, kkb2=kkb2, kks=kks, of=of, lfr=lfr)


# --------------------------------------------------------------------------- the extractor
def fdlp_spectrogram(signal, *, nfilters=20, order=50, fduration=0.5, frate=100, overlap_fraction=0.25,
                     coeff_num=50, coeff_range="1,20", fbank_type="mel,1", fbank=None,
                     gamma_weight="None", lifter=None, odd_mod_zero=False, rng=None,
                     srate=16000, window=np.hamming, return_trace=False):
    """One utterance through computeFDLPSpectrogram.py:172-227.  Returns float64 [Fr, nfilters].

    `rng` is a `random.Random` (or anything with randrange) standing for the module-level
    `random.randrange(2)` of line 225; one draw per segment i >= 1, after that segment's bands."""
    if rng is None:
        rng = _random
    g = derived_sizes(fduration, frate, overlap_fraction, srate)
    of = g["of"]
    if fbank is None:
        fbank = make_fbank(fbank_type, nfilters, int(2 * fduration * srate), srate)
    weights = build_weights(coeff_num, coeff_range, order, fduration, gamma_weight, lifter)

    tframes = signal.shape[0]
    time_frames = get_frames(signal, srate, g["lfr"], fduration, window)
    cos_trans = scipy.fftpack.dct(time_frames) / np.sqrt(2 * int(srate * fduration))
    frame_num, ndct = cos_trans.shape
    feats = np.zeros((nfilters, int(np.ceil(tframes * frate / srate))))
    Fr = feats.shape[1]
    ptr = 0
    ptrs = []
    M = 2 * int(fduration * frate)
    filt = fbank[:, 0:-1]
    for i in range(frame_num):
        band_dct = filt * cos_trans[i, :]                       # :190-191, all bands at once
        y = autocorr_circular(band_dct)                         # features.py:223-225
        xlpc = np.empty((nfilters, order + 1))
        gg = np.empty(nfilters)
        for j in range(nfilters):
            a = scipy.linalg.solve_toeplitz(y[j, 0:order], -y[j, 1:order + 1])   # features.py:226
            xlpc[j] = np.append(1, a)
            gg[j] = y[j, 0] + np.sum(xlpc[j] * y[j, 1:order + 2])                # features.py:228
        ms = modspec_batch(gg, xlpc, coeff_num)                 # :193
        ms = ms * weights                                       # :194-198
        if odd_mod_zero:
            ms[:, 1::2] = 0                                     # :199-200
        ms = scipy.fftpack.fft(ms, M, axis=1)                   # :201
        with np.errstate(over="ignore", invalid="ignore"):
            ms = np.abs(np.exp(ms))                             # :202
        kk = int(np.round(fduration * frate))
        kkb2 = int(np.round(fduration * frate / 2))
        with np.errstate(divide="ignore", invalid="ignore"):
            ms = ms[:, 0:kk] * np.hanning(kk) / window(kk)      # :205
        ptrs.append(ptr)
        if i == 0:                                              # :207-218
            if Fr < kkb2:
                feats[:, :] += ms[:, kkb2:kkb2 + Fr]
            else:
                feats[:, ptr:ptr + kkb2] += ms[:, kkb2:]
        elif i == frame_num - 1 or i == frame_num - 2:
            if ms.shape[1] >= Fr - ptr:
                feats[:, ptr:] += ms[:, :Fr - ptr]
            else:
                feats[:, ptr:ptr + kk] += ms
        else:
            feats[:, ptr:ptr + kk] += ms
        kks = int(np.round(fduration * frate * of))             # :220-225
        if i == 0:
            ptr = int(ptr + kks - kkb2)
        else:
            ptr = int(ptr + kks + rng.randrange(2))
    with np.errstate(invalid="ignore"):
        out = np.log(np.clip(feats.T, a_max=None, a_min=LOG_FLOOR))   # :227
    if return_trace:
        return out, dict(ptrs=ptrs, nseg=frame_num, Fr=Fr)
    return out


def fdlp_stages(signal, *, nfilters, order, fduration, overlap_fraction, coeff_num, fbank,
                srate=16000, window=np.hamming):
    """Per-stage intermediates of the hot loop for stage-by-stage kernel tests:
    frames, DCT, autocorrelation lags 0..p+1, LPC, gain, cepstrum (all float64)."""
    g = derived_sizes(fduration, 100, overlap_fraction, srate)
    time_frames = get_frames(signal, srate, g["lfr"], fduration, window)
    cos_trans = scipy.fftpack.dct(time_frames) / np.sqrt(2 * int(srate * fduration))
    nseg = cos_trans.shape[0]
    filt = fbank[:, 0:-1]
    r = np.empty((nseg, nfilters, order + 2))
    lpc = np.empty((nseg, nfilters, order + 1))
    gg = np.empty((nseg, nfilters))
    cep = np.empty((nseg, nfilters, coeff_num))
    for i in range(nseg):
        y = autocorr_circular(filt * cos_trans[i])
        r[i] = y[:, :order + 2]
        for j in range(nfilters):
            lpc[i, j], gg[i, j] = compute_lpc_fast(filt[j] * cos_trans[i], order)
        cep[i] = modspec_batch(gg[i], lpc[i], coeff_num)
    return dict(frames=time_frames, dct=cos_trans, r=r, lpc=lpc, gg=gg, cep=cep)


# --------------------------------------------------------------------------- augmentation
def add_noise_to_wav(sig, noise, snr, rand=None):
    """features.py:24-31 (add_noise_to_wav).  `rand` stands for the `np.random.rand()` draw.
    Energies are computed in the INPUT dtype: int16 arrays square and wrap in int16 exactly as
    numpy does in the reference."""
    if rand is None:
        rand = np.random.rand()
    rand_num = int(np.floor(rand * (len(noise) - len(sig))))
    ns = noise[rand_num:rand_num + len(sig)]
    with np.errstate(all="ignore"):
        E_s = np.mean(sig ** 2)
        E_n = np.mean(ns ** 2)
        alp = np.sqrt(E_s / (E_n * (10 ** (snr / 10))))
        return sig + alp * ns


DIFF_FIR = [1, 2, 3, 2, 0, -2, -5, -2, 0, 2, 3, 2, 1]  # computeFDLPSpectrogram.py:163


def diff_filter(sig):
    """computeFDLPSpectrogram.py:162-164 ('diff' mode): 13-tap FIR, mode='same'."""
    return scipy.signal.convolve(sig, DIFF_FIR, mode="same")
