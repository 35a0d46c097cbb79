"""numpy restatement of the reference MFCC extractor (TEST INFRASTRUCTURE, see oracle/__init__.py).
Follows /root/reference/src/featgen/computeMfccFeatures.py:45-133 and spliceFeats (features.py:157-169)."""
from __future__ import annotations

import numpy as np
import scipy.fftpack

from .fdlp import get_frames, make_fbank


def add_noise_normalised(sig, noise, snr, rand):
    """computeMfccFeatures.py:45-52 on signals already divided by 2^15 (float64: no integer wrap-around here).
    `rand` stands for the np.random.rand() draw."""
    off = int(np.floor(rand * (len(noise) - len(sig))))
    ns = noise[off:off + len(sig)]
    alp = np.sqrt(np.mean(sig ** 2) / (np.mean(ns ** 2) * (10 ** (snr / 10))))
    return sig + alp * ns


def splice_feats(feats, context):
    """features.py:157-169: +-context frames side by side, zero padded; the reference's loop stops `context`
    frames early, so the last `context` rows stay zero."""
    context = int(context)
    F, D = feats.shape
    out = np.zeros((F, D * (2 * context + 1)))
    padded = np.concatenate([np.zeros((context, D)), feats, np.zeros((context, D))], axis=0)
    for i in range(0, F - context):
        out[i, :] = padded[i:i + 2 * context + 1].reshape(-1)
    return out


def mfcc_features(signal, *, nfilters=30, nfft=1024, fduration=0.02, frate=100, context=None, srate=16000,
                  noise=None, snr=None, rand=None, ncep=13, window=np.hamming, reverb=None):
    """computeMfccFeatures.py:111-131.  `signal` (and `noise`) as read from the WAV (int16 or float32): both are
    divided by 2^15 (:40,:113).  Note the transform length: fft(frames, nfft/2+1) is an (nfft/2+1)-POINT DFT of
    the frame (513 points for nfft = 1024, frame truncated or zero-padded to that length) whose every bin --
    the mirrored upper half included -- meets the [nfilters, nfft/2+1] filterbank.  Returns float64 [F, 13]."""
    sig = np.asarray(signal) / np.power(2, 15)
    if noise is not None:
        sig = add_noise_normalised(sig, np.asarray(noise) / np.power(2, 15), float(snr), float(rand))
    if reverb is not None:
        from oracle.reverb import add_reverb
        sig = add_reverb(sig, reverb)                                    # :118-120
    fbank = make_fbank("mel,1", nfilters, nfft, srate)                   # :72  createFbank(nfilters, nfft, srate)
    time_frames = get_frames(sig, srate, frate, fduration, window)       # :121-122
    spec = np.abs(scipy.fftpack.fft(time_frames, int(nfft / 2 + 1), axis=1))
    with np.errstate(divide="ignore"):
        mel_energy = np.log10(np.matmul(spec, np.transpose(fbank)))      # :124-125
    feats = scipy.fftpack.dct(mel_energy, axis=1)[:, 0:ncep]             # :126-127 (type II, unnormalised)
    if context:
        feats = splice_feats(feats, context)                            # :129-130
    return feats
