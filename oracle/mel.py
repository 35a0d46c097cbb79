"""numpy restatement of the reference mel-spectrogram extractor (TEST INFRASTRUCTURE, see
oracle/__init__.py).  Follows /root/reference/src/featgen/computeMelSpectrum.py:147-157."""
from __future__ import annotations

import numpy as np
import scipy.fftpack

from .fdlp import get_frames, make_fbank


def mel_spectrogram(signal, *, nfilters=23, nfft=1024, fduration=0.02, frate=100, fbank_type="mel,1",
                    fbank=None, spectrum_type="log", srate=16000, window=np.hamming):
    """computeMelSpectrum.py:147-157: Hamming frames -> |FFT_nfft|[:nfft/2+1] @ fbank.T -> log10
    (or squared for spectrum_type='power').  Returns float64 [F, nfilters]."""
    if fbank is None:
        fbank = make_fbank(fbank_type, nfilters, nfft, srate)          # computeMelSpectrum.py:52-65
    time_frames = get_frames(signal, srate, frate, fduration, window)  # :147-148
    spec = np.abs(scipy.fftpack.fft(time_frames, nfft, axis=1)[:, :int(nfft / 2 + 1)])
    e = np.matmul(spec, np.transpose(fbank))
    if spectrum_type == "log":
        with np.errstate(divide="ignore"):
            return np.log10(e)                                         # :150-152
    if spectrum_type == "power":
        return np.power(e, 2)                                          # :153-155
    raise ValueError("Spectrum type not supported! ")
