"""ORACLE (test infrastructure, never imported by the product): numpy restatement of the reference's
reverberation augmentation, pinned on tests/golden/reverb_*.npz (outputs of the unmodified reference).

Follows addReverb, /root/reference/src/featgen/features.py:110-115, and the impulse-response handling of
computeFDLPSpectrogram.py:75-91.  Stated in terms of the alignment lag d instead of numpy's 'valid' index:

    full[n]  = sum_j reverb[j] sig[n - j]                      n = 0 .. T + R - 2          (:111)
    xx[d]    = sum_m sig[m] full[m + d]                        d = 0 .. R - 1              (:112, xxc[i], i = R-1-d)
    d*       = the LARGEST d attaining max xx                  (np.argmax takes the first i)
    indM     = d* + 1                                          (:113, len(xxc) - argmax)
    out      = full[indM : indM + T]                           (:114; T - 1 samples when indM == R)
"""
import numpy as np


def full_convolution(sig, reverb, exact=True):
    """Full linear convolution in float64.  For integer-valued `sig` and `reverb * 2^15` (int16 WAV data) the
    result is formed with int64 arithmetic and is exact; otherwise np.convolve in float64."""
    s = np.asarray(sig, dtype=np.float64)
    h = np.asarray(reverb, dtype=np.float64)
    hi = h * 32768.0
    if exact and np.all(s == np.rint(s)) and np.all(hi == np.rint(hi)) and \
            np.abs(s).max(initial=0) * np.abs(hi).sum() < 2.0 ** 62:
        return np.convolve(s.astype(np.int64), hi.astype(np.int64)).astype(np.float64) / 32768.0
    return np.convolve(s, h)


def alignment_lags(sig, full, R):
    """xx[d] for d = 0 .. R-1 (direct inner products)."""
    s = np.asarray(sig, dtype=np.float64)
    T = len(s)
    return np.array([np.dot(s, full[d:d + T]) for d in range(R)])


def add_reverb(sig, reverb, return_index=False):
    s = np.asarray(sig, dtype=np.float64)
    h = np.asarray(reverb, dtype=np.float64)
    T, R = len(s), len(h)
    full = full_convolution(s, h)
    xx = alignment_lags(s, full, R)
    best = xx.max()
    d_star = int(np.flatnonzero(xx == best)[-1])
    ind = d_star + 1
    out = full[ind:ind + T]
    return (out, ind) if return_index else out


def rir_channel(rir_wav_data):
    """computeFDLPSpectrogram.py:78-79: second channel of the recording, divided by 2^15."""
    return np.asarray(rir_wav_data)[:, 1] / np.power(2, 15)
