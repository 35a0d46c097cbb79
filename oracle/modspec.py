"""CPU oracle of the FDLP modulation-spectrum extractor -- TEST INFRASTRUCTURE ONLY.

numpy/scipy restatement of /root/reference/src/featgen/computeModulationSpectrum.py `getFeats` (lines 31-203):
`modulation_spectrum` is the real-modulation path, `modulation_spectrum_complex` the `--complex_modulation` one.
It reuses the helpers of oracle/fdlp.py, which restate features.py.  Pinned against outputs of the unmodified reference: tests/golden/modspec_*.npz
(tests/golden/make_golden_modspec.py), replayed by tests/test_oracle_golden.py.

Only tests/, __graft_entry__.smoke() and bench.py may import this package."""
import numpy as np
import scipy.fftpack

from . import fdlp as ofdlp


def sq_wind(N):
    """computeModulationSpectrum.py:27-28"""
    return np.ones(N)


def feature_selection(coeff_0, coeff_n, fduration, keep_even=False, compensate_noise=False):
    """Indices into the cepstrum and per-index scale of the emitted coefficients
    (computeModulationSpectrum.py:64-87,169-186): mod_spec[coeff_0-1:coeff_n], optionally times faxis, optionally
    every other one."""
    coeff_num = coeff_n - coeff_0 + 1
    idx = np.arange(coeff_0 - 1, coeff_n)
    scale = np.ones(coeff_n)
    if compensate_noise:
        fmax = coeff_num / (2 * fduration)            # :84-86
        scale = np.linspace(0, fmax, coeff_n)
    if keep_even:
        idx = idx[1::2] if coeff_0 % 2 == 0 else idx[0::2]   # :181-185
    return idx, scale[idx]


def modulation_spectrum(signal, *, nfilters=15, coeff_0=5, coeff_n=30, order=50, fduration=0.5, frate=100,
                        fbank_type="mel,1", keep_even=False, no_window=False, compensate_noise=False,
                        absolute_value=False, srate=16000, fbank=None):
    """Returns float64 [frame_num, nfilters * feat_len] (computeModulationSpectrum.py:143-198)."""
    window = sq_wind if no_window else np.hanning                      # :31, :88-90
    if fbank is None:
        fbank = ofdlp.make_fbank(fbank_type, nfilters, int(2 * fduration * srate), srate)   # :45-63
    frames = ofdlp.get_frames(signal, srate, frate, fduration, window)                        # :143-144
    cos_trans = scipy.fftpack.dct(frames) / np.sqrt(2 * int(srate * fduration))             # :150
    idx, scale = feature_selection(coeff_0, coeff_n, fduration, keep_even, compensate_noise)
    frame_num = cos_trans.shape[0]
    feats = np.zeros((frame_num, nfilters * len(idx)))
    for i in range(frame_num):
        band = fbank[:, 0:-1] * cos_trans[i, :][None, :]                                     # :162-163
        r = ofdlp.autocorr_circular(band)
        lpcs = np.zeros((nfilters, order + 1))
        ggs = np.zeros(nfilters)
        for j in range(nfilters):
            lpcs[j], ggs[j] = ofdlp.compute_lpc_fast(band[j], order)                          # :174
        ms = ofdlp.modspec_batch(ggs, lpcs, coeff_n)                                          # :175
        sel = ms[:, idx] * scale[None, :]                                                     # :176-181
        if absolute_value:
            sel = np.abs(sel)
        feats[i, :] = sel.reshape(-1)                                                         # :189-191
    return feats


def compute_lpc_complex(band, order):
    """features.py:222-230 with keepreal=False: complex circular autocorrelation, Hermitian Toeplitz solve
    (scipy takes the first row as conj(first column) when only the column is given)."""
    X = np.fft.fft(band, len(band))
    y = np.fft.ifft(X * np.conj(X))
    xlpc = scipy.linalg.solve_toeplitz(y[0:order], -y[1:order + 1])
    xlpc = np.append(1, xlpc)
    gg = y[0] + np.sum(xlpc * y[1:order + 2])
    return xlpc, gg


def modspec_from_lpc_complex(gg, xlpc, lim):
    """features.py:233-246 on complex input (the recursion is the same, log and sqrt take their principal values)."""
    x = np.array(xlpc, dtype=np.complex128)
    x[1:] = -x[1:]
    cep = np.zeros(lim, dtype=np.complex128)
    cep[0] = np.log(np.sqrt(gg))
    cep[1] = x[1]
    if x.shape[0] < lim:
        x = np.append(x, np.zeros(int(lim - x.shape[0] + 1)))
    for n in range(2, lim):
        aa = np.arange(1, n) / n
        cep[n] = np.sum(aa * np.flipud(x[1:n]) * cep[1:n]) + x[n]
    return cep


def modulation_spectrum_complex(signal, *, nfilters=15, coeff_0=5, coeff_n=30, order=50, fduration=0.5, frate=100,
                                fbank_type="mel,1", keep_even=False, no_window=False, compensate_noise=False,
                                absolute_value=False, srate=16000):
    """--complex_modulation (computeModulationSpectrum.py:45-48, :75-87, :150-152, :164-173): the frames' inverse
    FFT (first half) replaces the DCT, the filterbank has int(fduration*srate)/2 taps, the LPC model and its cepstrum
    are complex; a frame's band contributes |c| (--absolute_value) or real parts then imaginary parts."""
    window = sq_wind if no_window else np.hanning
    dur = int(fduration * srate)                                                              # :45-46
    fbank = ofdlp.make_fbank(fbank_type, nfilters, dur, srate)
    frames = ofdlp.get_frames(signal, srate, frate, fduration, window)
    spec = scipy.fftpack.ifft(frames)[:, :int(fduration * srate / 2)]                         # :150-151
    coeff_num = coeff_n - coeff_0 + 1
    scale = np.linspace(0, coeff_num / fduration, coeff_n) if compensate_noise else np.ones(coeff_n)   # :80-83
    if keep_even and not absolute_value:
        raise ValueError("the reference cannot run --keep_even with --complex_modulation unless --absolute_value "
                         "is given (each_feat has coeff_num/2 columns, temp2[::2] has coeff_num)")
    rows = []
    for i in range(spec.shape[0]):
        each = []
        for j in range(nfilters):
            band = fbank[j, 0:-1] * spec[i, :]
            xlpc, gg = compute_lpc_complex(band, order)
            ms = modspec_from_lpc_complex(gg, xlpc, coeff_n) * scale
            sel = ms[coeff_0 - 1:coeff_n]
            t = np.abs(sel) if absolute_value else np.append(np.real(sel), np.imag(sel))      # :168-172
            if keep_even:
                t = t[1::2] if coeff_0 % 2 == 0 else t[0::2]                                  # :181-185
            each.append(t)
        rows.append(np.concatenate(each))
    return np.array(rows)
