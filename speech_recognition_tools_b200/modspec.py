"""FDLP modulation-spectrum features: the host-side mirror of
`/root/reference/src/featgen/computeModulationSpectrum.py` (`getFeats`, lines 31-203), real-modulation path and
`--complex_modulation`.

Per utterance the reference takes 0.5 s frames every 10 ms (`getFrames(signal, srate, frate, fduration,
np.hanning)`), their DCT, and for every (frame, band) the sub-band LPC model and its cepstrum
(`computeLpcFast`, `computeModSpecFromLpc`); the feature vector of a frame is the selected cepstral
coefficients `mod_spec[coeff_0-1:coeff_n]` of all bands, band-major.  That is K1-K4 of the FDLP-spectrogram
path with a different window, a segment shift of `int(srate / frate)` samples and the cepstrum instead of the
envelope, so it runs on the same plan (`cepstrum_mode`) and kernels; `fdlp_modspec_forward` adds the gather.

`--complex_modulation` (:45-48, :150-152, :164-173) replaces the DCT by the first half of the frames' inverse FFT, the
filterbank by one over int(fduration*srate) points, and the real LPC model by a complex one: `fdlp_stage_frames` (the
plan's framing), a cuFFT transform through torch.fft (library code, like the reference's scipy call), then
`fdlp_modspec_complex_forward` (csrc/modspec_cplx.cuh: band product, circular autocorrelation, Hermitian Levinson-Durbin,
complex cepstrum, selection).

There is no CPU fallback: the extractor needs libfdlp_b200.so and a CUDA device."""
import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence, Union

import numpy as np
import torch

from . import _lib as L
from .fdlp import BatchLayout, FdlpExtractor, parse_fbank_type, segment_count


@dataclass
class ModSpecConfig:
    """argparse surface of computeModulationSpectrum.py:206-224 (same names and defaults)."""
    nfilters: int = 15
    coeff_0: int = 5
    coeff_n: int = 30
    keep_even: bool = False
    order: int = 50
    fduration: float = 0.5
    frate: int = 100
    fbank_type: str = "mel,1"
    no_window: bool = False
    complex_modulation: bool = False
    compensate_noise: bool = False
    absolute_value: bool = False
    set_unity_gain: bool = False  # parsed by the reference but never used (computeModulationSpectrum.py:31-203)
    srate: int = 16000
    tail_eps: float = 1e-12


def feature_selection(cfg: ModSpecConfig):
    """Cepstrum indices and scales of the emitted coefficients (computeModulationSpectrum.py:64-87,169-186)."""
    coeff_num = cfg.coeff_n - cfg.coeff_0 + 1
    if cfg.coeff_0 < 1 or coeff_num < 1:
        raise ValueError("need 1 <= coeff_0 <= coeff_n")
    idx = np.arange(cfg.coeff_0 - 1, cfg.coeff_n)
    scale = np.ones(cfg.coeff_n)
    if cfg.complex_modulation and cfg.keep_even and not cfg.absolute_value:
        # :64-71 sizes each_feat from coeff_num, :171-172 fills it with 2*coeff_num values: the reference stops with
        # "could not broadcast input array" on the first frame
        raise ValueError("--keep_even with --complex_modulation needs --absolute_value (the reference fails there: "
                         "computeModulationSpectrum.py:64-71 vs :171-185)")
    if cfg.compensate_noise:
        fmax = coeff_num / cfg.fduration if cfg.complex_modulation else coeff_num / (2 * cfg.fduration)   # :80-86
        scale = np.linspace(0, fmax, cfg.coeff_n)
    if cfg.keep_even:
        idx = idx[1::2] if cfg.coeff_0 % 2 == 0 else idx[0::2]
    return idx.astype(np.int32), scale[idx].astype(np.float64)


class ModSpecExtractor:
    """One plan on one device; call it with a list of mono waveforms -> list of [frame_num, nfilters*feat_len]."""

    def __init__(self, cfg: Optional[ModSpecConfig] = None, device=None, **kwargs):
        self.cfg = cfg if cfg is not None else ModSpecConfig(**kwargs)
        if not torch.cuda.is_available():
            raise RuntimeError("ModSpecExtractor needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self._lib = L.load()
        kind, fargs = parse_fbank_type(self.cfg.fbank_type)
        self._idx, self._scale = feature_selection(self.cfg)
        self.feat_len = int(len(self._idx))     # per band; twice that for complex cepstra without --absolute_value
        if self.cfg.complex_modulation and not self.cfg.absolute_value:
            self.feat_len *= 2
        self._weights = np.ones(self.cfg.coeff_n, dtype=np.float64)
        prm = L.FdlpParams()
        prm.srate, prm.nfilters, prm.order = self.cfg.srate, self.cfg.nfilters, self.cfg.order
        prm.coeff_num, prm.frate = self.cfg.coeff_n, self.cfg.frate
        prm.fduration, prm.overlap_fraction = float(self.cfg.fduration), 0.25
        prm.fbank_kind = kind
        prm.fbank_args = (C.c_double * 5)(*fargs)
        prm.weights = self._weights.ctypes.data_as(C.POINTER(C.c_double))
        prm.custom_fbank = None
        prm.dct_fp32 = 0
        prm.tail_eps = float(self.cfg.tail_eps)
        prm.window_kind = L.FDLP_WINDOW_SQUARE if self.cfg.no_window else L.FDLP_WINDOW_HANNING
        prm.seg_shift = int(self.cfg.srate / self.cfg.frate)  # getFrames: int(srate / frate)
        prm.cepstrum_mode = 1
        self._plan = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_plan_create(C.byref(prm), C.byref(self._plan)))
        g = L.FdlpGeometry()
        L.check(self._lib.fdlp_plan_get_geometry(self._plan, C.byref(g)))
        self.geom = {n: getattr(g, n) for n, _ in L.FdlpGeometry._fields_}
        self._idx_dev = torch.from_numpy(self._idx).to(self.device)
        self._scale_dev = torch.from_numpy(self._scale).to(self.device)
        if self.cfg.complex_modulation:
            self._init_complex(prm)

    def _init_complex(self, prm):
        """Filterbank of the complex path: createFbank(nfilters, int(fduration*srate), ...)[:, :-1] (:45-48, :162); its
        taps above tail_eps * max are the bands' supports."""
        N = self.geom["N"]
        if N % 2 or self.cfg.order + 2 > N // 2:
            raise ValueError("--complex_modulation needs an even int(fduration*srate) and order <= it/2 - 2")
        full = np.empty((self.cfg.nfilters, N // 2 + 1), dtype=np.float64)
        args = np.array([prm.fbank_args[i] for i in range(5)], dtype=np.float64)
        L.check(self._lib.fdlp_make_fbank(prm.fbank_kind, args.ctypes.data_as(C.c_void_p), self.cfg.nfilters, N,
                                          self.cfg.srate, full.ctypes.data_as(C.c_void_p)))
        fb = np.ascontiguousarray(full[:, :-1])
        keep = np.abs(fb) > self.cfg.tail_eps * np.abs(fb).max(axis=1, keepdims=True)
        if not keep.any(axis=1).all():
            raise ValueError("a band of the filterbank is empty")
        lo = keep.argmax(axis=1).astype(np.int32)
        hi = (keep.shape[1] - keep[:, ::-1].argmax(axis=1)).astype(np.int32)
        self._cfb = torch.from_numpy(fb).to(self.device)
        self._clo, self._chi = torch.from_numpy(lo).to(self.device), torch.from_numpy(hi).to(self.device)
        self._cwidth = int((hi - lo).max())

    def close(self):
        if getattr(self, "_plan", None) and self._plan.value:
            with torch.cuda.device(self.device):
                self._lib.fdlp_plan_destroy(self._plan)
            self._plan = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def layout(self, lengths: Sequence[int]) -> BatchLayout:
        g = self.geom
        lengths = np.asarray(lengths, dtype=np.int64)
        n_utt = len(lengths)
        if n_utt and lengths.min() < 2:
            raise ValueError("utterances need at least 2 samples")
        nsegs = np.array([segment_count(int(T), g["N"], g["S"]) for T in lengths], dtype=np.int64)
        seg_off = np.concatenate([[0], np.cumsum(nsegs)])
        return BatchLayout(
            lengths=lengths,
            utt_wav_off=np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64),
            utt_seg_off=seg_off.astype(np.int32),
            utt_frame_off=seg_off.astype(np.int64),  # one output row per segment
            seg_utt=np.repeat(np.arange(n_utt, dtype=np.int32), nsegs).astype(np.int32),
            seg_ptr=np.zeros(int(nsegs.sum()), dtype=np.int32),
            n_utt=n_utt, n_seg=int(nsegs.sum()), n_frames=int(nsegs.sum()),
            max_utt_frames=int(nsegs.max()) if n_utt else 0)

    def forward_packed(self, wav_dev: torch.Tensor, lay: BatchLayout, out: Optional[torch.Tensor] = None,
                       status: Optional[torch.Tensor] = None):
        """Packed device batch -> (feats [n_seg, nfilters*feat_len] float32, per-utterance status)."""
        lay.to(self.device)
        D = self.cfg.nfilters * self.feat_len
        if out is None:
            out = torch.empty((lay.n_seg, D), dtype=torch.float32, device=self.device)
        if status is None:
            status = torch.zeros(max(lay.n_utt, 1), dtype=torch.int32, device=self.device)
        b = L.FdlpBatch()
        d = lay.dev
        b.wav = wav_dev.data_ptr()
        b.wav_dtype = {torch.int16: L.FDLP_WAV_INT16, torch.float32: L.FDLP_WAV_FLOAT32,
                       torch.float64: L.FDLP_WAV_FLOAT64}[wav_dev.dtype]  # float64: the reverberated waveform
        b.n_utt, b.n_seg, b.n_frames, b.max_utt_frames = lay.n_utt, lay.n_seg, lay.n_frames, lay.max_utt_frames
        b.utt_wav_off = d["utt_wav_off"].data_ptr()
        b.utt_seg_off = d["utt_seg_off"].data_ptr()
        b.utt_frame_off = d["utt_frame_off"].data_ptr()
        b.seg_utt = d["seg_utt"].data_ptr()
        b.seg_ptr = d["seg_ptr"].data_ptr()
        b.noise_mode = L.FDLP_NOISE_NONE
        b.status = status.data_ptr()
        stream = torch.cuda.current_stream(self.device).cuda_stream
        if self.cfg.complex_modulation:
            N = self.geom["N"]
            frames = torch.empty((lay.n_seg, N), dtype=torch.float64, device=self.device)
            with torch.cuda.device(self.device):
                L.check(self._lib.fdlp_stage_frames(self._plan, C.byref(b), C.c_void_p(frames.data_ptr()), C.c_void_p(stream)))
                # ifft of a real frame: conj(rfft) / N; bins 0 .. N/2-1 (computeModulationSpectrum.py:150-151)
                spec = torch.fft.rfft(frames, dim=1).conj().resolve_conj()
                spec.mul_(1.0 / N)
                L.check(self._lib.fdlp_modspec_complex_forward(
                    C.c_void_p(spec.data_ptr()), spec.shape[1], lay.n_seg, C.c_void_p(d["seg_utt"].data_ptr()), N // 2,
                    C.c_void_p(self._cfb.data_ptr()), C.c_void_p(self._clo.data_ptr()), C.c_void_p(self._chi.data_ptr()),
                    self._cwidth, self.cfg.nfilters, self.cfg.order, self.cfg.coeff_n,
                    C.c_void_p(self._idx_dev.data_ptr()), C.c_void_p(self._scale_dev.data_ptr()), int(len(self._idx)),
                    1 if self.cfg.absolute_value else 0, C.c_void_p(out.data_ptr()), C.c_void_p(status.data_ptr()),
                    C.c_void_p(stream)))
            return out, status
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_modspec_forward(self._plan, C.byref(b), C.c_void_p(self._idx_dev.data_ptr()),
                                                   C.c_void_p(self._scale_dev.data_ptr()), self.feat_len,
                                                   1 if self.cfg.absolute_value else 0,
                                                   C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
        return out, status

    def __call__(self, wavs: Sequence[Union[np.ndarray, torch.Tensor]], return_status: bool = False,
                 max_segments: int = 20000, reverb=None, unity_gain: bool = False):
        """List of mono waveforms -> list of [frame_num_u, nfilters*feat_len] float32 device tensors.  Utterances are
        processed in runs of at most `max_segments` segments (a 10 s utterance has 1000 half-second segments).
        `reverb`: impulse response or reverb.Reverberator for --add_reverb (computeModulationSpectrum.py:94-105,146-148:
        addReverb before framing; the float64 result goes to the framing kernel).  `unity_gain`: the LPC gain is taken as
        1, i.e. c0 = log(sqrt(1)) = 0 (computeModulationSpectrum_segments.py:96-97; the other coefficients do not
        depend on the gain)."""
        if len(wavs) == 0:
            return ([], torch.zeros(0, dtype=torch.int32, device=self.device)) if return_status else []
        outs, sts = [], []
        i = 0
        while i < len(wavs):
            j, nseg = i, 0
            while j < len(wavs):
                n = segment_count(int(len(wavs[j])), self.geom["N"], self.geom["S"])
                if j > i and nseg + n > max_segments:
                    break
                nseg += n
                j += 1
            host, lengths = FdlpExtractor.pack(wavs[i:j])
            wav_dev = host.to(self.device, non_blocking=True)
            if reverb is not None:
                from .reverb import reverberate_packed
                wav_dev, lengths = reverberate_packed(reverb, wav_dev, lengths, self.device, torch.float64)
            lay = self.layout(lengths)
            feats, st = self.forward_packed(wav_dev, lay)
            if unity_gain and 0 in self._idx:
                pos = int(np.flatnonzero(self._idx == 0)[0])
                v = feats.view(lay.n_seg, self.cfg.nfilters, self.feat_len)
                v[:, :, pos] = 0.0
                if self.cfg.complex_modulation and not self.cfg.absolute_value:
                    v[:, :, len(self._idx) + pos] = 0.0
            outs.extend(feats[int(lay.utt_seg_off[u]):int(lay.utt_seg_off[u + 1])] for u in range(lay.n_utt))
            sts.append(st[:lay.n_utt])
            i = j
        if return_status:
            return outs, torch.cat(sts)
        return outs
