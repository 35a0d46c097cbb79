"""Host side of the mel-spectrogram baseline path (reference: compute_mel_spectrum,
src/featgen/computeMelSpectrum.py:40-170).  Same contract as fdlp.py: the host packs waveforms and does
the frame-count bookkeeping; every sample and feature is computed by libfdlp_b200.so kernels."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib as L
from .fdlp import FdlpExtractor, parse_fbank_type


@dataclass
class MelConfig:
    """computeMelSpectrum.py:20-37 minus I/O options."""
    nfilters: int = 23
    nfft: int = 1024
    fduration: float = 0.02
    frate: int = 100
    fbank_type: str = "mel,1"
    spectrum_type: str = "log"
    srate: int = 16000

    @classmethod
    def from_args(cls, args) -> "MelConfig":
        return cls(nfilters=args.nfilters, nfft=args.nfft, fduration=args.fduration, frate=args.frate,
                   fbank_type=args.fbank_type, spectrum_type=args.spectrum_type)


def mel_frame_count(T: int, flen: int, shift: int) -> int:
    """Frames getFrames yields (features.py:134-154) for a frame length `flen` and shift `shift`."""
    lim = T - 1 if flen % 2 == 0 else T
    return max(0, -(-lim // shift))


@dataclass
class MelLayout:
    lengths: np.ndarray
    utt_wav_off: np.ndarray
    utt_frame_off: np.ndarray
    n_utt: int = 0
    n_frames: int = 0
    dev: dict = field(default_factory=dict)

    def to(self, device):
        if self.dev.get("device") == device:
            return self
        self.dev = {"device": device,
                    "utt_wav_off": torch.from_numpy(self.utt_wav_off).to(device),
                    "utt_frame_off": torch.from_numpy(self.utt_frame_off).to(device)}
        return self


class MelExtractor:
    def __init__(self, cfg: Optional[MelConfig] = None, device=None, **kwargs):
        self.cfg = cfg if cfg is not None else MelConfig(**kwargs)
        if self.cfg.spectrum_type not in ("log", "power"):
            raise ValueError("Spectrum type not supported! ")
        if not torch.cuda.is_available():
            raise RuntimeError("MelExtractor needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self._lib = L.load()
        kind, fargs = parse_fbank_type(self.cfg.fbank_type)
        prm = L.MelParams()
        prm.srate, prm.nfilters, prm.nfft, prm.frate = self.cfg.srate, self.cfg.nfilters, self.cfg.nfft, self.cfg.frate
        prm.fduration = float(self.cfg.fduration)
        prm.fbank_kind = kind
        prm.fbank_args = (C.c_double * 5)(*fargs)
        prm.custom_fbank = None
        prm.power = 1 if self.cfg.spectrum_type == "power" else 0
        self._plan = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(self._lib.mel_plan_create(C.byref(prm), C.byref(self._plan)))
        self.flen = int(self.cfg.srate * self.cfg.fduration)
        self.shift = int(self.cfg.srate / self.cfg.frate)

    def close(self):
        if getattr(self, "_plan", None) and self._plan.value:
            with torch.cuda.device(self.device):
                self._lib.mel_plan_destroy(self._plan)
            self._plan = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def filterbank(self) -> np.ndarray:
        out = np.empty((self.cfg.nfilters, self.cfg.nfft // 2 + 1), dtype=np.float64)
        L.check(self._lib.mel_plan_get_fbank(self._plan, out.ctypes.data_as(C.c_void_p)))
        return out

    def layout(self, lengths: Sequence[int]) -> MelLayout:
        lengths = np.asarray(lengths, dtype=np.int64)
        if len(lengths) and lengths.min() < 2:
            raise ValueError("utterances need at least 2 samples")
        nfr = np.array([mel_frame_count(int(T), self.flen, self.shift) for T in lengths], dtype=np.int64)
        return MelLayout(lengths=lengths,
                         utt_wav_off=np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64),
                         utt_frame_off=np.concatenate([[0], np.cumsum(nfr)]).astype(np.int64),
                         n_utt=len(lengths), n_frames=int(nfr.sum()))

    def forward_packed(self, wav_dev: torch.Tensor, lay: MelLayout, feats: Optional[torch.Tensor] = None,
                       cmvn: Optional[torch.Tensor] = None):
        lay.to(self.device)
        if feats is None:
            feats = torch.empty((lay.n_frames, self.cfg.nfilters), dtype=torch.float32, device=self.device)
        b = L.MelBatch()
        b.wav = wav_dev.data_ptr()
        b.wav_dtype = L.FDLP_WAV_INT16 if wav_dev.dtype == torch.int16 else L.FDLP_WAV_FLOAT32
        b.n_utt, b.n_frames = lay.n_utt, lay.n_frames
        b.utt_wav_off = lay.dev["utt_wav_off"].data_ptr()
        b.utt_frame_off = lay.dev["utt_frame_off"].data_ptr()
        b.feats = feats.data_ptr()
        b.cmvn_acc = cmvn.data_ptr() if cmvn is not None else None
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            L.check(self._lib.mel_forward(self._plan, C.byref(b), C.c_void_p(stream)))
        return feats

    def mix_noise(self, wav_dev: torch.Tensor, lengths, noise):
        """--add_noise on the mel path (computeMelSpectrum.py:134-141): `noise` is 'diff' (13-tap FIR,
        fdlp_diff_filter) or a fdlp.NoiseSpec (add_noise_to_wav, features.py:24-31: fdlp_noise_mix).  Returns the float32
        packed batch the framing kernel then reads."""
        from .fdlp import noise_to_device
        lengths = np.asarray(lengths, dtype=np.int64)
        off = torch.from_numpy(np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64)).to(self.device)
        out = torch.empty(wav_dev.numel(), dtype=torch.float32, device=self.device)
        dt = L.FDLP_WAV_INT16 if wav_dev.dtype == torch.int16 else L.FDLP_WAV_FLOAT32
        if wav_dev.dtype not in (torch.int16, torch.float32):
            raise TypeError("waveforms must be int16 or float32")
        stream = torch.cuda.current_stream(self.device).cuda_stream
        max_len = int(lengths.max()) if len(lengths) else 0
        with torch.cuda.device(self.device):
            if isinstance(noise, str):
                if noise != "diff":
                    raise ValueError("noise must be a NoiseSpec or the string 'diff'")
                L.check(self._lib.fdlp_diff_filter(C.c_void_p(wav_dev.data_ptr()), dt, C.c_void_p(off.data_ptr()),
                                                    len(lengths), max_len, C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
            else:
                nd = noise_to_device(noise, lengths, self.device)
                alpha = torch.empty(max(len(lengths), 1), dtype=torch.float64, device=self.device)
                L.check(self._lib.fdlp_noise_mix(C.c_void_p(wav_dev.data_ptr()), dt, C.c_void_p(off.data_ptr()),
                                                  len(lengths), max_len, C.c_void_p(nd["samples"].data_ptr()),
                                                  C.c_void_p(nd["off"].data_ptr()), C.c_void_p(nd["snr"].data_ptr()),
                                                  nd["mode"], C.c_void_p(alpha.data_ptr()), C.c_void_p(out.data_ptr()),
                                                  C.c_void_p(stream)))
        return out

    def __call__(self, wavs, cmvn: Optional[torch.Tensor] = None, reverb=None, noise=None):
        """`noise`: 'diff' or a fdlp.NoiseSpec for --add_noise (applied first, computeMelSpectrum.py:134-141);
        `reverb`: impulse response or reverb.Reverberator for --add_reverb (:143-145); the augmented waveform reaches the
        framing kernel as float32."""
        if len(wavs) == 0:
            return []
        host, lengths = FdlpExtractor.pack(wavs)
        wav_dev = host.to(self.device, non_blocking=True)
        if noise is not None:
            wav_dev = self.mix_noise(wav_dev, lengths, noise)
        if reverb is not None:
            from .reverb import reverberate_packed
            wav_dev, lengths = reverberate_packed(reverb, wav_dev, lengths, self.device, torch.float32)
        lay = self.layout(lengths)
        feats = self.forward_packed(wav_dev, lay, cmvn=cmvn)
        return [feats[int(lay.utt_frame_off[u]):int(lay.utt_frame_off[u + 1])] for u in range(lay.n_utt)]


def mel_spectrogram(wavs, *, device=None, **cfg_kwargs):
    ex = MelExtractor(MelConfig(**cfg_kwargs), device=device)
    try:
        return ex(wavs)
    finally:
        ex.close()
