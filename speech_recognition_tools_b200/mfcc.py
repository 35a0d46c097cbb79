"""Host side of the MFCC path (reference: extractMelEnergyFeats, src/featgen/computeMfccFeatures.py:57-133).
Same contract as mel.py: the host packs waveforms and does the frame-count bookkeeping; every sample and
feature is computed by libfdlp_b200.so kernels (csrc/mfcc.cuh)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib as L
from .fdlp import FdlpExtractor, NoiseSpec
from .mel import MelLayout, mel_frame_count

NCEP = 13  # computeMfccFeatures.py:127  mfcc_feats[:, 0:13]


@dataclass
class MfccConfig:
    """computeMfccFeatures.py:139-150 minus I/O options."""
    nfilters: int = 30
    nfft: int = 1024
    fduration: float = 0.02
    frate: int = 100
    context: Optional[int] = None
    srate: int = 16000

    @classmethod
    def from_args(cls, args) -> "MfccConfig":
        return cls(nfilters=args.nfilters, nfft=args.nfft, fduration=args.fduration, frate=args.frate,
                   context=args.context)


class MfccExtractor:
    def __init__(self, cfg: Optional[MfccConfig] = None, device=None, **kwargs):
        self.cfg = cfg if cfg is not None else MfccConfig(**kwargs)
        if not torch.cuda.is_available():
            raise RuntimeError("MfccExtractor needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self._lib = L.load()
        prm = L.MfccParams()
        prm.srate, prm.nfilters, prm.nfft, prm.frate = self.cfg.srate, self.cfg.nfilters, self.cfg.nfft, self.cfg.frate
        prm.fduration = float(self.cfg.fduration)
        prm.ncep = NCEP
        self._plan = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(self._lib.mfcc_plan_create(C.byref(prm), C.byref(self._plan)))
        self.flen = int(self.cfg.srate * self.cfg.fduration)
        self.shift = int(self.cfg.srate / self.cfg.frate)
        self.context = int(self.cfg.context) if self.cfg.context else 0  # `if args.context:` (:129)
        self.dim = NCEP * (2 * self.context + 1)

    def set_tensor_cores(self, enable: bool = True) -> bool:
        """Choose between the tensor-core kernel (default where the configuration fits: nfft 1024, 20 ms frames, at most
        32 bands) and the FP32-pipe kernel; returns whether the tensor-core kernel will run."""
        return bool(L.check_count(self._lib.mfcc_plan_set_tensor_cores(self._plan, 1 if enable else 0)))

    def close(self):
        if getattr(self, "_plan", None) and self._plan.value:
            with torch.cuda.device(self.device):
                self._lib.mfcc_plan_destroy(self._plan)
            self._plan = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def layout(self, lengths: Sequence[int]) -> MelLayout:
        lengths = np.asarray(lengths, dtype=np.int64)
        if len(lengths) and lengths.min() < 2:
            raise ValueError("utterances need at least 2 samples")
        nfr = np.array([mel_frame_count(int(T), self.flen, self.shift) for T in lengths], dtype=np.int64)
        return MelLayout(lengths=lengths,
                         utt_wav_off=np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64),
                         utt_frame_off=np.concatenate([[0], np.cumsum(nfr)]).astype(np.int64),
                         n_utt=len(lengths), n_frames=int(nfr.sum()))

    def _noise_to_device(self, noise: Optional[NoiseSpec], lay: MelLayout):
        """--add_noise 'type,snr' (computeMfccFeatures.py:45-52): offsets floor(rand * (len(noise) - T)); the
        energies are those of the normalised float signals, so NoiseSpec.mode is not consulted."""
        if noise is None:
            return None
        ns = noise.samples
        if not isinstance(ns, torch.Tensor):
            ns = torch.from_numpy(np.ascontiguousarray(ns))
        if ns.dtype != torch.int16:
            raise TypeError("noise recording must be int16 (scipy.io.wavfile.read of noises/<type>.wav)")
        rand = noise.rand
        if rand is None or isinstance(rand, (int, np.integer)):
            rand = np.random.RandomState(0 if rand is None else int(rand))
        if isinstance(rand, np.random.RandomState):
            draws = np.array([rand.rand() for _ in range(lay.n_utt)])
        else:
            draws = np.asarray(list(rand), dtype=np.float64)
        Tn = int(ns.numel())
        if lay.n_utt and Tn <= int(lay.lengths.max()):
            raise ValueError("noise recording must be longer than every utterance")
        off = np.floor(draws * (Tn - lay.lengths)).astype(np.int64)
        snr = np.broadcast_to(np.asarray(noise.snr_db, dtype=np.float64), (lay.n_utt,)).copy()
        return dict(samples=ns.to(self.device, non_blocking=True), off=torch.from_numpy(off).to(self.device),
                    snr=torch.from_numpy(snr).to(self.device))

    def forward_packed(self, wav_dev: torch.Tensor, lay: MelLayout, feats: Optional[torch.Tensor] = None,
                       noise_dev=None):
        lay.to(self.device)
        if feats is None:
            feats = torch.empty((lay.n_frames, self.dim), dtype=torch.float32, device=self.device)
        if wav_dev.dtype not in (torch.int16, torch.float32):
            raise TypeError("waveforms must be int16 or float32")
        b = L.MfccBatch()
        b.wav = wav_dev.data_ptr()
        b.wav_dtype = L.FDLP_WAV_INT16 if wav_dev.dtype == torch.int16 else L.FDLP_WAV_FLOAT32
        b.n_utt, b.n_frames = lay.n_utt, lay.n_frames
        b.utt_wav_off = lay.dev["utt_wav_off"].data_ptr()
        b.utt_frame_off = lay.dev["utt_frame_off"].data_ptr()
        b.feats = feats.data_ptr()
        b.context = self.context
        if noise_dev is not None:
            b.noise = noise_dev["samples"].data_ptr()
            b.utt_noise_off = noise_dev["off"].data_ptr()
            b.utt_snr_db = noise_dev["snr"].data_ptr()
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            L.check(self._lib.mfcc_forward(self._plan, C.byref(b), C.c_void_p(stream)))
        return feats

    def __call__(self, wavs, noise: Optional[NoiseSpec] = None, reverb=None):
        """List of mono waveforms (int16 / float32 as read from the WAV) -> list of [F_u, 13 (2 context + 1)]
        float32 device tensors.  `reverb`: impulse response or reverb.Reverberator (computeMfccFeatures.py:118-120;
        the reference reverberates signal / 2^15, which is the reverberated signal / 2^15 exactly)."""
        if len(wavs) == 0:
            return []
        host, lengths = FdlpExtractor.pack(wavs)
        wav_dev = host.to(self.device, non_blocking=True)
        if reverb is not None:
            from .reverb import reverberate_packed
            nd = None
            if noise is not None:
                # the mix happens before the reverberation (:115-120).  Its coefficient is a ratio of mean squares, the
                # same for signal / 2^15 and noise / 2^15 as for the raw samples: the reverberation kernel's
                # float-energy mix on the raw samples, times 2^-15 in the MFCC kernel, is the reference's waveform
                if wav_dev.dtype != torch.int16:
                    raise NotImplementedError("noise mixing followed by reverberation needs int16 samples")
                nd = dict(self._noise_to_device(noise, self.layout(lengths)), mode=L.FDLP_NOISE_FLOAT)
                noise = None
            wav_dev, lengths = reverberate_packed(reverb, wav_dev, lengths, self.device, torch.float32, nd)
        lay = self.layout(lengths)
        noise_dev = self._noise_to_device(noise, lay)
        feats = self.forward_packed(wav_dev, lay, noise_dev=noise_dev)
        return [feats[int(lay.utt_frame_off[u]):int(lay.utt_frame_off[u + 1])] for u in range(lay.n_utt)]


def mfcc(wavs, *, device=None, **cfg_kwargs):
    ex = MfccExtractor(MfccConfig(**cfg_kwargs), device=device)
    try:
        return ex(wavs)
    finally:
        ex.close()
