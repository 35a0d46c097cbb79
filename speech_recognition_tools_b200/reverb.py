"""--add_reverb: the room-impulse-response augmentation of the extractors, on the GPU.

Mirrors `addReverb(sig, reverb)` (reference src/featgen/features.py:110-115) and the impulse-response handling of
computeFDLPSpectrogram.py:75-91 (`rir = read('./RIR/RIR_<room>_AnglA.wav')[1][:, 1] / 2**15`):

    out  = np.convolve(sig, reverb)
    xxc  = np.correlate(sig, out, 'valid')
    indM = len(xxc) - np.argmax(xxc)
    out  = out[indM:indM + len(sig)]

The kernels (csrc/reverb.cuh) evaluate both the convolution and the alignment correlation as partitioned block
FFTs in float64.  There is no CPU fallback."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence, Union

import numpy as np
import torch

from . import _lib as L

# computeFDLPSpectrogram.py:76-87 (computeMfccFeatures.py:81-88 knows only the first and the last)
RIR_FILES = {
    "small_room": "./RIR/RIR_SmallRoom1_near_AnglA.wav",
    "medium_room": "./RIR/RIR_MediumRoom1_far_AnglA.wav",
    "large_room": "./RIR/RIR_LargeRoom1_far_AnglA.wav",
}


def load_rir(add_reverb: str, rir_dir: Optional[str] = None) -> Optional[np.ndarray]:
    """The reference's --add_reverb handling: None for 'clean' (or no flag), ValueError for an unknown room, else the
    second channel of the recording divided by 2^15 (float64).  `rir_dir` replaces the reference's hard-wired ./RIR."""
    if not add_reverb or add_reverb == "clean":
        return None
    if add_reverb not in RIR_FILES:
        raise ValueError("Invalid type of reverberation!")
    path = RIR_FILES[add_reverb]
    if rir_dir is not None:
        path = os.path.join(rir_dir, os.path.basename(path))
    from scipy.io.wavfile import read
    _, rir = read(path)  # FileNotFoundError when the recordings are not there, as in the reference
    return rir[:, 1] / np.power(2, 15)


def block_layout(lengths: Sequence[int], R: int, block: int, pairs_per_cta: int = 1):
    """Index tables of a packed batch for fdlp_reverb_forward (include/fdlp_b200.h): utterance u has
    ceil((T_u + R - 1) / block) blocks of the full convolution, two to a complex transform:
    D_u = ceil(blocks / 2) pairs and ceil(D_u / pairs_per_cta) convolution CTAs.
    Returns (utt_pair_off [n+1], pair_utt [sum D], utt_conv_off [n+1], conv_utt [sum CTAs]), all int32."""
    lengths = np.asarray(lengths, dtype=np.int64)
    nb = (lengths + R - 1 + block - 1) // block
    D = (nb + 1) // 2
    nc = (D + pairs_per_cta - 1) // pairs_per_cta
    ids = np.arange(len(lengths), dtype=np.int32)
    return (np.concatenate([[0], np.cumsum(D)]).astype(np.int32), np.repeat(ids, D).astype(np.int32),
            np.concatenate([[0], np.cumsum(nc)]).astype(np.int32), np.repeat(ids, nc).astype(np.int32))


class Reverberator:
    """One impulse response on one device.  `forward_packed` reverberates a packed device batch; calling the object
    with a list of waveforms returns a list of float64 device tensors."""

    def __init__(self, rir: np.ndarray, device=None, block: int = 0):
        if not torch.cuda.is_available():
            raise RuntimeError("Reverberator needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self._lib = L.load()
        self.rir = np.ascontiguousarray(np.asarray(rir, dtype=np.float64).reshape(-1))
        if self.rir.size < 1:
            raise ValueError("empty impulse response")
        self._plan = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_reverb_plan_create(self.rir.ctypes.data_as(C.c_void_p), int(self.rir.size),
                                                      int(block), C.byref(self._plan)))
        b, p, m = C.c_int32(), C.c_int32(), C.c_int32()
        L.check(self._lib.fdlp_reverb_plan_get_block(self._plan, C.byref(b), C.byref(p), C.byref(m)))
        self.block, self.partitions, self.pairs_per_cta = int(b.value), int(p.value), int(m.value)

    def close(self):
        if getattr(self, "_plan", None) and self._plan.value:
            with torch.cuda.device(self.device):
                self._lib.fdlp_reverb_plan_destroy(self._plan)
            self._plan = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_profiling(self, enable: bool = True):
        L.check(self._lib.fdlp_reverb_plan_set_profiling(self._plan, 1 if enable else 0))

    KERNELS = ("rv_fwd_kernel", "rv_conv_kernel", "rv_corr_kernel", "rv_align_kernel")

    def kernel_ms(self):
        out = (C.c_float * 4)()
        L.check(self._lib.fdlp_reverb_plan_get_kernel_ms(self._plan, out))
        return dict(zip(self.KERNELS, [float(v) for v in out]))

    def scratch_bytes(self, n_pair: int, n_utt: int) -> int:
        return int(self._lib.fdlp_reverb_plan_scratch_bytes(self._plan, int(n_pair), int(n_utt)))

    def block_layout(self, lengths: Sequence[int]):
        return block_layout(lengths, self.rir.size, self.block, self.pairs_per_cta)

    def forward_packed(self, wav_dev: torch.Tensor, lengths: Sequence[int], utt_wav_off_dev: Optional[torch.Tensor] = None,
                       noise_dev=None, out_dtype=torch.float64):
        """Reverberate a packed device batch (int16 or float32).  `noise_dev` is FdlpExtractor._noise_to_device's
        dictionary: the noise is mixed in first, as the reference does.  Returns (out, indM [n_utt] int32,
        out_len [n_utt] int64); `out` has the packed layout of the input."""
        if wav_dev.dtype not in (torch.int16, torch.float32):
            raise TypeError("waveforms must be int16 or float32")
        if out_dtype not in (torch.float32, torch.float64):
            raise TypeError("out_dtype must be float32 or float64")
        lengths = np.asarray(lengths, dtype=np.int64)
        n_utt = len(lengths)
        dev = self.device
        out = torch.empty(wav_dev.numel(), dtype=out_dtype, device=dev)  # rv_align_kernel writes every sample
        ind = torch.zeros(max(n_utt, 1), dtype=torch.int32, device=dev)
        out_len = torch.zeros(max(n_utt, 1), dtype=torch.int64, device=dev)
        if n_utt == 0:
            return out, ind[:0], out_len[:0]
        if utt_wav_off_dev is None:
            utt_wav_off_dev = torch.from_numpy(np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64)).to(dev)
        tabs = [torch.from_numpy(t).to(dev) for t in self.block_layout(lengths)]
        b = L.ReverbBatch()
        b.wav = wav_dev.data_ptr()
        b.wav_dtype = L.FDLP_WAV_INT16 if wav_dev.dtype == torch.int16 else L.FDLP_WAV_FLOAT32
        b.n_utt = n_utt
        b.utt_wav_off = utt_wav_off_dev.data_ptr()
        b.max_len = int(lengths.max())
        b.n_pair, b.n_conv = int(tabs[1].numel()), int(tabs[3].numel())
        b.utt_pair_off, b.pair_utt = tabs[0].data_ptr(), tabs[1].data_ptr()
        b.utt_conv_off, b.conv_utt = tabs[2].data_ptr(), tabs[3].data_ptr()
        if noise_dev is None:
            b.noise_mode = L.FDLP_NOISE_NONE
        else:
            b.noise_mode = noise_dev["mode"]
            b.noise = noise_dev["samples"].data_ptr()
            b.utt_noise_off = noise_dev["off"].data_ptr()
            b.utt_snr_db = noise_dev["snr"].data_ptr()
        b.out = out.data_ptr()
        b.out_dtype = L.FDLP_WAV_FLOAT64 if out_dtype == torch.float64 else L.FDLP_WAV_FLOAT32
        b.ind = ind.data_ptr()
        b.out_len = out_len.data_ptr()
        stream = torch.cuda.current_stream(dev).cuda_stream
        with torch.cuda.device(dev):
            L.check(self._lib.fdlp_reverb_forward(self._plan, C.byref(b), C.c_void_p(stream)))
        self._keep = (tabs, utt_wav_off_dev)  # alive until the next call: the launch is asynchronous
        return out, ind, out_len

    def __call__(self, wavs: Sequence[Union[np.ndarray, torch.Tensor]], out_dtype=torch.float64):
        """List of mono int16 / float32 waveforms -> list of reverberated device tensors (length T_u, or T_u - 1 in
        the reference's corner case indM == len(rir))."""
        if len(wavs) == 0:
            return []
        ts = [w if isinstance(w, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(w)) for w in wavs]
        lengths = [int(t.numel()) for t in ts]
        wav_dev = torch.cat([t.reshape(-1) for t in ts]).to(self.device)
        out, _, out_len = self.forward_packed(wav_dev, lengths, out_dtype=out_dtype)
        off = np.concatenate([[0], np.cumsum(lengths)])
        n = out_len.cpu().numpy()
        return [out[int(off[u]):int(off[u]) + int(n[u])] for u in range(len(lengths))]


def reverberate_packed(reverb, wav_dev: torch.Tensor, lengths, device, out_dtype=torch.float64, noise_dev=None,
                       utt_wav_off_dev=None):
    """Shared by the extractors' --add_reverb paths: `reverb` is a Reverberator or an impulse response.  Returns the
    reverberated packed batch and its per-utterance lengths; utterances that come back one sample short (indM ==
    len(rir), features.py:113-114) are re-packed."""
    rv = reverb if isinstance(reverb, Reverberator) else Reverberator(reverb, device=device)
    lengths = np.asarray(lengths, dtype=np.int64)
    out, _, out_len = rv.forward_packed(wav_dev, lengths, utt_wav_off_dev, noise_dev, out_dtype)
    n = out_len.cpu().numpy()
    if np.array_equal(n, lengths):
        return out, lengths
    off = np.concatenate([[0], np.cumsum(lengths)])
    return torch.cat([out[int(off[u]):int(off[u]) + int(n[u])] for u in range(len(lengths))]), n
