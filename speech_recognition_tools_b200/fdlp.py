"""Host side of the B200 FDLP-spectrogram extractor.

Mirrors the reference's `getFeats` (src/featgen/computeFDLPSpectrogram.py:29-237): same options, same
frame/segment geometry, same output layout (frames x sub-bands, natural log).  The host does only what
the reference does once per job or per utterance outside the hot loop -- option parsing, the integer
segment/frame bookkeeping and the overlap-add pointer trace with its `randrange(2)` jitter
(:220-225) -- and hands packed device buffers to libfdlp_b200.so through the C ABI
(include/fdlp_b200.h).  All arithmetic on samples and features happens in the CUDA kernels; there is no
CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import math
import random as _random
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Union

import numpy as np
import torch

from . import _lib as L

LOG_FLOOR = math.log(1e-14)


# --------------------------------------------------------------------------- configuration
@dataclass
class FdlpConfig:
    """The reference's argparse surface (computeFDLPSpectrogram.py:241-262) minus I/O options."""
    nfilters: int = 20
    coeff_num: int = 50
    coeff_range: str = "1,20"
    order: int = 50
    fduration: float = 0.5
    frate: int = 100
    overlap_fraction: float = 0.25
    fbank_type: str = "mel,1"
    odd_mod_zero: bool = False
    gamma_weight: str = "None"
    lifter: Optional[Sequence[float]] = None  # contents of --lifter_config (first line, comma separated)
    srate: int = 16000
    dct_fp32: bool = False
    # the autocorrelation skips filter taps below tail_eps * max|F_j| (0.0: only exactly-zero taps); see DESIGN.md
    tail_eps: float = 1e-12
    # envelopes cross from the envelope kernel to the overlap-add kernel as float32 (fdlp_params.env_f64 = 0); a value
    # outside the float32 range (extreme lifter / gamma scalings) raises FDLP_ST_ENV_RANGE in the status word
    env_f64: bool = False

    @classmethod
    def from_args(cls, args) -> "FdlpConfig":
        lifter = None
        if getattr(args, "lifter_config", None):
            with open(args.lifter_config, "r") as fid:  # computeFDLPSpectrogram.py:43-46
                lifter = [float(x) for x in fid.readline().strip().split(",")]
        gw = args.gamma_weight
        if not isinstance(gw, str):
            gw = ",".join(gw)
        return cls(nfilters=args.nfilters, coeff_num=args.coeff_num, coeff_range=args.coeff_range, order=args.order,
                   fduration=args.fduration, frate=args.frate, overlap_fraction=args.overlap_fraction,
                   fbank_type=args.fbank_type, odd_mod_zero=bool(args.odd_mod_zero), gamma_weight=gw, lifter=lifter)


def parse_fbank_type(fbank_type: str):
    """computeFDLPSpectrogram.py:49-63: 'mel,warp' or 'cochlear,om_w,alp,fixed,bet,warp'."""
    tok = fbank_type.strip().split(",")
    if tok[0] == "mel":
        if len(tok) < 2:
            raise ValueError("Mel filter bank not configured properly....")
        return L.FDLP_FBANK_MEL, [float(tok[1]), 0, 0, 0, 0]
    if tok[0] == "cochlear":
        if len(tok) < 6:
            raise ValueError("Cochlear filter bank not configured properly....")
        return L.FDLP_FBANK_COCHLEAR, [float(tok[1]), float(tok[2]), float(int(tok[3])), float(tok[4]), float(tok[5])]
    raise ValueError("Invalid type of filter bank, use mel or cochlear with proper configuration")


def build_weights(cfg: FdlpConfig) -> np.ndarray:
    """Mask (computeFDLPSpectrogram.py:94-103) * lifter (:195-196) * gamma weights (:107-118,197-198), with
    `ms[1::2] = 0` (:199-200) folded in: one float64 vector of length coeff_num."""
    lo, hi = (int(v) for v in cfg.coeff_range.split(","))
    w = np.array([1.0 if (lo <= i <= hi) else 0.0 for i in range(cfg.coeff_num)])
    if cfg.lifter is not None:
        lifter = np.asarray(cfg.lifter, dtype=np.float64)
        if lifter.shape[0] != cfg.coeff_num:
            raise ValueError("operands could not be broadcast together with shapes (%d,) (%d,)"
                             % (cfg.coeff_num, lifter.shape[0]))
        w = w * lifter
    g = cfg.gamma_weight.strip().split(",")
    if g[0] != "None":
        import scipy.stats
        if cfg.order != cfg.coeff_num:
            raise ValueError("operands could not be broadcast together with shapes (%d,) (%d,)"
                             % (cfg.coeff_num, cfg.order))
        x = np.linspace(0, cfg.order - 1, cfg.order)
        scale, shape, pk_required = float(g[0]), float(g[1]), float(g[2])
        pk_required = pk_required * (2 * cfg.fduration)
        pk = (shape - 1) * scale
        loc = -pk + pk_required
        w = w * (scipy.stats.gamma.pdf(x, a=shape, loc=loc, scale=scale) * 3 * scale)
    if cfg.odd_mod_zero:
        w = w.copy()
        w[1::2] = 0.0
    return np.ascontiguousarray(w, dtype=np.float64)


def segment_count(T: int, N: int, S: int) -> int:
    """Number of frames getFrames yields (features.py:134-154): idx = sp_b + k*S while idx + sp_f < T + 2*extend."""
    if N % 2 == 0:
        lim = T - 1   # k*S < T - 1
    else:
        lim = T
    return max(0, -(-lim // S))


def frame_count(T: int, frate: int, srate: int) -> int:
    """computeFDLPSpectrogram.py:182: int(np.ceil(tframes * frate / srate))."""
    return int(np.ceil(T * frate / srate))


def ola_pointers(nseg: int, Fr: int, kk: int, kkb2: int, kks: int, rng) -> List[int]:
    """Overlap-add pointer trace, computeFDLPSpectrogram.py:207-225, raising where numpy would.

    One `rng.randrange(2)` draw per segment i >= 1 (including the last), exactly as line 225."""
    ptrs = []
    ptr = 0
    for i in range(nseg):
        ptrs.append(ptr)
        if i == 0:
            if Fr >= kkb2 and (kk - kkb2) != kkb2:
                raise ValueError("operands could not be broadcast together with shapes (%d,) (%d,) "
                                 "(odd fduration*frate is not supported by the reference either)" % (kkb2, kk - kkb2))
        elif i >= nseg - 2:
            n = Fr - ptr
            if n < 0 and (kk + n) > 1:
                raise ValueError("operands could not be broadcast together with shapes (0,) (%d,): "
                                 "overlap-add pointer ran past the utterance (reference raises too)" % (kk + n))
        else:
            if ptr + kk > Fr:
                raise ValueError("operands could not be broadcast together with shapes (%d,) (%d,): "
                                 "overlap-add pointer ran past the utterance (reference raises too)"
                                 % (max(0, Fr - ptr), kk))
        if i == 0:
            ptr = int(ptr + kks - kkb2)
        else:
            ptr = int(ptr + kks + rng.randrange(2))
    return ptrs


def _randrange2_bulk(rng, n: int) -> np.ndarray:
    """`[rng.randrange(2) for _ in range(n)]` for a Mersenne-Twister `random.Random`, from two bulk `getrandbits`
    calls instead of n Python calls, leaving the generator in exactly the state the loop leaves it in.
    CPython: randrange(2) -> _randbelow(2) -> k = (2).bit_length() = 2; r = getrandbits(2) until r < 2, and
    getrandbits(k <= 32) is the top k bits of one 32-bit output; getrandbits(32 m) is m outputs, least significant
    word first.  Checked against the loop once per process on a copy of the generator (other interpreters / subclasses:
    the loop itself)."""
    if n <= 0:
        return np.zeros(0, dtype=np.int64)
    inst = getattr(rng, "mt_instance", rng)  # a wrapper around the module-level generator exposes it
    if type(inst) is not _random.Random or not _bulk_ok():
        rr = rng.randrange
        return np.fromiter((rr(2) for _ in range(n)), dtype=np.int64, count=n)
    state = inst.getstate()
    words = 2 * n + 64
    while True:
        raw = inst.getrandbits(32 * words).to_bytes(4 * words, "little")
        top = np.frombuffer(raw, dtype="<u4") >> 30
        acc = np.flatnonzero(top < 2)
        if acc.size >= n:
            break
        inst.setstate(state)
        words *= 2
    used = int(acc[n - 1]) + 1       # outputs the loop would have consumed
    inst.setstate(state)
    inst.getrandbits(32 * used)      # advance by exactly that many
    return top[acc[:n]].astype(np.int64)


_BULK_OK = None


def _bulk_ok() -> bool:
    global _BULK_OK
    if _BULK_OK is None:
        _BULK_OK = True  # (set first: the check below goes through _randrange2_bulk itself)
        a, b = _random.Random(987654321), _random.Random(987654321)
        fast = _randrange2_bulk(a, 257)
        slow = [b.randrange(2) for _ in range(257)]
        _BULK_OK = list(fast) == slow and a.getstate() == b.getstate()
    return _BULK_OK


def ola_pointers_batch(nsegs: np.ndarray, frs: np.ndarray, kk: int, kkb2: int, kks: int, rngs) -> np.ndarray:
    """`ola_pointers` for a whole batch at once (utterance-major): the same draws in the same order -- one
    `randrange(2)` per segment i >= 1 of every utterance, the last included (computeFDLPSpectrogram.py:225) -- the
    same pointer trace and the same ValueError where the reference's slice assignment would raise; the arithmetic
    around the draws is numpy instead of a Python loop per segment (1000 utterances: 0.1 s -> 5 ms)."""
    nsegs = np.asarray(nsegs, dtype=np.int64)
    frs = np.asarray(frs, dtype=np.int64)
    total = int(nsegs.sum())
    if total == 0:
        return np.zeros(0, dtype=np.int32)
    has = nsegs > 0
    if (kk - kkb2) != kkb2 and np.any(has & (frs >= kkb2)):
        raise ValueError("operands could not be broadcast together with shapes (%d,) (%d,) "
                         "(odd fduration*frate is not supported by the reference either)" % (kkb2, kk - kkb2))
    seg_off = np.concatenate([[0], np.cumsum(nsegs)])
    first = seg_off[:-1][has]
    is_first = np.zeros(total, dtype=bool)
    is_first[first] = True
    ndraw = np.maximum(nsegs - 1, 0)
    if all(r is rngs[0] for r in rngs):  # one generator consumed sequentially over the utterances
        bits = _randrange2_bulk(rngs[0], int(ndraw.sum()))
    else:
        bits = np.concatenate([_randrange2_bulk(r, int(n)) for r, n in zip(rngs, ndraw)] + [np.zeros(0, dtype=np.int64)])
    inc = np.full(total, kks, dtype=np.int64)   # added after segment i: kks + randrange(2) for i >= 1 ...
    inc[~is_first] += bits
    inc[is_first] = kks - kkb2                   # ... and kks - kkb2 after segment 0
    excl = np.cumsum(inc) - inc
    ptr = excl - np.repeat(excl[first], nsegs[has])
    i = np.arange(total) - np.repeat(seg_off[:-1][has], nsegs[has])
    nseg_r, fr_r = np.repeat(nsegs[has], nsegs[has]), np.repeat(frs[has], nsegs[has])
    late = (i > 0) & (i >= nseg_r - 2)
    n_left = fr_r - ptr
    bad = (late & (n_left < 0) & (kk + n_left > 1)) | ((i > 0) & ~late & (ptr + kk > fr_r))
    if bad.any():
        b = int(np.argmax(bad))
        if late[b]:
            raise ValueError("operands could not be broadcast together with shapes (0,) (%d,): "
                             "overlap-add pointer ran past the utterance (reference raises too)" % (kk + n_left[b]))
        raise ValueError("operands could not be broadcast together with shapes (%d,) (%d,): "
                         "overlap-add pointer ran past the utterance (reference raises too)"
                         % (max(0, int(n_left[b])), kk))
    return ptr.astype(np.int32)


def batch_layout(lengths, geom, frate: int, srate: int, jitter=None) -> "BatchLayout":
    """Index tables of one packed batch from the utterance lengths: segment / frame counts (features.py:134-154,
    computeFDLPSpectrogram.py:182) and the overlap-add pointer trace with its jitter (:207-225).  `geom` needs N, S, kk,
    kkb2, kks.  Pure host arithmetic (no device access): the CLI pipeline runs it while the GPU works on the previous batch."""
    lengths = np.asarray(lengths, dtype=np.int64)
    n_utt = len(lengths)
    if n_utt and lengths.min() < 2:
        raise ValueError("utterances need at least 2 samples")
    rngs = _as_rngs(jitter, n_utt)
    lim = lengths - 1 if geom["N"] % 2 == 0 else lengths
    nsegs = np.maximum(0, -(-lim // geom["S"])).astype(np.int64)
    frs = np.ceil(lengths * frate / srate).astype(np.int64)
    seg_ptr = ola_pointers_batch(nsegs, frs, geom["kk"], geom["kkb2"], geom["kks"], rngs)
    return BatchLayout(
        lengths=lengths,
        utt_wav_off=np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64),
        utt_seg_off=np.concatenate([[0], np.cumsum(nsegs)]).astype(np.int32),
        utt_frame_off=np.concatenate([[0], np.cumsum(frs)]).astype(np.int64),
        seg_utt=np.repeat(np.arange(n_utt, dtype=np.int32), nsegs).astype(np.int32),
        seg_ptr=seg_ptr,
        n_utt=n_utt, n_seg=int(nsegs.sum()), n_frames=int(frs.sum()),
        max_utt_frames=int(frs.max()) if n_utt else 0)


# --------------------------------------------------------------------------- batch bookkeeping
@dataclass
class BatchLayout:
    """Host-side index tables of one packed batch (what fdlp_batch points at)."""
    lengths: np.ndarray
    utt_wav_off: np.ndarray
    utt_seg_off: np.ndarray
    utt_frame_off: np.ndarray
    seg_utt: np.ndarray
    seg_ptr: np.ndarray
    n_utt: int = 0
    n_seg: int = 0
    n_frames: int = 0
    max_utt_frames: int = 0
    dev: dict = field(default_factory=dict)

    def to(self, device):
        if self.dev.get("device") == device:
            return self
        # one pinned staging buffer, one H2D copy, then typed views
        parts = [self.utt_wav_off.astype(np.int64), self.utt_frame_off.astype(np.int64),
                 self.utt_seg_off.astype(np.int32), self.seg_utt.astype(np.int32), self.seg_ptr.astype(np.int32)]
        sizes = [p.nbytes for p in parts]
        padded = [(s + 15) // 16 * 16 for s in sizes]
        host = torch.empty(sum(padded), dtype=torch.uint8, pin_memory=torch.cuda.is_available())
        hv = host.numpy()
        off = 0
        offs = []
        for p, s, ps in zip(parts, sizes, padded):
            hv[off:off + s] = p.view(np.uint8).reshape(-1)
            offs.append(off)
            off += ps
        d = host.to(device, non_blocking=True)
        names = ["utt_wav_off", "utt_frame_off", "utt_seg_off", "seg_utt", "seg_ptr"]
        dtypes = [torch.int64, torch.int64, torch.int32, torch.int32, torch.int32]
        self.dev = {"device": device, "_buf": d, "_host": host}
        for name, o, s, dt in zip(names, offs, sizes, dtypes):
            self.dev[name] = d[o:o + s].view(dt)
        return self


def split_layout(lay: "BatchLayout", n_chunks: int, taper: bool = False):
    """Cut a packed batch into up to `n_chunks` runs of whole utterances with about equal segment counts
    (`taper`: segment counts in the ratio 1 : 2 : 3 ... 3 : 2 : 1, so that the first upload and the last
    download of a copy / compute pipeline, which nothing overlaps, are short).
    Returns [(u0, u1, sub_layout)]; the sub-layout's offsets are relative to the chunk's first sample / segment /
    frame, so each chunk is a self-contained fdlp_batch over a slice of the packed buffers."""
    n_chunks = max(1, min(int(n_chunks), lay.n_utt))
    wts = np.array([min(c + 1, n_chunks - c, 3) if taper else 1 for c in range(n_chunks)], dtype=np.float64)
    cum = np.cumsum(wts) / wts.sum()
    bounds = [0]
    for c in range(1, n_chunks):
        target = lay.n_seg * cum[c - 1]
        u = int(np.searchsorted(lay.utt_seg_off, target, side="left"))
        u = min(max(u, bounds[-1] + 1), lay.n_utt - (n_chunks - c))
        bounds.append(u)
    bounds.append(lay.n_utt)
    out = []
    for u0, u1 in zip(bounds[:-1], bounds[1:]):
        s0, s1 = int(lay.utt_seg_off[u0]), int(lay.utt_seg_off[u1])
        f0, f1 = int(lay.utt_frame_off[u0]), int(lay.utt_frame_off[u1])
        frs = np.diff(lay.utt_frame_off[u0:u1 + 1])
        sub = BatchLayout(
            lengths=lay.lengths[u0:u1],
            utt_wav_off=(lay.utt_wav_off[u0:u1 + 1] - lay.utt_wav_off[u0]).astype(np.int64),
            utt_seg_off=(lay.utt_seg_off[u0:u1 + 1] - s0).astype(np.int32),
            utt_frame_off=(lay.utt_frame_off[u0:u1 + 1] - f0).astype(np.int64),
            seg_utt=(lay.seg_utt[s0:s1] - u0).astype(np.int32),
            seg_ptr=lay.seg_ptr[s0:s1].astype(np.int32),
            n_utt=u1 - u0, n_seg=s1 - s0, n_frames=f1 - f0,
            max_utt_frames=int(frs.max()) if u1 > u0 else 0)
        out.append((u0, u1, sub))
    return out


def _as_rngs(jitter, n_utt):
    """`jitter`: a random.Random consumed sequentially over the utterances (what the reference does within
    one process), an int seed for such a generator, or a sequence of per-utterance seeds."""
    if jitter is None:
        jitter = 0
    if isinstance(jitter, _random.Random) or hasattr(jitter, "randrange"):
        return [jitter] * n_utt
    if isinstance(jitter, (int, np.integer)):
        r = _random.Random(int(jitter))
        return [r] * n_utt
    seeds = list(jitter)
    if len(seeds) != n_utt:
        raise ValueError("need one jitter seed per utterance")
    return [s if isinstance(s, _random.Random) else _random.Random(int(s)) for s in seeds]


# --------------------------------------------------------------------------- noise
@dataclass
class NoiseSpec:
    """--add_noise 'type,snr' (features.py:24-44): `samples` is the int16 noise recording, `snr_db` a scalar or
    per-utterance array, `rand` the per-utterance uniform draws standing for `np.random.rand()` (or a
    np.random.RandomState consumed sequentially, or an int seed).  A multi-condition batch (REVERB-shaped streams,
    one noise type per stream) passes a LIST of recordings as `samples` and `which[u]` = the recording of utterance u;
    the reference takes one `type,snr` per job, so such a batch equals one reference job per stream."""
    samples: Union[np.ndarray, torch.Tensor, Sequence[Union[np.ndarray, torch.Tensor]]]
    snr_db: Union[float, Sequence[float]]
    rand: Union[None, int, np.random.RandomState, Sequence[float]] = None
    mode: str = "int16_wrap"  # 'int16_wrap' = the reference's numpy dtype semantics; 'float' = true energies
    which: Optional[Sequence[int]] = None


def noise_to_device(noise: Optional["NoiseSpec"], lengths, device):
    """Device tables of --add_noise for one batch: the recording(s), per-utterance excerpt offsets
    floor(rand * (len(noise) - T)) (features.py:25) and SNRs."""
    if noise is None:
        return None
    lengths = np.asarray(lengths, dtype=np.int64)
    recs = noise.samples if isinstance(noise.samples, (list, tuple)) else [noise.samples]
    recs = [r if isinstance(r, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(r)) for r in recs]
    if any(r.dtype != torch.int16 or r.dim() != 1 for r in recs):
        raise TypeError("noise recording must be int16 (scipy.io.wavfile.read of noises/<type>.wav)")
    n_utt = len(lengths)
    which = np.zeros(n_utt, dtype=np.int64) if noise.which is None else np.asarray(noise.which, dtype=np.int64)
    if which.shape != (n_utt,) or (n_utt and (which.min() < 0 or which.max() >= len(recs))):
        raise ValueError("NoiseSpec.which needs one recording index per utterance")
    rec_len = np.array([int(r.numel()) for r in recs], dtype=np.int64)
    rec_base = np.concatenate([[0], np.cumsum(rec_len)])[:-1]
    ns = recs[0] if len(recs) == 1 else torch.cat(recs)
    rand = noise.rand
    if rand is None or isinstance(rand, (int, np.integer)):
        rand = np.random.RandomState(0 if rand is None else int(rand))
    if isinstance(rand, np.random.RandomState):
        draws = np.array([rand.rand() for _ in range(n_utt)])
    else:
        draws = np.asarray(list(rand), dtype=np.float64)
    Tn = rec_len[which]
    if n_utt and np.any(Tn <= lengths):
        raise ValueError("noise recording must be longer than every utterance (features.py:25-26)")
    off = rec_base[which] + np.floor(draws * (Tn - lengths)).astype(np.int64)  # features.py:25
    snr = np.broadcast_to(np.asarray(noise.snr_db, dtype=np.float64), (n_utt,)).copy()
    mode = {"int16_wrap": L.FDLP_NOISE_INT16_WRAP, "float": L.FDLP_NOISE_FLOAT}[noise.mode]
    return dict(mode=mode, samples=ns.to(device, non_blocking=True),
                off=torch.from_numpy(off).to(device), snr=torch.from_numpy(snr).to(device))


# --------------------------------------------------------------------------- extractor
def host_threads(cap: int) -> int:
    """Worker threads of the host-side copy stages: the cores this process may use, shared between the ranks torchrun
    started on this node (LOCAL_WORLD_SIZE), at least 2, at most `cap`.  On the 8-GPU boxes of this pool (32 vCPUs) that
    is 4 per rank: more threads than cores only delays the thread that drives the GPU."""
    import os
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 4
    ranks = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1") or 1))
    return max(2, min(cap, cores // ranks))


class FdlpExtractor:
    """One plan (filterbank, twiddles, scratch) on one device; call it with a list of waveforms."""

    def __init__(self, cfg: Optional[FdlpConfig] = None, device=None, **kwargs):
        self.cfg = cfg if cfg is not None else FdlpConfig(**kwargs)
        if not torch.cuda.is_available():
            raise RuntimeError("FdlpExtractor needs a CUDA device (sm_100a); there is no CPU fallback")
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self._lib = L.load()
        kind, fargs = parse_fbank_type(self.cfg.fbank_type)
        self._weights = build_weights(self.cfg)
        prm = L.FdlpParams()
        prm.srate, prm.nfilters, prm.order = self.cfg.srate, self.cfg.nfilters, self.cfg.order
        prm.coeff_num, prm.frate = self.cfg.coeff_num, self.cfg.frate
        prm.fduration, prm.overlap_fraction = float(self.cfg.fduration), float(self.cfg.overlap_fraction)
        prm.fbank_kind = kind
        prm.fbank_args = (C.c_double * 5)(*fargs)
        prm.weights = self._weights.ctypes.data_as(C.POINTER(C.c_double))
        prm.custom_fbank = None
        prm.dct_fp32 = 1 if self.cfg.dct_fp32 else 0
        prm.tail_eps = float(self.cfg.tail_eps)
        prm.env_f64 = 1 if self.cfg.env_f64 else 0
        self._plan = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_plan_create(C.byref(prm), C.byref(self._plan)))
        g = L.FdlpGeometry()
        L.check(self._lib.fdlp_plan_get_geometry(self._plan, C.byref(g)))
        self.geom = {n: getattr(g, n) for n, _ in L.FdlpGeometry._fields_}

    def close(self):
        if getattr(self, "_pack_pool", None) is not None:
            self._pack_pool.shutdown(wait=False)
            self._pack_pool = None
        if getattr(self, "_plan", None) and self._plan.value:
            with torch.cuda.device(self.device):
                self._lib.fdlp_plan_destroy(self._plan)
            self._plan = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- introspection
    @property
    def nfilters(self):
        return self.cfg.nfilters

    def filterbank(self) -> np.ndarray:
        out = np.empty((self.cfg.nfilters, self.geom["N"]), dtype=np.float64)
        L.check(self._lib.fdlp_plan_get_fbank(self._plan, out.ctypes.data_as(C.c_void_p)))
        return out

    def scratch_bytes(self, n_seg: int) -> int:
        return int(self._lib.fdlp_plan_scratch_bytes(self._plan, int(n_seg)))

    def set_profiling(self, enable: bool = True):
        L.check(self._lib.fdlp_plan_set_profiling(self._plan, 1 if enable else 0))

    def set_overlap(self, enable: bool = True):
        L.check(self._lib.fdlp_plan_set_overlap(self._plan, 1 if enable else 0))

    def set_autocorr_mode(self, mode: str = "auto"):
        """'auto': family decomposition when the filterbank allows it (cochlear); 'direct': per-band kernel only."""
        L.check(self._lib.fdlp_plan_set_autocorr_mode(self._plan, {"auto": 0, "direct": 1}[mode]))

    def autocorr_ms(self):
        """CUDA-event durations of the three autocorrelation kernels of the most recent profiled forward."""
        out = (C.c_float * 3)()
        L.check(self._lib.fdlp_plan_get_autocorr_ms(self._plan, out))
        return dict(family_blocks=out[0], combine=out[1], direct=out[2])

    KERNELS = ("dct_segments_kernel", "family_block_autocorr_kernel", "family_scan_kernel", "autocorr_direct",
               "lpc_cepstrum", "envelope_dmma_kernel", "ola_log_kernel")

    def kernel_ms(self):
        """Per-kernel CUDA-event durations of the most recent profiled forward, keyed by KERNELS."""
        out = (C.c_float * 7)()
        L.check(self._lib.fdlp_plan_get_kernel_ms(self._plan, out))
        return dict(zip(self.KERNELS, [float(v) for v in out]))

    def stage_ms(self):
        """Per-stage CUDA-event durations of the most recent forward (needs set_profiling(True))"""
        out = (C.c_float * 4)()
        L.check(self._lib.fdlp_plan_get_stage_ms(self._plan, out))
        return dict(dct=out[0], autocorr=out[1], lpc_env=out[2], ola=out[3])

    # ---- bookkeeping
    def layout(self, lengths: Sequence[int], jitter=None) -> BatchLayout:
        return batch_layout(lengths, self.geom, self.cfg.frate, self.cfg.srate, jitter)

    @staticmethod
    def pack(wavs: Sequence[Union[np.ndarray, torch.Tensor]]):
        """Concatenate 1-D waveforms (all int16 or all float32) into one pinned host tensor."""
        ts = [w if isinstance(w, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(w)) for w in wavs]
        for t in ts:
            if t.dim() != 1:
                raise ValueError("waveforms must be mono 1-D arrays (the reference pads every axis otherwise)")
        dt = ts[0].dtype
        if dt not in (torch.int16, torch.float32) or any(t.dtype != dt for t in ts):
            raise TypeError("waveforms must all be int16 or all float32 (what scipy.io.wavfile.read returns)")
        total = sum(int(t.numel()) for t in ts)
        host = torch.empty(total, dtype=dt, pin_memory=torch.cuda.is_available())
        torch.cat(ts, out=host)
        return host, [int(t.numel()) for t in ts]

    def _batch_struct(self, wav_dev, lay, feats, cmvn, status, noise_dev):
        b = L.FdlpBatch()
        d = lay.dev
        b.wav = wav_dev.data_ptr()
        b.wav_dtype = {torch.int16: L.FDLP_WAV_INT16, torch.float32: L.FDLP_WAV_FLOAT32,
                       torch.float64: L.FDLP_WAV_FLOAT64}[wav_dev.dtype]
        b.n_utt, b.n_seg, b.n_frames, b.max_utt_frames = lay.n_utt, lay.n_seg, lay.n_frames, lay.max_utt_frames
        b.utt_wav_off = d["utt_wav_off"].data_ptr()
        b.utt_seg_off = d["utt_seg_off"].data_ptr()
        b.utt_frame_off = d["utt_frame_off"].data_ptr()
        b.seg_utt = d["seg_utt"].data_ptr()
        b.seg_ptr = d["seg_ptr"].data_ptr()
        if noise_dev is None:
            b.noise_mode = L.FDLP_NOISE_NONE
        else:
            b.noise_mode = noise_dev["mode"]
            b.noise = noise_dev["samples"].data_ptr()
            b.utt_noise_off = noise_dev["off"].data_ptr()
            b.utt_snr_db = noise_dev["snr"].data_ptr()
        b.feats = feats.data_ptr() if feats is not None else None
        b.cmvn_acc = cmvn.data_ptr() if cmvn is not None else None
        b.status = status.data_ptr() if status is not None else None
        return b

    def _noise_to_device(self, noise: Optional[NoiseSpec], lay):
        return noise_to_device(noise, lay.lengths, self.device)

    # ---- the hot path
    def forward_packed(self, wav_dev: torch.Tensor, lay: BatchLayout, feats: Optional[torch.Tensor] = None,
                       cmvn: Optional[torch.Tensor] = None, status: Optional[torch.Tensor] = None,
                       noise_dev=None):
        """Run the whole path on an already-resident packed batch.  Returns (feats [n_frames, J] fp32, status)."""
        lay.to(self.device)
        if feats is None:
            feats = torch.empty((lay.n_frames, self.cfg.nfilters), dtype=torch.float32, device=self.device)
        if status is None:
            status = torch.zeros(max(lay.n_utt, 1), dtype=torch.int32, device=self.device)
        b = self._batch_struct(wav_dev, lay, feats, cmvn, status, noise_dev)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_forward(self._plan, C.byref(b), C.c_void_p(stream)))
        return feats, status

    def diff_filter(self, wav_dev: torch.Tensor, lay: BatchLayout) -> torch.Tensor:
        """--add_noise diff (computeFDLPSpectrogram.py:162-164): the 13-tap FIR `convolve(signal, a, mode='same')`
        applied per utterance to a packed device batch.  Returns a float32 packed batch (exact for int16 input)."""
        lay.to(self.device)
        out = torch.empty(wav_dev.numel(), dtype=torch.float32, device=self.device)
        dt = L.FDLP_WAV_INT16 if wav_dev.dtype == torch.int16 else L.FDLP_WAV_FLOAT32
        if wav_dev.dtype not in (torch.int16, torch.float32):
            raise TypeError("waveforms must be int16 or float32")
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_diff_filter(C.c_void_p(wav_dev.data_ptr()), dt,
                                                C.c_void_p(lay.dev["utt_wav_off"].data_ptr()), lay.n_utt,
                                                int(lay.lengths.max()) if lay.n_utt else 0,
                                                C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
        return out

    def extract_host(self, host_wav: torch.Tensor, lay: BatchLayout, feats_host: Optional[torch.Tensor] = None,
                     n_chunks: int = 6, cmvn: Optional[torch.Tensor] = None, status: Optional[torch.Tensor] = None,
                     wav_dev: Optional[torch.Tensor] = None, feats: Optional[torch.Tensor] = None,
                     taper: bool = True, noise_dev=None, fill=None):
        """End-to-end call for a packed batch that lives in (pinned) HOST memory: samples go host -> device,
        features device -> host, and both copies overlap the kernels.  The batch is cut into `n_chunks` runs of
        whole utterances; chunk c+1 is uploaded on a copy stream while chunk c computes on the current stream
        and the features of chunk c-1 are downloaded on a third stream.  Results are identical to one
        forward_packed over the whole batch (utterances are independent).  `fill(u0, u1, w0, w1)`, if given, is called
        right before chunk [u0, u1) is uploaded and must have put its samples into host_wav[w0:w1] when it returns
        (extract_list packs a list of arrays this way, chunk by chunk, while the device is busy).
        Returns (feats_dev, feats_host, status); the current stream waits for the last download."""
        dev = self.device
        J = self.cfg.nfilters
        if wav_dev is None:
            wav_dev = torch.empty(host_wav.numel(), dtype=host_wav.dtype, device=dev)
        if feats is None:
            feats = torch.empty((lay.n_frames, J), dtype=torch.float32, device=dev)
        if feats_host is None:
            feats_host = torch.empty((lay.n_frames, J), dtype=torch.float32, pin_memory=True)
        if status is None:
            status = torch.zeros(max(lay.n_utt, 1), dtype=torch.int32, device=dev)
        if not hasattr(self, "_s_in"):
            self._s_in, self._s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        cur = torch.cuda.current_stream(dev)
        chunks = split_layout(lay, n_chunks, taper)
        ev0 = torch.cuda.Event()
        ev0.record(cur)
        self._s_in.wait_event(ev0)
        self._s_out.wait_event(ev0)
        for u0, u1, sub in chunks:
            w0, w1 = int(lay.utt_wav_off[u0]), int(lay.utt_wav_off[u1])
            f0, f1 = int(lay.utt_frame_off[u0]), int(lay.utt_frame_off[u1])
            if fill is not None:
                fill(u0, u1, w0, w1)  # the host packs this chunk while the device works on the previous ones
            with torch.cuda.stream(self._s_in):
                wav_dev[w0:w1].copy_(host_wav[w0:w1], non_blocking=True)
                # the framing kernel uses 16-byte loads when the batch base pointer is 16-byte aligned:
                # hand it the slice from the aligned sample below w0 and shift the offsets instead
                sub.utt_wav_off = sub.utt_wav_off + (w0 % 8)
                sub.to(dev)
                e = torch.cuda.Event()
                e.record(self._s_in)
            cur.wait_event(e)
            nd = None
            if noise_dev is not None:  # per-utterance noise tables of this chunk (the recording itself is shared)
                nd = dict(noise_dev, off=noise_dev["off"][u0:u1], snr=noise_dev["snr"][u0:u1])
            self.forward_packed(wav_dev[w0 - (w0 % 8):w1], sub, feats=feats[f0:f1], cmvn=cmvn, status=status[u0:u1],
                                noise_dev=nd)
            done = torch.cuda.Event()
            done.record(cur)
            self._s_out.wait_event(done)
            with torch.cuda.stream(self._s_out):
                feats_host[f0:f1].copy_(feats[f0:f1], non_blocking=True)
        self._last_upload = e if chunks else None
        fin = torch.cuda.Event()
        fin.record(self._s_out)
        cur.wait_event(fin)
        return feats, feats_host, status

    def extract_list(self, wavs: Sequence[np.ndarray], jitter=None, feats_host: Optional[torch.Tensor] = None,
                     cmvn: Optional[torch.Tensor] = None, status: Optional[torch.Tensor] = None, n_chunks: int = 6,
                     wav_dev: Optional[torch.Tensor] = None, feats: Optional[torch.Tensor] = None,
                     threads: Optional[int] = None):
        """The whole host side of one batch: a list of (unpinned) int16 / float32 numpy waveforms in, pinned-host
        features out.  Layout (segment / frame tables, jitter draws) first -- it needs the lengths only -- then the
        utterances are packed into a grow-only pinned staging buffer chunk by chunk, by `threads` workers (numpy
        copies release the GIL), each chunk right before its upload, so packing overlaps the device work of the
        chunks before it.  Returns (lay, feats_dev, feats_host, status)."""
        from concurrent.futures import ThreadPoolExecutor
        if threads is None:
            threads = host_threads(8)
        wavs = [np.ascontiguousarray(w) for w in wavs]
        dt = wavs[0].dtype
        if dt not in (np.int16, np.float32) or any(w.dtype != dt or w.ndim != 1 for w in wavs):
            raise TypeError("waveforms must be mono and all int16 or all float32 (what scipy.io.wavfile.read returns)")
        lay = batch_layout([len(w) for w in wavs], self.geom, self.cfg.frate, self.cfg.srate, jitter)
        total = int(lay.utt_wav_off[-1])
        if getattr(self, "_last_upload", None) is not None:
            self._last_upload.synchronize()  # the previous call's uploads have left the staging buffer
        tdt = torch.int16 if dt == np.int16 else torch.float32
        st = getattr(self, "_staging", None)
        if st is None or st.dtype != tdt or st.numel() < total:
            self._staging = st = torch.empty(int(total * 1.1) + 1024, dtype=tdt, pin_memory=True)
        if getattr(self, "_pack_pool", None) is None:
            self._pack_pool = ThreadPoolExecutor(max_workers=max(1, threads))
        flat = st.numpy()
        off = lay.utt_wav_off

        def put(lo, hi):
            for u in range(lo, hi):
                flat[off[u]:off[u + 1]] = wavs[u]

        # the packer runs ahead of the uploads in a thread of its own: chunk after chunk, each spread over the pool;
        # extract_host's `fill` hook only waits for the chunk it is about to upload
        import threading
        chunks = split_layout(lay, n_chunks, True)
        ready = {u0: threading.Event() for u0, _, _ in chunks}
        failure = []

        def packer():
            try:
                for u0, u1, _ in chunks:
                    step = max(4, (u1 - u0 + threads - 1) // threads)
                    list(self._pack_pool.map(lambda lo, u1=u1, step=step: put(lo, min(lo + step, u1)), range(u0, u1, step)))
                    ready[u0].set()
            except BaseException as e:  # noqa: BLE001
                failure.append(e)
                for ev in ready.values():
                    ev.set()

        th = threading.Thread(target=packer, daemon=True)
        th.start()

        def fill(u0, u1, w0, w1):
            ready[u0].wait()
            if failure:
                raise failure[0]

        f, fh, s_ = self.extract_host(st[:total], lay, feats_host=feats_host, n_chunks=n_chunks, cmvn=cmvn, status=status,
                                      wav_dev=wav_dev, feats=feats, fill=fill, taper=True)
        th.join()
        return lay, f, fh, s_

    def reverberate(self, wav_dev: torch.Tensor, lay: BatchLayout, reverb, noise_dev=None):
        """--add_reverb (computeFDLPSpectrogram.py:168-170): noise mixing (if any) then addReverb, on the device.
        `reverb` is a reverb.Reverberator or the float64 impulse response.  Returns (float64 packed batch,
        per-utterance lengths): an utterance comes back one sample shorter in the reference's corner case
        indM == len(rir), and the packed batch is then re-packed to the new lengths."""
        from .reverb import reverberate_packed
        lay.to(self.device)
        return reverberate_packed(reverb, wav_dev, lay.lengths, self.device, torch.float64, noise_dev,
                                  lay.dev["utt_wav_off"])

    def __call__(self, wavs: Sequence[Union[np.ndarray, torch.Tensor]], jitter=None,
                 noise: Optional[NoiseSpec] = None, cmvn: Optional[torch.Tensor] = None,
                 return_status: bool = False, reverb=None):
        """Extract FDLP spectrograms for a list of mono waveforms.  Returns a list of [Fr_u, nfilters]
        float32 device tensors (views into one buffer), in input order.  `reverb`: impulse response (float64 array,
        reverb.load_rir) or a reverb.Reverberator for --add_reverb."""
        if len(wavs) == 0:  # an empty scp shard: nothing to extract
            empty_status = torch.zeros(0, dtype=torch.int32, device=self.device)
            return ([], empty_status) if return_status else []
        host, lengths = self.pack(wavs)
        # the overlap-add jitter draws belong to the final lengths: with reverberation the bookkeeping of the
        # augmentation stage uses a throw-away layout
        lay = self.layout(lengths, jitter if reverb is None else 0)
        wav_dev = host.to(self.device, non_blocking=True)
        if isinstance(noise, str):
            if noise != "diff":
                raise ValueError("noise must be a NoiseSpec or the string 'diff'")
            wav_dev = self.diff_filter(wav_dev, lay)
            noise = None
        noise_dev = self._noise_to_device(noise, lay)
        if reverb is not None:
            wav_dev, lengths = self.reverberate(wav_dev, lay, reverb, noise_dev)
            noise_dev = None
            lay = self.layout(lengths, jitter)
        feats, status = self.forward_packed(wav_dev, lay, cmvn=cmvn, noise_dev=noise_dev)
        outs = [feats[int(lay.utt_frame_off[u]):int(lay.utt_frame_off[u + 1])] for u in range(lay.n_utt)]
        if return_status:
            return outs, status
        return outs

    # ---- stage-level access for the parity tests (same kernels, intermediate buffers exposed)
    def stage_dct(self, wavs, noise: Optional[NoiseSpec] = None):
        host, lengths = self.pack(wavs)
        lay = self.layout(lengths, 0).to(self.device)
        wav_dev = host.to(self.device)
        noise_dev = self._noise_to_device(noise, lay)
        out = torch.empty((lay.n_seg, self.geom["N"]), dtype=torch.float64, device=self.device)
        b = self._batch_struct(wav_dev, lay, None, None, None, noise_dev)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_stage_dct(self._plan, C.byref(b), C.c_void_p(out.data_ptr()), C.c_void_p(stream)))
        return out, lay

    def stage_autocorr(self, dct: torch.Tensor):
        dct = dct.contiguous()
        n_seg = dct.shape[0]
        r = torch.empty((n_seg, self.cfg.nfilters, self.geom["nlags"]), dtype=torch.float64, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_stage_autocorr(self._plan, C.c_void_p(dct.data_ptr()), n_seg,
                                                  C.c_void_p(r.data_ptr()), C.c_void_p(stream)))
        return r

    def stage_lpc_envelope(self, r: torch.Tensor):
        r = r.contiguous()
        n_seg, J = r.shape[0], self.cfg.nfilters
        kw = dict(dtype=torch.float64, device=self.device)
        lpc = torch.empty((n_seg, J, self.cfg.order + 1), **kw)
        gain = torch.empty((n_seg, J), **kw)
        cep = torch.empty((n_seg, J, self.geom["L"]), **kw)
        env = torch.empty((n_seg, J, self.geom["kk"]), **kw)
        st = torch.zeros(n_seg, dtype=torch.int32, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            L.check(self._lib.fdlp_stage_lpc_envelope(
                self._plan, C.c_void_p(r.data_ptr()), n_seg, C.c_void_p(lpc.data_ptr()), C.c_void_p(gain.data_ptr()),
                C.c_void_p(cep.data_ptr()), C.c_void_p(env.data_ptr()), C.c_void_p(st.data_ptr()), C.c_void_p(stream)))
        return dict(lpc=lpc, gain=gain, cep=cep, env=env, status=st)


def fdlp_spectrogram(wavs, *, jitter=None, noise: Optional[NoiseSpec] = None, device=None, **cfg_kwargs):
    """Functional form: `fdlp_spectrogram([wav0, wav1, ...], nfilters=80, order=150, ...)` ->
    list of [Fr, nfilters] float32 tensors on the device."""
    ex = FdlpExtractor(FdlpConfig(**cfg_kwargs), device=device)
    try:
        return ex(wavs, jitter=jitter, noise=noise)
    finally:
        ex.close()
