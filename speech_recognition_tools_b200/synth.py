"""Deterministic synthetic 16 kHz waveforms.

Counter-based: sample n of utterance `utt` under `seed` depends only on (seed, utt, n), through a
splitmix64 hash done in integer arithmetic, so numpy on any host produces the same int16 stream
(no dependence on a library RNG's stream stability).  Two families:
  * white      -- sum of four 16-bit uniforms (Irwin-Hall), scaled to sigma ~ 3000 LSB;
  * speechlike -- 39 harmonics of a 120+-20 Hz f0, 4 Hz AM, 0.3 Hz pause gate, white floor sigma_f
                  (ill-conditioned sub-band envelopes: the case that needs FP64).
Neither family ever produces an all-zero segment (the reference raises LinAlgError on those).
"""
from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x += np.uint64(0x9E3779B97F4A7C15)
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def _hash(seed: int, utt: int, n: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        key = (np.uint64(seed) * np.uint64(0xD1342543DE82EF95)) ^ (np.uint64(utt) * np.uint64(0x2545F4914F6CDD1D))
        return splitmix64(key + n.astype(np.uint64))


def white_int16(seed: int, utt: int, T: int, scale_q16: int = 5197) -> np.ndarray:
    """Irwin-Hall(4) of 16-bit uniforms, centred, times scale_q16/65536 (5197 -> sigma ~ 3000)."""
    h = _hash(seed, utt, np.arange(T, dtype=np.uint64))
    s = ((h & np.uint64(0xFFFF)) + ((h >> np.uint64(16)) & np.uint64(0xFFFF))
         + ((h >> np.uint64(32)) & np.uint64(0xFFFF)) + (h >> np.uint64(48))).astype(np.int64) - 2 * 65535
    v = (s * scale_q16 + 32768) >> 16
    return np.clip(v, -32768, 32767).astype(np.int16)


def speechlike_int16(seed: int, utt: int, T: int, sigma_f: float = 30.0, srate: int = 16000) -> np.ndarray:
    t = np.arange(T, dtype=np.float64) / srate
    ph0 = (utt % 17) * 0.37 + (seed % 13) * 0.11
    f0 = 120.0 + 20.0 * np.sin(2 * np.pi * 0.7 * t + ph0)
    phi = 2 * np.pi * np.cumsum(f0) / srate
    harm = np.zeros(T)
    for h in range(1, 40):
        harm += np.sin(h * phi) / h
    am = (0.5 + 0.5 * np.sin(2 * np.pi * 4.0 * t + ph0)) ** 2 * (np.sin(2 * np.pi * 0.3 * t + 2 * ph0) > -0.3)
    floor = white_int16(seed ^ 0x5EED, utt, T).astype(np.float64) * (sigma_f / 3000.0)
    x = np.round(6000.0 * am * harm + floor)
    return np.clip(x, -32768, 32767).astype(np.int16)


def make_utterance(kind: str, seed: int, utt: int, T: int, sigma_f: float = 30.0) -> np.ndarray:
    if kind == "white":
        return white_int16(seed, utt, T)
    if kind == "speechlike":
        return speechlike_int16(seed, utt, T, sigma_f)
    raise ValueError("unknown synthetic family %r" % kind)


def wsj_shaped_lengths(n_utt: int, seed: int = 1234, lo: int = 32000, hi: int = 240000) -> np.ndarray:
    """BASELINE.json configs[1]: utterance lengths uniform in 2..15 s (cfg #2)."""
    return np.random.default_rng(seed).integers(lo, hi + 1, size=n_utt)


def noise_int16(name: str, T: int, seed: int = 7) -> np.ndarray:
    """Synthetic stand-in for noises/<name>.wav (absent from the reference snapshot,
    /root/reference/.MISSING_LARGE_BLOBS): white noise through a one-pole low-pass whose corner
    depends on the name, int16."""
    names = ["babble", "buccaneer1", "buccaneer2", "car", "destroyerops", "f16", "factory1", "factory2",
             "m109", "machinegun", "pink", "street", "volvo", "white"]
    idx = names.index(name) if name in names else len(name) % len(names)
    w = white_int16(seed, 1000 + idx, T).astype(np.float64)
    a = 0.0 if name == "white" else 0.5 + 0.035 * idx
    if a > 0:
        import scipy.signal
        w = scipy.signal.lfilter([1 - a], [1, -a], w) * (1.0 / np.sqrt((1 - a) / (1 + a)))
    return np.clip(np.round(w), -32768, 32767).astype(np.int16)
