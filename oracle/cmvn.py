"""Global CMVN statistics (TEST INFRASTRUCTURE, see oracle/__init__.py).

The reference computes them with the external Kaldi binary `compute-cmvn-stats`
(/root/reference/e2e/wsj/run_fdlp_e1.sh:280); Kaldi is not vendored and not pinned (README.md:5-12),
so this restates Kaldi's documented stats layout -- PARITY UNPINNED:
    stats[0, d] = sum_t x[t, d]      stats[0, D] = number of frames
    stats[1, d] = sum_t x[t, d]^2    stats[1, D] = 0
accumulated in float64 over every frame of every utterance.  `apply_cmvn` is Kaldi's
`apply-cmvn --norm-vars=true` (mean/variance normalisation with the global stats)."""
from __future__ import annotations

import numpy as np


def cmvn_stats(feature_mats):
    mats = [np.asarray(m, dtype=np.float64) for m in feature_mats]
    D = mats[0].shape[1]
    stats = np.zeros((2, D + 1))
    for m in mats:
        stats[0, :D] += m.sum(axis=0)
        stats[1, :D] += (m * m).sum(axis=0)
        stats[0, D] += m.shape[0]
    return stats


def apply_cmvn(feats, stats, norm_vars=True):
    D = feats.shape[1]
    n = stats[0, D]
    mean = stats[0, :D] / n
    out = np.asarray(feats, dtype=np.float64) - mean
    if norm_vars:
        var = stats[1, :D] / n - mean * mean
        out = out / np.sqrt(np.maximum(var, 1e-20))
    return out
