"""Shared scp / wav ingest for the CLI shims (reference: computeFDLPSpectrogram.py:119-154)."""
from __future__ import annotations

import sys

import numpy as np
from scipy.io.wavfile import read


def iter_scp(scp_path, scp_type, srate=16000):
    """Yield (uttid, signal) in scp order; unreadable entries are reported on stderr and skipped
    (the reference skips them silently, computeFDLPSpectrogram.py:131-152) -- a corrupt or empty stream, for which
    scipy raises ValueError, included: one bad scp line must not abort the shard."""
    if scp_type not in ("wav", "segment"):
        raise ValueError("Invalid type of scp type, it should be either wav or segment")
    from .pipeline import parse_scp, read_slow
    for uttid, inwav in parse_scp(scp_path):
        try:
            sr, signal = read_slow(inwav, scp_type)
        except Exception as e:  # noqa: BLE001
            print("%s: skipping %s (%r)" % (sys.argv[0], uttid, e), file=sys.stderr)
            continue
        assert sr == srate, "Input file has different sampling rate."
        if signal.ndim != 1:
            raise ValueError("%s: only mono input is supported (the reference pads every axis of a "
                             "multi-channel array)" % uttid)
        if signal.dtype not in (np.int16, np.float32):
            signal = signal.astype(np.float32)
        yield uttid, signal


def load_noise(noise_type):
    """features.py:34-44 (load_noise): noises/<type>.wav relative to the working directory."""
    import os
    noise_file = "noises/" + noise_type + ".wav"
    if not os.path.isfile(noise_file):
        print("Noise file " + noise_file + " not found!")
        sys.exit(1)
    sr, noise = read(noise_file)
    return noise


def write_lens(outfile, lens):
    with open(outfile + ".len", "w+") as f:
        for key, n in lens.items():
            f.write("{:s} {:d}".format(key, n))
            f.write("\n")
