#!/usr/bin/env python
"""Drop-in for /root/reference/src/featgen/computeMfccFeatures.py (same flags :139-150, same outputs),
computed by libfdlp_b200.so on the GPU.  `extractMelEnergyFeats(args)` keeps the reference's name."""
import argparse
import os
import sys

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from speech_recognition_tools_b200.cli.common import iter_scp, load_noise  # noqa: E402
from speech_recognition_tools_b200.fdlp import NoiseSpec  # noqa: E402
from speech_recognition_tools_b200.kaldi_io import write_ark_scp  # noqa: E402
from speech_recognition_tools_b200.reverb import Reverberator, load_rir  # noqa: E402
from speech_recognition_tools_b200.mfcc import MfccConfig, MfccExtractor  # noqa: E402

MAX_SAMPLES_PER_BATCH = 64 * 1024 * 1024


def get_args(argv=None):
    parser = argparse.ArgumentParser("Extract Mel Energy Features")
    parser.add_argument("scp", help='"scp" list')
    parser.add_argument("outfile", help="output file")
    parser.add_argument("--nfilters", type=int, default=30, help="number of filters (30)")
    parser.add_argument("--fduration", type=float, default=0.02, help="Window length (0.02 sec)")
    parser.add_argument("--frate", type=int, default=100, help="Frame rate (100 Hz)")
    parser.add_argument("--context", type=int, help="Frame Context")
    parser.add_argument("--nfft", type=int, default=1024, help="Number of points of computing FFT")
    parser.add_argument("--add_reverb", help='input "clean" OR "small_room" OR "large_room"')
    parser.add_argument("--add_noise", help='Specify "type of noise, snr" if not specify none', default="none")
    parser.add_argument("--kaldi_cmd", help="Kaldi command to use to get ark files")
    return parser.parse_args(argv)


def extractMelEnergyFeats(args, srate=16000, window=np.hamming, return_feats=False):
    if window is not np.hamming:
        raise NotImplementedError("only the Hamming window of the reference is implemented")
    rir = None
    if args.add_reverb:
        if args.add_reverb == "clean":
            print("%s: No reverberation added!" % sys.argv[0])
        elif args.add_reverb in ("small_room", "large_room"):  # computeMfccFeatures.py:81-88 has no medium room
            rir = load_rir(args.add_reverb)
        else:
            raise ValueError("Invalid type of reverberation!")
    noise = None
    if args.add_noise != "none":
        noise_info = args.add_noise.strip().split(",")
        noise = (load_noise(noise_info[0]), float(noise_info[1]))
    ex = MfccExtractor(MfccConfig.from_args(args))
    rir_plan = [Reverberator(rir, device=ex.device)] if rir is not None else []
    all_feats = {}

    def flush(batch):
        if batch:
            nspec = None
            if noise is not None:  # one np.random.rand() per utterance, in scp order (computeMfccFeatures.py:46)
                nspec = NoiseSpec(samples=noise[0], snr_db=noise[1], rand=[np.random.rand() for _ in batch])
            outs = ex([b[1] for b in batch], noise=nspec, reverb=rir_plan[0] if rir_plan else None)
            for (k, _), o in zip(batch, outs):
                all_feats[k] = o.cpu().numpy()

    batch, total, dtype = [], 0, None
    for uttid, signal in iter_scp(args.scp, "wav", srate):
        print("%s: Computing Features for file: %s" % (sys.argv[0], uttid))
        sys.stdout.flush()
        if batch and (total + len(signal) > MAX_SAMPLES_PER_BATCH or signal.dtype != dtype):
            flush(batch)
            batch, total = [], 0
        batch.append((uttid, signal))
        total += len(signal)
        dtype = signal.dtype
    flush(batch)
    ex.close()
    write_ark_scp(all_feats.items(), args.outfile + ".ark", args.outfile + ".scp")
    if return_feats:
        return all_feats


if __name__ == "__main__":
    extractMelEnergyFeats(get_args())
