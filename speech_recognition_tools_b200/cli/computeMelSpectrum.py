#!/usr/bin/env python
"""Drop-in for /root/reference/src/featgen/computeMelSpectrum.py (same flags :20-37, same outputs),
computed by libfdlp_b200.so on the GPU.  `compute_mel_spectrum(args)` keeps the reference's signature."""
import argparse
import os
import sys

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from speech_recognition_tools_b200.cli.common import iter_scp, load_noise, write_lens  # noqa: E402
from speech_recognition_tools_b200.fdlp import NoiseSpec  # noqa: E402
from speech_recognition_tools_b200.kaldi_io import write_ark_scp  # noqa: E402
from speech_recognition_tools_b200.reverb import Reverberator, load_rir  # noqa: E402
from speech_recognition_tools_b200.mel import MelConfig, MelExtractor  # noqa: E402

MAX_SAMPLES_PER_BATCH = 64 * 1024 * 1024


def get_args(argv=None):
    parser = argparse.ArgumentParser("Extract Mel Energy Features")
    parser.add_argument("scp", help="scp file")
    parser.add_argument("outfile", help="output file")
    parser.add_argument("--scp_type", default="wav", help="scp type can be 'wav' or 'segment'")
    parser.add_argument("--spectrum_type", default="log", help="log/power For log spectrum or energy spectrum")
    parser.add_argument("--nfilters", type=int, default=23, help="number of filters (30)")
    parser.add_argument("--fduration", type=float, default=0.02, help="Window length (0.02 sec)")
    parser.add_argument("--frate", type=int, default=100, help="Frame rate (100 Hz)")
    parser.add_argument("--nfft", type=int, default=1024, help="Number of points of computing FFT")
    parser.add_argument("--add_reverb", help='input "clean" OR "small_room" OR "large_room"')
    parser.add_argument("--fbank_type", type=str, default="mel,1",
                        help="mel,warp_fact OR cochlear,om_w,alpa,fixed,beta,warp_fact")
    parser.add_argument("--write_utt2num_frames", action="store_true", help="Set to write utt2num_frames")
    parser.add_argument("--add_noise", help='Specify "type of noise, snr"')
    return parser.parse_args(argv)


def compute_mel_spectrum(args, srate=16000, window=np.hamming, return_feats=False):
    if window is not np.hamming:
        raise NotImplementedError("only the Hamming window of the reference recipes is implemented")
    if args.spectrum_type not in ("log", "power"):
        print("Spectrum type not supported! ")
        sys.exit(1)
    noise = None  # computeMelSpectrum.py:69-72,134-141
    if args.add_noise and args.add_noise != "clean":
        if args.add_noise == "diff":
            noise = "diff"
        else:
            noise_info = args.add_noise.strip().split(",")
            noise = (load_noise(noise_info[0]), float(noise_info[1]))
    rir = load_rir(args.add_reverb)  # computeMelSpectrum.py:74-90: ./RIR/RIR_<room>_AnglA.wav, channel 1, / 2^15
    ex = MelExtractor(MelConfig.from_args(args))
    reverberator = Reverberator(rir, device=ex.device) if rir is not None else None
    all_feats, all_lens = {}, {}

    def flush(batch):
        if batch:
            nspec = noise
            if isinstance(noise, tuple):
                draws = [np.random.rand() for _ in batch]  # features.py:25, one draw per utterance in scp order
                nspec = NoiseSpec(samples=noise[0], snr_db=noise[1], rand=draws, mode="int16_wrap")
            outs = ex([b[1] for b in batch], reverb=reverberator, noise=nspec)
            for (k, _), o in zip(batch, outs):
                all_feats[k] = o.cpu().numpy()
                all_lens[k] = int(o.shape[0])

    batch, total, dtype = [], 0, None
    for uttid, signal in iter_scp(args.scp, args.scp_type, srate):
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
    if args.write_utt2num_frames:
        write_lens(args.outfile, all_lens)
    if return_feats:
        return all_feats


if __name__ == "__main__":
    compute_mel_spectrum(get_args())
