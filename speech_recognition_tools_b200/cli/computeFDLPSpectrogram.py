#!/usr/bin/env python
"""Drop-in for /root/reference/src/featgen/computeFDLPSpectrogram.py: same positional arguments, flags,
defaults (:241-262) and outputs (<outfile>.ark/.scp Kaldi binary float matrices, <outfile>.len), computed
by libfdlp_b200.so on the GPU.  `getFeats(args)` keeps the reference's signature.

Randomness: the overlap-add jitter is drawn from the module-level `random` generator and the noise offset
from `np.random`, in scp order, exactly like the reference -- seed them the same way to reproduce it."""
import argparse
import os
import random
import sys
import time

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from speech_recognition_tools_b200 import _lib  # noqa: E402
from speech_recognition_tools_b200.cli.common import iter_scp, load_noise, write_lens  # noqa: E402
from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor, NoiseSpec, segment_count  # noqa: E402
from speech_recognition_tools_b200.kaldi_io import write_ark_scp  # noqa: E402
from speech_recognition_tools_b200.reverb import Reverberator, load_rir  # noqa: E402

MAX_SEGMENTS_PER_BATCH = 4096


class _GlobalRandom:
    """`random.randrange(2)` of computeFDLPSpectrogram.py:225 on the module-level generator."""

    @staticmethod
    def randrange(n):
        return random.randrange(n)


def getFeats(args, srate=16000, window=np.hamming, return_feats=False):
    if window is not np.hamming:
        raise NotImplementedError("only the Hamming window of the reference recipes is implemented")
    cfg = FdlpConfig.from_args(args)
    add_noise, add_reverb = args.add_noise, args.add_reverb
    # computeFDLPSpectrogram.py:75-91: ./RIR/RIR_<room>_AnglA.wav, second channel, / 2^15; ValueError for an unknown
    # room, the file-not-found error of scipy's reader when the recordings are not in the working directory
    rir = load_rir(add_reverb)
    if add_reverb == "clean":
        print("%s: No reverberation added!" % sys.argv[0])
    reverberator = None
    noise = None
    if add_noise and add_noise not in ("clean",):
        if add_noise == "diff":
            noise = "diff"  # computeFDLPSpectrogram.py:162-164: 13-tap FIR instead of a noise recording
        else:
            noise_info = add_noise.strip().split(",")
            noise = (load_noise(noise_info[0]), float(noise_info[1]))
    if args.odd_mod_zero:
        print("%s: Ignoring odd modulations... " % sys.argv[0])

    ex = FdlpExtractor(cfg)
    g = ex.geom
    all_feats, all_lens, bad = {}, {}, []

    def flush(batch):
        if not batch:
            return
        ids = [b[0] for b in batch]
        wavs = [b[1] for b in batch]
        nspec = None
        if noise == "diff":
            nspec = "diff"
        elif noise is not None:
            draws = [np.random.rand() for _ in wavs]  # features.py:25, one draw per utterance in scp order
            nspec = NoiseSpec(samples=noise[0], snr_db=noise[1], rand=draws, mode="int16_wrap")
        nonlocal reverberator
        if rir is not None and reverberator is None:
            reverberator = Reverberator(rir, device=ex.device)
        outs, status = ex(wavs, jitter=_GlobalRandom(), noise=nspec, return_status=True, reverb=reverberator)
        st = status.cpu().numpy()
        for u, (k, o) in enumerate(zip(ids, outs)):
            all_feats[k] = o.cpu().numpy()
            all_lens[k] = int(o.shape[0])
            if st[u] & _lib.FDLP_ST_SINGULAR:
                bad.append(k)

    batch, nseg, dtype = [], 0, None
    for uttid, signal in iter_scp(args.scp, args.scp_type, srate):
        print("%s: Computing Features for file: %s" % (sys.argv[0], uttid))
        sys.stdout.flush()
        n = segment_count(len(signal), g["N"], g["S"])
        if batch and (nseg + n > MAX_SEGMENTS_PER_BATCH or signal.dtype != dtype):
            flush(batch)
            batch, nseg = [], 0
        batch.append((uttid, signal))
        nseg += n
        dtype = signal.dtype
    flush(batch)
    ex.close()
    if bad:
        # the reference raises numpy.linalg.LinAlgError('Singular principal minor') at the first such utterance
        raise np.linalg.LinAlgError("Singular principal minor in utterance(s): %s" % " ".join(bad))
    write_ark_scp(all_feats.items(), args.outfile + ".ark", args.outfile + ".scp")
    if args.write_utt2num_frames:
        write_lens(args.outfile, all_lens)
    if return_feats:
        return all_feats


def get_parser():
    parser = argparse.ArgumentParser("Extract FDLP Spectrogram.")
    parser.add_argument("scp", help='"scp" list')
    parser.add_argument("outfile", help="output file")
    parser.add_argument("--scp_type", default="wav", help="scp type can be 'wav' or 'segment'")
    parser.add_argument("--nfilters", type=int, default=20, help="number of filters (15)")
    parser.add_argument("--coeff_num", type=int, default=50, help="Total Number of coefficients to compute")
    parser.add_argument("--coeff_range", type=str, default="1,20", help="Range of Modulation coefficients to keep")
    parser.add_argument("--order", type=int, default=50, help="LPC filter order (50)")
    parser.add_argument("--fduration", type=float, default=0.5, help="Window length (0.5 sec)")
    parser.add_argument("--frate", type=int, default=100, help="Frame rate (100 Hz)")
    parser.add_argument("--overlap_fraction", type=float, default=0.25, help="Fraction of Overlap for OLA")
    parser.add_argument("--kaldi_cmd", default="copy-feats", help="accepted for compatibility; the ark is written natively")
    parser.add_argument("--add_reverb", help='input "clean" OR "small_room" OR "large_room"')
    parser.add_argument("--fbank_type", type=str, default="mel,1",
                        help="mel,warp_fact OR cochlear,om_w,alpa,fixed,beta,warp_fact")
    parser.add_argument("--odd_mod_zero", action="store_true", help="Ignore the odd modulation coefficients")
    parser.add_argument("--gamma_weight", type=str, default="None", help="Configured as scale,shape,pk")
    parser.add_argument("--lifter_config", type=str, default=None, help="Configuration for general liftering")
    parser.add_argument("--write_utt2num_frames", action="store_true", help="Set to write utt2num_frames")
    parser.add_argument("--add_noise", help='Specify "type of noise, snr"')
    return parser


if __name__ == "__main__":
    args = get_parser().parse_args()
    start_time = time.time()
    print("%s: Extracting features...." % sys.argv[0])
    sys.stdout.flush()
    getFeats(args)
    print("Execution Time: {t:.3f} seconds".format(t=time.time() - start_time))
    sys.stdout.flush()
