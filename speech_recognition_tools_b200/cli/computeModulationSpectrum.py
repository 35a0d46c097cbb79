#!/usr/bin/env python
"""Drop-in for /root/reference/src/featgen/computeModulationSpectrum.py (same flags :206-224, same outputs),
computed by libfdlp_b200.so on the GPU.  `getFeats(args)` keeps the reference's signature; `--complex_modulation`
(complex LPC on the frames' ifft) is supported (modspec.py)."""
import argparse
import os
import sys

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from speech_recognition_tools_b200.cli.common import iter_scp  # noqa: E402
from speech_recognition_tools_b200.kaldi_io import write_ark_scp  # noqa: E402
from speech_recognition_tools_b200.modspec import ModSpecConfig, ModSpecExtractor  # noqa: E402
from speech_recognition_tools_b200.reverb import Reverberator, load_rir  # noqa: E402

MAX_UTTS_PER_BATCH = 64


def get_args(argv=None):
    parser = argparse.ArgumentParser("Extract FDLP Modulation Spectral Features.")
    parser.add_argument("scp", help='"scp" list')
    parser.add_argument("outfile", help="output file")
    parser.add_argument("--scp_type", default="wav", help="scp type can be 'wav' or 'segment'")
    parser.add_argument("--nfilters", type=int, default=15, help="number of filters (15)")
    parser.add_argument("--coeff_0", type=int, default=5, help="starting coefficient")
    parser.add_argument("--coeff_n", type=int, default=30, help="ending coefficient")
    parser.add_argument("--keep_even", action="store_true", help="Keep only even coefficients")
    parser.add_argument("--order", type=int, default=50, help="LPC filter order (50)")
    parser.add_argument("--fduration", type=float, default=0.5, help="Window length (0.5 sec)")
    parser.add_argument("--frate", type=int, default=100, help="Frame rate (100 Hz)")
    parser.add_argument("--add_reverb", help='input "clean" OR "small_room" OR "large_room"')
    parser.add_argument("--fbank_type", type=str, default="mel,1",
                        help="mel,warp_fact OR cochlear,om_w,alpa,fixed,beta,warp_fact")
    parser.add_argument("--set_unity_gain", action="store_true", help="Set LPC gain to 1 (True)")
    parser.add_argument("--no_window", action="store_true", help="Keeps the square window")
    parser.add_argument("--complex_modulation", action="store_true", help="Computes modulation by fft and not dct")
    parser.add_argument("--compensate_noise", action="store_true", help="Compensate 1/f noise in modulation spectrum")
    parser.add_argument("--absolute_value", action="store_true", help="Compute absolute value of modulation spectrum")
    parser.add_argument("--kaldi_cmd", help="Kaldi command to use to get ark files")
    return parser.parse_args(argv)


def getFeats(args, srate=16000, window=np.hanning, return_feats=False):
    if window is not np.hanning:
        raise NotImplementedError("only np.hanning (and --no_window) of the reference are implemented")
    # computeModulationSpectrum.py:94-105: ./RIR/RIR_{SmallRoom1_near,LargeRoom1_far}_AnglA.wav, channel 1, / 2^15
    # (this script knows no medium_room); ValueError for any other name
    if args.add_reverb and args.add_reverb not in ("clean", "small_room", "large_room"):
        raise ValueError("Invalid type of reverberation!")
    rir = load_rir(args.add_reverb)
    if args.add_reverb == "clean":
        print("%s: No reverberation added!" % sys.argv[0])
    cfg = ModSpecConfig(nfilters=args.nfilters, coeff_0=args.coeff_0, coeff_n=args.coeff_n, keep_even=args.keep_even,
                        order=args.order, fduration=args.fduration, frate=args.frate, fbank_type=args.fbank_type,
                        no_window=args.no_window, complex_modulation=args.complex_modulation,
                        compensate_noise=args.compensate_noise, absolute_value=args.absolute_value,
                        set_unity_gain=getattr(args, "set_unity_gain", False), srate=srate)
    ex = ModSpecExtractor(cfg)
    reverberator = Reverberator(rir, device=ex.device) if rir is not None else None
    all_feats = {}
    batch = []

    def flush():
        if batch:
            outs = ex([b[1] for b in batch], reverb=reverberator)
            for (k, _), o in zip(batch, outs):
                all_feats[k] = o.cpu().numpy()
            del batch[:]

    dtype = None
    for uttid, signal in iter_scp(args.scp, args.scp_type, srate):
        print("%s: Computing Features for file: %s" % (sys.argv[0], uttid))
        sys.stdout.flush()
        if batch and (len(batch) >= MAX_UTTS_PER_BATCH or signal.dtype != dtype):
            flush()
        dtype = signal.dtype
        batch.append((uttid, signal))
    flush()
    ex.close()
    write_ark_scp(all_feats.items(), args.outfile + ".ark", args.outfile + ".scp")  # dict2Ark, features.py:63-69
    if return_feats:
        return all_feats
    return None


if __name__ == "__main__":
    import time
    a = get_args()
    t0 = time.time()
    print("%s: Extracting features...." % sys.argv[0])
    getFeats(a)
    print("Execution Time: {t:.3f} seconds".format(t=time.time() - t0))
