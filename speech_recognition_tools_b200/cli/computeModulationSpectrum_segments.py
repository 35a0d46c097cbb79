#!/usr/bin/env python
"""Drop-in for /root/reference/src/featgen/computeModulationSpectrum_segments.py (same positional arguments and flags,
:120-131; same outputs), computed by libfdlp_b200.so on the GPU.  `getFeats(args)` keeps the reference's signature.

The reference (:24-118) walks a Kaldi `segments` file (`seg_id wav_id t_beg t_end`), cuts `signal_big[int(t_beg * sr) :
int(t_end * sr)]` out of the recording named in the scp, divides by 2^15, optionally reverberates, and emits for every
10 ms frame the first `nmodulations` cepstral coefficients of the sub-band LPC models of a 0.5 s hanning window
(`--set_unity_gain`: the gain is taken as 1, i.e. c0 = 0), `nfilters x nmodulations` values per frame, one matrix per
segment.  Called by recipes/timit/local_pyspeech/make_modspec_feats.sh:56-66 for data directories with a segments file."""
import argparse
import io
import os
import subprocess
import sys

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from speech_recognition_tools_b200.kaldi_io import write_ark_scp  # noqa: E402
from speech_recognition_tools_b200.modspec import ModSpecConfig, ModSpecExtractor  # noqa: E402
from speech_recognition_tools_b200.reverb import Reverberator, load_rir  # noqa: E402

MAX_SEGMENTS_PER_BATCH = 64


def get_args(argv=None):
    parser = argparse.ArgumentParser("Extract FDLP Modulation Spectral Features with segment files")
    parser.add_argument("scp", help='"scp" list')
    parser.add_argument("segment", help="segment file")
    parser.add_argument("outfile", help="output file")
    parser.add_argument("--nfilters", type=int, default=15, help="number of filters (15)")
    parser.add_argument("--nmodulations", type=int, default=12,
                        help="number of modulations of the modulation spectrum (12)")
    parser.add_argument("--order", type=int, default=50, help="LPC filter order (50)")
    parser.add_argument("--fduration", type=float, default=0.5, help="Window length (0.5 sec)")
    parser.add_argument("--frate", type=int, default=100, help="Frame rate (100 Hz)")
    parser.add_argument("--add_reverb", help='input "clean" OR "small_room" OR "large_room"')
    parser.add_argument("--set_unity_gain", action="store_true", help="Set LPC gain to 1 (True)")
    parser.add_argument("--kaldi_cmd", help="Kaldi command to use to get ark files")
    return parser.parse_args(argv)


def _read_wav(inwav, srate):
    from scipy.io.wavfile import read
    if inwav[-1] == "|":  # :70-73 (no sample-rate check on this branch in the reference either)
        proc = subprocess.run(inwav[:-1], shell=True, stdout=subprocess.PIPE)
        return read(io.BytesIO(proc.stdout))
    sr, signal = read(inwav)
    assert sr == srate, "Input file has different sampling rate."
    return sr, signal


def getFeats(args, srate=16000, window=np.hanning, return_feats=False):
    if window is not np.hanning:
        raise NotImplementedError("only np.hanning of the reference is implemented")
    if args.add_reverb and args.add_reverb not in ("clean", "small_room", "large_room"):
        raise ValueError("Invalid type of reverberation!")  # :37-48
    rir = load_rir(args.add_reverb)
    if args.add_reverb == "clean":
        print("%s: No reverberation added!" % sys.argv[0])
    # createFbank(nfilters, int(2 * fduration * srate), srate) (:35): mel bands, warp 1; mod_spec[0 : nmodulations] (:98-99)
    cfg = ModSpecConfig(nfilters=args.nfilters, coeff_0=1, coeff_n=args.nmodulations, order=args.order,
                        fduration=args.fduration, frate=args.frate, fbank_type="mel,1", srate=srate)
    ex = ModSpecExtractor(cfg)
    reverberator = Reverberator(rir, device=ex.device) if rir is not None else None
    wav_ids, wav_locs = [], []
    with open(args.scp, "r") as fid:  # :53-59
        for line in fid:
            tokens = line.strip().split()
            if tokens:
                wav_ids.append(tokens[0])
                wav_locs.append(" ".join(tokens[1:]))
    all_feats, batch = {}, []

    def flush():
        if batch:
            outs = ex([b[1] for b in batch], reverb=reverberator, unity_gain=bool(args.set_unity_gain))
            for (k, _), o in zip(batch, outs):
                all_feats[k] = o.cpu().numpy()
            del batch[:]

    wav_in_buffer, signal_big, sr = "", None, srate
    with open(args.segment, "r") as fid_s:
        for line_s in fid_s:
            token_s = line_s.strip().split()
            if not token_s:
                continue
            seg_id, wav_id = token_s[0], token_s[1]
            if wav_in_buffer != wav_id:  # :68-78: the recording stays in memory until another one is named
                wav_in_buffer = wav_id
                sr, signal_big = _read_wav(wav_locs[wav_ids.index(wav_id)], srate)
            t_beg, t_end = int(float(token_s[2]) * sr), int(float(token_s[3]) * sr)
            # signal / 2^15 (:82): a power-of-two scale, exact in float32 for int16 and float32 samples alike
            signal = (signal_big[t_beg:t_end].astype(np.float32) * np.float32(2.0 ** -15))
            print("%s: Computing Features for file: %s and segment: %s" % (sys.argv[0], wav_id, seg_id))
            sys.stdout.flush()
            batch.append((seg_id, signal))
            if len(batch) >= MAX_SEGMENTS_PER_BATCH:
                flush()
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
    sys.stdout.flush()
    getFeats(a)
    print("Execution Time: {t:.3f} seconds".format(t=time.time() - t0))
    sys.stdout.flush()
