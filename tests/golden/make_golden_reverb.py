#!/usr/bin/env python
"""Generate tests/golden/reverb_*.npz and fdlp_addreverb_*.npz by running the UNMODIFIED reference
(/root/reference/src/featgen/features.py addReverb, and getFeats with --add_reverb).  Build container only.

The room recordings ./RIR/RIR_*_AnglA.wav are absent from the reference snapshot; synthetic two-channel int16
recordings with those names (speech_recognition_tools_b200.synth.rir_int16) are written into the working directory
the reference is run from.  Only the reference's OUTPUTS are stored."""
import json
import os
import sys
import tempfile

import numpy as np
from scipy.io import wavfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as G  # noqa: E402  (imports the reference modules and installs the capture shims)

from speech_recognition_tools_b200 import synth  # noqa: E402
from speech_recognition_tools_b200.reverb import RIR_FILES  # noqa: E402

ref_features = G.ref_features


def unit_case(name, spec, room, R, rir_seed=11):
    x = G.make_input(spec)
    rir = synth.rir_int16(room, R, rir_seed)[:, 1] / np.power(2, 15)
    out = ref_features.addReverb(x, rir)
    full = np.convolve(x, rir)
    xxc = np.correlate(x, full, "valid")
    ind = len(xxc) - int(np.argmax(xxc))
    meta = dict(name=name, input=spec, room=room, R=R, rir_seed=rir_seed, input_sha1=G.sha(x), ind=ind)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), out=out, xxc_head=xxc[:8], xxc_max=xxc.max(),
                        meta=np.array(json.dumps(meta)))
    print("%-28s T=%d R=%d ind=%d out=%s" % (name, len(x), R, ind, out.shape))


def feat_case(name, params, spec, room, R, seed=0, rir_seed=11):
    x = G.make_input(spec)
    with tempfile.TemporaryDirectory() as d:
        os.makedirs(os.path.join(d, "RIR"))
        wavfile.write(os.path.join(d, RIR_FILES[room]), 16000, synth.rir_int16(room, R, rir_seed))
        out = G.run_reference("fdlp", dict(params, add_reverb=room), x, seed, d)
    meta = dict(name=name, kind="fdlp", params=dict(params, add_reverb=room), input=spec, seed=seed, noise=None,
                input_sha1=G.sha(x), room=room, R=R, rir_seed=rir_seed)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), feats=out, meta=np.array(json.dumps(meta)))
    print("%-28s out=%s min=%.5f max=%.5f" % (name, out.shape, np.nanmin(out), np.nanmax(out)))


def mel_case(name, params, spec, room, R, rir_seed=11):
    x = G.make_input(spec)
    with tempfile.TemporaryDirectory() as d:
        os.makedirs(os.path.join(d, "RIR"))
        wavfile.write(os.path.join(d, RIR_FILES[room]), 16000, synth.rir_int16(room, R, rir_seed))
        out = G.run_reference("mel", dict(params, add_reverb=room), x, 0, d)
    meta = dict(name=name, kind="mel", params=dict(params, add_reverb=room), input=spec, seed=0, noise=None,
                input_sha1=G.sha(x), room=room, R=R, rir_seed=rir_seed)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), feats=out, meta=np.array(json.dumps(meta)))
    print("%-28s out=%s min=%.5f max=%.5f" % (name, out.shape, np.nanmin(out), np.nanmax(out)))


def mfcc_case(name, params, spec, room, R, rir_seed=11):
    import argparse
    import make_golden_mfcc as M  # imports the reference MFCC script and installs its capture shim
    x = G.make_input(spec)
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as d:
        os.makedirs(os.path.join(d, "RIR"))
        wavfile.write(os.path.join(d, RIR_FILES[room]), 16000, synth.rir_int16(room, R, rir_seed))
        wav = os.path.join(d, "u.wav")
        wavfile.write(wav, 16000, x)
        scp = os.path.join(d, "wav.scp")
        with open(scp, "w") as f:
            f.write("utt1 %s\n" % wav)
        os.chdir(d)
        try:
            args = argparse.Namespace(scp=scp, outfile=os.path.join(d, "out"), **dict(M.DEFAULTS, **params, add_reverb=room))
            M.ref_mfcc.extractMelEnergyFeats(args)
        finally:
            os.chdir(cwd)
        out = M._captured["utt1"]
    meta = dict(name=name, kind="mfcc", params=params, input=spec, noise=None, input_sha1=G.sha(x), room=room, R=R,
                rir_seed=rir_seed)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), feats=out, meta=np.array(json.dumps(meta)))
    print("%-28s out=%s min=%.5f max=%.5f" % (name, out.shape, np.nanmin(out), np.nanmax(out)))


def main():
    unit_case("reverb_small_white", dict(kind="white", seed=30, utt=0, T=3000), "small_room", 700)
    unit_case("reverb_medium_speech", dict(kind="speechlike", seed=31, utt=1, T=20000, sigma_f=30.0), "medium_room", 5000)
    unit_case("reverb_large_white", dict(kind="white", seed=32, utt=2, T=30001), "large_room", 9001)
    unit_case("reverb_short_T_lt_R", dict(kind="white", seed=33, utt=3, T=500), "small_room", 1200)
    unit_case("reverb_tiny_rir", dict(kind="white", seed=34, utt=4, T=4000), "other", 3)
    feat_case("fdlp_addreverb_cfg1", {}, dict(kind="white", seed=35, utt=5, T=40001), "small_room", 4000, seed=3)
    feat_case("fdlp_addreverb_wsj", G.WSJ, dict(kind="speechlike", seed=36, utt=6, T=64000, sigma_f=30.0),
              "large_room", 12000, seed=4)
    mel_case("mel_addreverb_recipe", G.MEL_RECIPE, dict(kind="speechlike", seed=37, utt=7, T=30000, sigma_f=30.0),
             "medium_room", 6000)
    mfcc_case("mfcc_addreverb_default", {}, dict(kind="speechlike", seed=38, utt=8, T=20000, sigma_f=30.0),
              "large_room", 5000)


if __name__ == "__main__":
    main()
