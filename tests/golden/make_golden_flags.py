#!/usr/bin/env python
"""Round-2 goldens for the flags the shims used to refuse, from the UNMODIFIED reference:

* `computeMelSpectrum.py --add_noise type,snr` and `--add_noise diff` (computeMelSpectrum.py:69-72,134-141);
* `computeModulationSpectrum.py --add_reverb small_room` (:94-105,146-148) with a synthetic ./RIR recording;
* `computeModulationSpectrum_segments.py` (:24-118): Kaldi `segments` file, --nmodulations, --set_unity_gain, --add_reverb.

    python tests/golden/make_golden_flags.py

Same rules as make_golden.py, whose machinery this reuses: only OUTPUTS on seeded synthetic inputs are stored."""
import argparse
import json
import os
import random
import sys
import tempfile

import numpy as np
from scipy.io import wavfile

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402  (puts /root/reference/src/featgen on sys.path)
import computeModulationSpectrum as ref_mod  # noqa: E402
import computeModulationSpectrum_segments as ref_seg  # noqa: E402

from speech_recognition_tools_b200 import synth  # noqa: E402

_cap = {}


def _capture(feat_dict, outfile, kaldi_cmd="copy-feats"):
    _cap.clear()
    _cap.update({k: np.array(v) for k, v in feat_dict.items()})


ref_mod.dict2Ark = _capture
ref_seg.dict2Ark = _capture

MOD_DEFAULTS = dict(scp_type="wav", nfilters=15, coeff_0=5, coeff_n=30, keep_even=False, order=50, fduration=0.5,
                    frate=100, add_reverb=None, fbank_type="mel,1", set_unity_gain=False, no_window=False,
                    complex_modulation=False, compensate_noise=False, absolute_value=False, kaldi_cmd=None)
SEG_DEFAULTS = dict(nfilters=15, nmodulations=12, order=50, fduration=0.5, frate=100, add_reverb=None,
                    set_unity_gain=False, kaldi_cmd=None)


def _save(name, arrays, meta):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), meta=np.array(json.dumps(meta)), **arrays)
    print(name, {k: v.shape for k, v in arrays.items()})


def modspec_reverb():
    spec = dict(kind="speechlike", seed=81, utt=0, T=20000, sigma_f=30.0)
    x = mg.make_input(spec)
    room, R, rs = "small_room", 700, 4
    with tempfile.TemporaryDirectory() as d:
        os.makedirs(os.path.join(d, "RIR"))
        wavfile.write(os.path.join(d, "RIR", "RIR_SmallRoom1_near_AnglA.wav"), 16000, synth.rir_int16(room, R, rs))
        wavfile.write(os.path.join(d, "u.wav"), 16000, x)
        with open(os.path.join(d, "wav.scp"), "w") as f:
            f.write("utt1 %s\n" % os.path.join(d, "u.wav"))
        cwd = os.getcwd()
        os.chdir(d)
        try:
            params = dict(nfilters=6, coeff_0=1, coeff_n=12, order=20, add_reverb=room)
            args = argparse.Namespace(scp=os.path.join(d, "wav.scp"), outfile=os.path.join(d, "o"), **dict(MOD_DEFAULTS, **params))
            ref_mod.getFeats(args)
        finally:
            os.chdir(cwd)
    meta = dict(name="modspec_addreverb", params=params, input=spec, input_sha1=mg.sha(x), room=room, R=R, rir_seed=rs)
    _save("modspec_addreverb", dict(feats=_cap["utt1"]), meta)


def segments_cases():
    spec = dict(kind="speechlike", seed=82, utt=0, T=64000, sigma_f=30.0)
    x = mg.make_input(spec)
    spec2 = dict(kind="white", seed=83, utt=1, T=40000)
    x2 = mg.make_input(spec2)
    segs = [("seg_a", "w1", 0.25, 1.3), ("seg_b", "w1", 1.5, 3.1), ("seg_c", "w2", 0.0, 0.9), ("seg_d", "w1", 2.0, 2.8)]
    room, R, rs = "large_room", 900, 5
    for name, params in (("modsegs_default", dict()),
                         ("modsegs_unity_reverb", dict(nfilters=8, nmodulations=9, order=24, set_unity_gain=True,
                                                       add_reverb=room))):
        with tempfile.TemporaryDirectory() as d:
            os.makedirs(os.path.join(d, "RIR"))
            wavfile.write(os.path.join(d, "RIR", "RIR_LargeRoom1_far_AnglA.wav"), 16000, synth.rir_int16(room, R, rs))
            wavfile.write(os.path.join(d, "w1.wav"), 16000, x)
            wavfile.write(os.path.join(d, "w2.wav"), 16000, x2)
            with open(os.path.join(d, "wav.scp"), "w") as f:
                f.write("w1 %s\nw2 %s\n" % (os.path.join(d, "w1.wav"), os.path.join(d, "w2.wav")))
            with open(os.path.join(d, "segments"), "w") as f:
                for s in segs:
                    f.write("%s %s %g %g\n" % s)
            cwd = os.getcwd()
            os.chdir(d)
            try:
                args = argparse.Namespace(scp=os.path.join(d, "wav.scp"), segment=os.path.join(d, "segments"),
                                          outfile=os.path.join(d, "o"), **dict(SEG_DEFAULTS, **params))
                ref_seg.getFeats(args)
            finally:
                os.chdir(cwd)
        meta = dict(name=name, params=params, inputs=[spec, spec2], input_sha1=[mg.sha(x), mg.sha(x2)],
                    segments=[list(s) for s in segs], room=room, R=R, rir_seed=rs)
        _save(name, {k: v for k, v in _cap.items()}, meta)


def mel_noise():
    mg.case("mel_noise_babble10", "mel", dict(mg.MEL_RECIPE, add_noise="babble,10"),
            dict(kind="speechlike", seed=84, utt=0, T=30000, sigma_f=30.0), seed=21, noise=dict(name="babble", T=90000))
    mg.case("mel_noise_diff", "mel", dict(add_noise="diff"), dict(kind="white", seed=85, utt=1, T=12000), seed=22)
    mg.case("mel_noise_float32", "mel", dict(add_noise="pink,0"),
            dict(kind="white", seed=86, utt=2, T=16000, dtype="float32"), seed=23, noise=dict(name="pink", T=50000))


if __name__ == "__main__":
    random.seed(0)
    mel_noise()
    modspec_reverb()
    segments_cases()
