#!/usr/bin/env python
"""Generate tests/golden/modspec_*.npz by running the UNMODIFIED reference modulation-spectrum extractor
(/root/reference/src/featgen/computeModulationSpectrum.py).  Build container only.  Same shims as
make_golden.py: `dict2Ark` is replaced by an in-memory capture, the input is a WAV written with scipy."""
import argparse
import hashlib
import json
import os
import sys
import tempfile

import numpy as np
from scipy.io import wavfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/src/featgen")

import computeModulationSpectrum as ref_ms  # noqa: E402  (the reference)

from speech_recognition_tools_b200 import synth  # noqa: E402

_captured = {}


def _capture(feat_dict, outfile, kaldi_cmd="copy-feats"):
    _captured.clear()
    _captured.update({k: np.array(v) for k, v in feat_dict.items()})


ref_ms.dict2Ark = _capture

DEFAULTS = dict(scp_type="wav", nfilters=15, coeff_0=5, coeff_n=30, keep_even=False, order=50, fduration=0.5,
                frate=100, add_reverb=None, fbank_type="mel,1", set_unity_gain=False, no_window=False,
                complex_modulation=False, compensate_noise=False, absolute_value=False, kaldi_cmd="copy-feats")


def case(name, params, spec):
    x = synth.make_utterance(spec["kind"], spec["seed"], spec["utt"], spec["T"], spec.get("sigma_f", 30.0))
    with tempfile.TemporaryDirectory() as d:
        wav = os.path.join(d, "u.wav")
        wavfile.write(wav, 16000, x)
        scp = os.path.join(d, "wav.scp")
        with open(scp, "w") as f:
            f.write("utt1 %s\n" % wav)
        args = argparse.Namespace(scp=scp, outfile=os.path.join(d, "out"), **dict(DEFAULTS, **params))
        ref_ms.getFeats(args)
    out = _captured["utt1"]
    meta = dict(name=name, kind="modspec", params=params, input=spec,
                input_sha1=hashlib.sha1(np.ascontiguousarray(x).tobytes()).hexdigest())
    np.savez_compressed(os.path.join(HERE, name + ".npz"), feats=out, meta=np.array(json.dumps(meta)))
    print("%-28s out=%s min=%.5f max=%.5f" % (name, out.shape, np.nanmin(out), np.nanmax(out)))


def complex_cases():
    """--complex_modulation: ifft frames, complex Hermitian LPC, complex cepstrum (computeModulationSpectrum.py:150-151,
    :164-173).  keep_even without --absolute_value is not here: the reference sizes each_feat from coeff_num but fills
    it with 2*coeff_num values there and stops with a ValueError."""
    case("modspec_complex_default", dict(complex_modulation=True), dict(kind="white", seed=34, utt=4, T=6400))
    case("modspec_complex_abs_comp", dict(complex_modulation=True, compensate_noise=True, absolute_value=True, coeff_0=1,
                                          coeff_n=20, nfilters=8), dict(kind="white", seed=35, utt=5, T=3000))
    case("modspec_complex_cochlear_even", dict(complex_modulation=True, fbank_type="cochlear,0.2,2.5,1,2.5,1", nfilters=10,
                                               coeff_0=6, coeff_n=25, keep_even=True, absolute_value=True, order=40),
         dict(kind="speechlike", seed=36, utt=6, T=4001, sigma_f=30.0))
    case("modspec_complex_square_short", dict(complex_modulation=True, no_window=True, nfilters=6, coeff_0=5, coeff_n=15),
         dict(kind="white", seed=37, utt=7, T=500))
    case("modspec_complex_speech_pad", dict(complex_modulation=True, order=20, coeff_0=1, coeff_n=30, fduration=0.25,
                                            nfilters=12), dict(kind="speechlike", seed=38, utt=8, T=5000, sigma_f=60.0))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "complex":
        return complex_cases()
    case("modspec_default_white", {}, dict(kind="white", seed=30, utt=0, T=6400))
    case("modspec_cochlear_even", dict(fbank_type="cochlear,0.2,2.5,1,2.5,1", nfilters=10, coeff_0=6, coeff_n=25,
                                       keep_even=True, order=40),
         dict(kind="speechlike", seed=31, utt=1, T=4001, sigma_f=30.0))
    case("modspec_compensate_abs", dict(compensate_noise=True, absolute_value=True, coeff_0=1, coeff_n=20, nfilters=8),
         dict(kind="white", seed=32, utt=2, T=3000))
    case("modspec_square_window_short", dict(no_window=True, nfilters=6, coeff_0=5, coeff_n=15, keep_even=True),
         dict(kind="white", seed=33, utt=3, T=500))
    complex_cases()


if __name__ == "__main__":
    main()
