#!/usr/bin/env python
"""Generate tests/golden/mfcc_*.npz by running the UNMODIFIED reference MFCC extractor
(/root/reference/src/featgen/computeMfccFeatures.py).  Build container only.  Same shims as make_golden.py:
`getKaldiArk` is replaced by an in-memory capture, the input is a WAV written with scipy."""
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

import computeMfccFeatures as ref_mfcc  # noqa: E402  (the reference)

from speech_recognition_tools_b200 import synth  # noqa: E402

_captured = {}


def _capture(feat_dict, outfile, kaldi_cmd="copy-feats"):
    _captured.clear()
    _captured.update({k: np.array(v) for k, v in feat_dict.items()})


ref_mfcc.getKaldiArk = _capture

DEFAULTS = dict(nfilters=30, fduration=0.02, frate=100, context=None, nfft=1024, add_reverb=None, add_noise="none",
                kaldi_cmd="copy-feats")


def case(name, params, spec, noise=None):
    x = synth.make_utterance(spec["kind"], spec["seed"], spec["utt"], spec["T"], spec.get("sigma_f", 30.0))
    cwd = os.getcwd()
    with tempfile.TemporaryDirectory() as d:
        wav = os.path.join(d, "u.wav")
        wavfile.write(wav, 16000, x)
        scp = os.path.join(d, "wav.scp")
        with open(scp, "w") as f:
            f.write("utt1 %s\n" % wav)
        p = dict(DEFAULTS, **params)
        meta_noise = None
        if noise is not None:  # the reference reads noises/<type>.wav relative to the working directory
            os.makedirs(os.path.join(d, "noises"))
            nz = synth.noise_int16(noise["type"], noise["T"], seed=noise["seed"])
            wavfile.write(os.path.join(d, "noises", noise["type"] + ".wav"), 16000, nz)
            p["add_noise"] = "%s,%g" % (noise["type"], noise["snr"])
            np.random.seed(noise["np_seed"])
            rand = float(np.random.rand())
            np.random.seed(noise["np_seed"])
            meta_noise = dict(noise, rand=rand)
            os.chdir(d)
        try:
            args = argparse.Namespace(scp=scp, outfile=os.path.join(d, "out"), **p)
            ref_mfcc.extractMelEnergyFeats(args)
        finally:
            os.chdir(cwd)
    out = _captured["utt1"]
    meta = dict(name=name, kind="mfcc", params=params, input=spec, noise=meta_noise,
                input_sha1=hashlib.sha1(np.ascontiguousarray(x).tobytes()).hexdigest())
    np.savez_compressed(os.path.join(HERE, name + ".npz"), feats=out, meta=np.array(json.dumps(meta)))
    print("%-28s out=%s min=%.5f max=%.5f" % (name, out.shape, np.nanmin(out), np.nanmax(out)))


def main():
    case("mfcc_default_white", {}, dict(kind="white", seed=40, utt=0, T=24000))
    case("mfcc_default_speech", {}, dict(kind="speechlike", seed=41, utt=1, T=31999, sigma_f=30.0))
    case("mfcc_nfft512_context", dict(nfft=512, nfilters=23, context=2), dict(kind="speechlike", seed=42, utt=2, T=9000,
                                                                             sigma_f=300.0))
    case("mfcc_long_frame_trunc", dict(nfft=256, nfilters=20, fduration=0.025), dict(kind="white", seed=43, utt=3, T=8001))
    case("mfcc_noise_snr10", {}, dict(kind="speechlike", seed=44, utt=4, T=20000, sigma_f=30.0),
         noise=dict(type="babble", T=60000, seed=9, snr=10.0, np_seed=123))
    case("mfcc_short", {}, dict(kind="white", seed=45, utt=5, T=150))


if __name__ == "__main__":
    main()
