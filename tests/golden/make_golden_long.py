#!/usr/bin/env python
"""Round-2 goldens: the timed configurations that had no reference comparison (missing #1/#2 and
next-round item 1).  Same rules as make_golden.py (whose machinery this reuses): the UNMODIFIED reference under
/root/reference/src/featgen produces the outputs, only outputs on seeded synthetic inputs are stored.

    python tests/golden/make_golden_long.py

* BASELINE configs[3] at its largest shape: fduration 3.0, 128 cochlear bands, LPC order 300, 300 cepstral
  coefficients (N = 48000: the two-CTA float64 DCT, two lag groups, the 16-lane LPC kernel) on an 8.5 s
  speech-like utterance; the 2.0 s variant on the same input.
* LibriSpeech-shaped lengths (configs[4]) with the WSJ recipe: a 35 s speech-like and a 30 s white utterance
  (31 / 27 segments: pointer drift, the last-two-segments branch, computeFDLPSpectrogram.py:207-227).
* configs[2]-shaped streams: REVERB recipe parameters with a different noise type / SNR per stream (the reference
  takes one `type,snr` per job, features.py:24-44, so every stream is one reference run; the CUDA test mixes all
  of them in ONE ragged batch with per-stream arrays).
Features are stored as float32 (the product's output type; the comparison tolerance is 5e-6).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402

CFG4_3S = dict(nfilters=128, order=300, fduration=3.0, coeff_num=300, coeff_range="0,300", overlap_fraction=0.25,
               fbank_type="cochlear,1,1,1,2.5,1", add_reverb="clean", add_noise="clean")
CFG4_2S = dict(CFG4_3S, fduration=2.0)

# name, noise type, snr, kind, T, sigma_f  (stream lengths are ragged on purpose: 1, 2, 3 and 5 segments)
REVERB_STREAMS = [
    ("fdlp_rvb_stream0_babble5", "babble", 5, "speechlike", 47000, 300.0),
    ("fdlp_rvb_stream1_pink20", "pink", 20, "white", 17500, 30.0),
    ("fdlp_rvb_stream2_car0", "car", 0, "speechlike", 36001, 30.0),
    ("fdlp_rvb_stream3_white35", "white", 35, "speechlike", 83000, 300.0),
]


def main():
    only = set(sys.argv[1:])

    def want(name):
        return not only or name in only

    f32 = dict(feats_dtype="float32")
    if want("fdlp_cfg4_3s_128b_o300"):
        mg.case("fdlp_cfg4_3s_128b_o300", "fdlp", CFG4_3S,
                dict(kind="speechlike", seed=51, utt=0, T=136000, sigma_f=30.0), seed=11, **f32)
    if want("fdlp_cfg4_2s_128b_o300"):
        mg.case("fdlp_cfg4_2s_128b_o300", "fdlp", CFG4_2S,
                dict(kind="speechlike", seed=51, utt=0, T=136000, sigma_f=30.0), seed=12, **f32)
    if want("fdlp_wsj_speech30_35s"):
        mg.case("fdlp_wsj_speech30_35s", "fdlp", mg.WSJ,
                dict(kind="speechlike", seed=52, utt=1, T=560000, sigma_f=30.0), seed=13, **f32)
    if want("fdlp_wsj_white_30s"):
        mg.case("fdlp_wsj_white_30s", "fdlp", mg.WSJ, dict(kind="white", seed=53, utt=2, T=480001), seed=14, **f32)
    for i, (name, ntype, snr, kind, T, sf) in enumerate(REVERB_STREAMS):
        if want(name):
            mg.case(name, "fdlp", dict(mg.REVERB, add_noise="%s,%d" % (ntype, snr)),
                    dict(kind=kind, seed=54, utt=i, T=T, sigma_f=sf), seed=20 + i,
                    noise=dict(name=ntype, T=150000 + 1000 * i), **f32)


if __name__ == "__main__":
    main()
