import os, sys, random
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200 import synth
from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor
from oracle import fdlp as ofdlp
WSJ = dict(nfilters=80, order=150, fduration=1.5, frate=100, overlap_fraction=0.25, coeff_num=100,
           coeff_range="0,100", fbank_type="cochlear,1,1,1,2.5,1")
fb = ofdlp.make_fbank(WSJ["fbank_type"], 80, 48000, 16000)
exs = {e: FdlpExtractor(FdlpConfig(tail_eps=e, **WSJ)) for e in (0.0, 1e-14, 1e-12, 1e-10, 1e-8)}
for e, ex in exs.items():
    print("eps", e, "support fraction", ex.geom["support_taps"] / (80 * 24000))
for kind, sf in (("speechlike", 3.0), ("speechlike", 30.0), ("white", 0)):
    x = synth.make_utterance(kind, 5, 1, 60000, sf)
    if kind == "speechlike" and sf == 3.0:
        # extra spectral tilt: strong 100 Hz hum, 80 dB above the floor
        x = np.clip(x.astype(np.float64) + 20000 * np.sin(2 * np.pi * 100 * np.arange(len(x)) / 16000), -32768, 32767).astype(np.int16)
    ref = ofdlp.fdlp_spectrogram(x, fbank=fb, rng=random.Random(1), **WSJ)
    fl = ref <= np.log(1e-14) + 1e-6
    base = None
    for e, ex in exs.items():
        got = ex([x], jitter=[1])[0].cpu().numpy().astype(np.float64)
        if base is None: base = got
        print(kind, sf, "eps", e, "vs oracle %.3e" % np.abs(got - ref)[~fl].max(), "vs eps0 %.3e" % np.abs(got - base)[~fl].max())
