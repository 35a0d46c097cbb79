#!/usr/bin/env python
"""Developer probe: per-stage device times of fdlp_forward on a WSJ-shaped batch (run on the GPU box)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200 import synth  # noqa: E402
from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor  # noqa: E402

WSJ = dict(nfilters=80, order=150, fduration=1.5, frate=100, overlap_fraction=0.25, coeff_num=100,
           coeff_range="0,100", fbank_type="cochlear,1,1,1,2.5,1")


def main():
    n_utt = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    cfg = dict(WSJ)
    for kv in sys.argv[2:]:
        k, v = kv.split("=")
        cfg[k] = type(WSJ.get(k, 0))(v) if k in WSJ and not isinstance(WSJ[k], str) else v
    lengths = synth.wsj_shaped_lengths(n_utt, seed=1234)
    wavs = [synth.white_int16(1, u, int(T)) for u, T in enumerate(lengths)]
    ex = FdlpExtractor(FdlpConfig(**cfg))
    host, lens = ex.pack(wavs)
    lay = ex.layout(lens, jitter=0).to(ex.device)
    wav = host.to(ex.device)
    feats = torch.empty((lay.n_frames, cfg["nfilters"]), dtype=torch.float32, device=ex.device)
    ex.set_profiling(True)
    acc = {}
    for it in range(6):
        ex.forward_packed(wav, lay, feats=feats)
        s = dict(ex.stage_ms())
        s.update({"ac_" + k: v for k, v in ex.autocorr_ms().items()})
        if it >= 2:
            for k, v in s.items():
                acc[k] = acc.get(k, 0) + v / 4
    ex.set_profiling(False)
    for ov in (0, 1):
        ex.set_overlap(bool(ov))
        for _ in range(2):
            ex.forward_packed(wav, lay, feats=feats)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            ex.forward_packed(wav, lay, feats=feats)
        e1.record()
        torch.cuda.synchronize()
        print("overlap=%d  %.2f ms/step" % (ov, e0.elapsed_time(e1) / 5))
    g = ex.geom
    flops = 2.0 * g["nlags"] * g["support_taps"] * lay.n_seg
    audio = lengths.sum() / 16000.0
    tot = sum(v for k, v in acc.items() if not k.startswith("ac_"))
    print("geom", {k: g[k] for k in ("support_taps", "family_taps", "direct_taps", "nfam", "block_len")})
    print("n_seg=%d audio=%.0fs stage_ms=%s total=%.2fms xRT=%.0f autocorr=%.2f TFLOP/s (visited taps, support %.3f)" %
          (lay.n_seg, audio, {k: round(v, 3) for k, v in acc.items()}, tot, audio / tot * 1e3,
           flops / acc["autocorr"] / 1e9, g["support_taps"] / (cfg["nfilters"] * g["N"])))


if __name__ == "__main__":
    main()
