#!/usr/bin/env python
"""Developer probe: MFCC kernel throughput on the bench batch (run on the GPU box)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200 import synth  # noqa: E402
from speech_recognition_tools_b200.fdlp import FdlpExtractor  # noqa: E402
from speech_recognition_tools_b200.mfcc import MfccConfig, MfccExtractor  # noqa: E402


def main():
    n_utt = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    lengths = synth.wsj_shaped_lengths(n_utt, seed=1234)
    wavs = [synth.white_int16(1, u, int(T)) for u, T in enumerate(lengths)]
    for cfg in (dict(), dict(nfft=512, nfilters=23), dict(context=5)):
        ex = MfccExtractor(MfccConfig(**cfg))
        host, lens = FdlpExtractor.pack(wavs)
        lay = ex.layout(lens).to(ex.device)
        wav = host.to(ex.device)
        feats = torch.empty((lay.n_frames, ex.dim), dtype=torch.float32, device=ex.device)
        for _ in range(3):
            ex.forward_packed(wav, lay, feats=feats)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            ex.forward_packed(wav, lay, feats=feats)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        audio = lengths.sum() / 16000.0
        P = ex.cfg.nfft // 2 + 1
        fma = 2.0 * min(ex.flen, P) * (P // 2 + 1) * lay.n_frames
        print("cfg=%s frames=%d %.3f ms/step xRT=%.0f  DFT %.1f TFLOP/s fp32" %
              (cfg, lay.n_frames, ms, audio / ms * 1e3, 2 * fma / ms / 1e9))
        ex.close()


if __name__ == "__main__":
    main()
