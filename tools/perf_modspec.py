"""Throughput of the modulation-spectrum extractor (reference defaults: 15 mel bands, order 50, 0.5 s frames every
10 ms, coefficients 5..30) on synthetic utterances; prints audio-seconds per second (device-resident input)."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from speech_recognition_tools_b200 import synth  # noqa: E402
from speech_recognition_tools_b200.fdlp import FdlpExtractor  # noqa: E402
from speech_recognition_tools_b200.modspec import ModSpecConfig, ModSpecExtractor  # noqa: E402

n_utt = int(sys.argv[1]) if len(sys.argv) > 1 else 40
fb = sys.argv[2] if len(sys.argv) > 2 else "mel,1"
ex = ModSpecExtractor(ModSpecConfig(fbank_type=fb), device="cuda:0")
rng = np.random.default_rng(1)
wavs = [synth.make_utterance("white" if i % 2 else "speechlike", 3, i, int(T)) for i, T in
        enumerate(rng.integers(32000, 160000, n_utt))]
host, lengths = FdlpExtractor.pack(wavs)
lay = ex.layout(lengths)
wav_dev = host.cuda()
out = torch.empty((lay.n_seg, ex.cfg.nfilters * ex.feat_len), dtype=torch.float32, device="cuda:0")
for _ in range(3):
    ex.forward_packed(wav_dev, lay, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    ex.forward_packed(wav_dev, lay, out=out)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
audio = sum(lengths) / 16000.0
print("modspec fbank=%s utts=%d audio=%.0fs segments=%d  %.2f ms/step  %.0f xRT  geom=%s" %
      (fb, n_utt, audio, lay.n_seg, ms, audio / (ms * 1e-3), {k: ex.geom[k] for k in ("N", "S", "L", "nfam", "family_taps", "direct_taps", "support_taps")}))
