#!/usr/bin/env python
"""Developer stress run of the tensor-core MFCC kernel (GPU box): random ragged batches, frame counts around the 128-frame
tile, repeated launches back to back; every result against the FP32-pipe kernel."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200.mfcc import MfccConfig, MfccExtractor  # noqa: E402


def main():
    rng = np.random.default_rng(7)
    ex = MfccExtractor(MfccConfig(), device="cuda:0")
    worst = 0.0
    frames_seen = set()
    for it in range(int(sys.argv[1]) if len(sys.argv) > 1 else 150):
        n_utt = int(rng.integers(1, 40))
        target = [1, 2, 127, 128, 129, 255, 256, 257, 1000, 20000][it % 10]
        lens = rng.integers(2, 2 * 160 * max(2, target) // n_utt + 3, n_utt)
        wavs = [(rng.standard_normal(int(T)) * rng.uniform(10, 8000)).astype(np.int16) for T in lens]
        assert ex.set_tensor_cores(True)
        a = torch.cat(ex(wavs))
        for _ in range(3):          # back-to-back launches on the same plan
            a2 = torch.cat(ex(wavs))
            assert torch.equal(a, a2), "tensor-core kernel is not deterministic"
        ex.set_tensor_cores(False)
        b = torch.cat(ex(wavs))
        frames_seen.add(int(a.shape[0]))
        fin = torch.isfinite(b)
        assert torch.equal(fin, torch.isfinite(a)), "different finite pattern"
        if fin.any():
            worst = max(worst, float((a[fin] - b[fin]).abs().max()))
    torch.cuda.synchronize()
    print("iterations ok; frame counts from %d to %d; max |tc - fp32| = %.2e" % (min(frames_seen), max(frames_seen), worst))
    assert worst < 2e-4


if __name__ == "__main__":
    main()
