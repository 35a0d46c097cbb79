"""Throughput of the --add_reverb kernels on the bench batch (1000 WSJ-shaped white utterances generated on the
device) with a synthetic room response of R samples; per-kernel CUDA-event times for several block sizes, and the
reverberated FDLP step (reverb + WSJ-recipe extraction).  Prints one JSON line per setting."""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from speech_recognition_tools_b200 import synth  # noqa: E402
from speech_recognition_tools_b200.reverb import Reverberator  # noqa: E402

n_utt = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
R = int(sys.argv[2]) if len(sys.argv) > 2 else 16000
blocks = [int(b) for b in sys.argv[3].split(",")] if len(sys.argv) > 3 else [1024, 2048, 4096]
lengths = synth.wsj_shaped_lengths(n_utt)
wav_dev, off = synth.white_int16_device(9, np.arange(n_utt), lengths, device="cuda:0")
rir = synth.rir_int16("large_room", R)[:, 1] / 2.0 ** 15
audio = float(lengths.sum()) / 16000.0
for B in blocks:
    rv = Reverberator(rir, device="cuda:0", block=B)
    rv.set_profiling(True)
    for _ in range(2):
        rv.forward_packed(wav_dev, lengths, off)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    acc = {}
    e0.record()
    for _ in range(n):
        rv.forward_packed(wav_dev, lengths, off)
        for k, v in rv.kernel_ms().items():
            acc[k] = acc.get(k, 0.0) + v / n
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    pair_off, pair_utt, conv_off, conv_utt = rv.block_layout(lengths)
    n_pair, n_conv = len(pair_utt), len(conv_utt)
    F, P, M = 2 * rv.block, rv.partitions, rv.pairs_per_cta
    rows = n_pair + n_utt * (P - 1)
    # spectrum traffic the kernels request (16 F bytes per row): fwd writes the window spectra; conv reads P + M - 1
    # of them (+ the P response spectra, L2-resident) per CTA and writes one per pair; corr reads 2 D rows per (utt, q)
    print(json.dumps(dict(workload="add_reverb", n_utt=n_utt, audio_s=audio, R=R, block=rv.block, partitions=P,
                          pairs_per_cta=M, n_pair=n_pair, ms_per_step=ms, xRT=audio / (ms * 1e-3), kernel_ms=acc,
                          scratch_gb=rv.scratch_bytes(n_pair, n_utt) / 1e9,
                          spectrum_gb=dict(fwd=rows * F * 16e-9, conv=(n_conv * (2 * P + M - 1) + n_pair) * F * 16e-9,
                                           corr=(2.0 * n_pair * -(-P // 8) + 2.0 * n_utt * P) * F * 16e-9),
                          transforms=dict(fwd=rows, conv=2 * n_pair, corr=n_utt * P))))
    rv.close()
    del rv
    torch.cuda.empty_cache()
