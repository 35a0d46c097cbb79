#!/usr/bin/env python
"""Throughput of the extractor on the other workloads SURVEY.md section 8(d) names (run on the GPU box).

bench.py times configs[1] (the WSJ recipe batch); this probe times, with the same CUDA-event method and the
batch resident in HBM, the argparse defaults (#1), the REVERB recipe with noise mixing on 125 x 8 channel streams
(#3), the long-segment stress settings (#4) and a LibriSpeech-shaped length distribution (#5, a bounded sample
extrapolated to 1000 h).  One JSON line per workload; results are copied into profiles/.
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200 import synth  # noqa: E402
from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor, NoiseSpec  # noqa: E402

COCH = "cochlear,1,1,1,2.5,1"
DEFAULTS = dict(nfilters=20, order=50, fduration=0.5, frate=100, overlap_fraction=0.25, coeff_num=50,
                coeff_range="1,20", fbank_type="mel,1")
WSJ = dict(nfilters=80, order=150, fduration=1.5, frate=100, overlap_fraction=0.25, coeff_num=100,
           coeff_range="0,100", fbank_type=COCH)
REVERB = dict(WSJ, coeff_num=450, coeff_range="1,450")


def long_cfg(fdur, J, p, C):
    return dict(nfilters=J, order=p, fduration=fdur, frate=100, overlap_fraction=0.25, coeff_num=C,
                coeff_range="0,%d" % C, fbank_type=COCH)


def time_steps(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def run(name, cfg, wavs, noise=None, steps=5, warmup=3, note="", rir=None):
    t0 = time.time()
    ex = FdlpExtractor(FdlpConfig(**cfg))
    t_plan = time.time() - t0
    host, lens = ex.pack(wavs)
    lay = ex.layout(lens, jitter=0).to(ex.device)
    wav = host.to(ex.device)
    noise_dev = ex._noise_to_device(noise, lay) if noise is not None else None
    feats = torch.empty((lay.n_frames, cfg["nfilters"]), dtype=torch.float32, device=ex.device)
    rv_ms = None
    if rir is not None:  # --add_reverb: noise mix + reverberation (float64 waveform), then the extraction
        from speech_recognition_tools_b200.reverb import Reverberator
        rv = Reverberator(rir, device=ex.device)

        def both():
            w64, _, _ = rv.forward_packed(wav, lens, lay.dev["utt_wav_off"], noise_dev)
            ex.forward_packed(w64, lay, feats=feats)
        rv_ms = time_steps(lambda: rv.forward_packed(wav, lens, lay.dev["utt_wav_off"], noise_dev), steps, warmup)
        total_ms = time_steps(both, steps, warmup)
        wav, _, _ = rv.forward_packed(wav, lens, lay.dev["utt_wav_off"], noise_dev)
        noise_dev = None
        note = (note + "; " if note else "") + "reverberation alone %.2f ms, with the extraction %.2f ms" % (rv_ms, total_ms)
    ms = time_steps(lambda: ex.forward_packed(wav, lay, feats=feats, noise_dev=noise_dev), steps, warmup)
    ex.set_profiling(True)
    ex.forward_packed(wav, lay, feats=feats, noise_dev=noise_dev)
    stage = {k: round(v, 3) for k, v in ex.stage_ms().items()}
    ex.set_profiling(False)
    audio = float(np.sum(lens)) / 16000.0
    _, status = ex.forward_packed(wav, lay, feats=feats, noise_dev=noise_dev)
    out = {"workload": name, "params": {k: cfg[k] for k in ("nfilters", "order", "fduration", "coeff_num")},
           "utterances": len(wavs), "audio_s": round(audio, 1), "segments": int(lay.n_seg),
           "ms_per_step": round(ms + (rv_ms or 0.0), 3), "xRT": round(audio / (ms + (rv_ms or 0.0)) * 1e3),
           "reverb_ms": None if rv_ms is None else round(rv_ms, 3), "stage_ms": stage,
           "plan_create_s": round(t_plan, 2), "dct_is_fp64": int(ex.geom["dct_is_fp64"]),
           "flagged_utts": int((status != 0).sum().item()), "finite": bool(torch.isfinite(feats).all().item()),
           "note": note}
    print(json.dumps(out), flush=True)
    del ex, feats, wav
    torch.cuda.empty_cache()
    return out


def cfg3_streams(n_src=125, n_ch=8):
    """125 sources x 8 channels: each channel a delayed, attenuated copy of the source (int16)."""
    lengths = synth.wsj_shaped_lengths(n_src, seed=77)
    wavs = []
    for s, T in enumerate(lengths):
        src = synth.make_utterance("speechlike" if s % 2 else "white", 3, s, int(T) + 64, 30.0).astype(np.float64)
        for c in range(n_ch):
            d, gain = 7 * c, 1.0 - 0.06 * c
            wavs.append(np.round(gain * src[d:d + int(T)]).astype(np.int16))
    return wavs


def librispeech_lengths(n, seed=5):
    rng = np.random.default_rng(seed)
    sigma = 0.45
    d = np.clip(rng.lognormal(np.log(12.3) - sigma * sigma / 2, sigma, n), 1.0, 35.0)
    return (d * 16000).astype(np.int64)


def main():
    which = set(sys.argv[1:]) or {"1", "3", "4", "5"}
    res = []
    if "1" in which:
        x = synth.white_int16(0, 0, 160000)
        res.append(run("cfg1 argparse defaults, 1 utterance of 10 s", DEFAULTS, [x], steps=20,
                       note="launch/latency-bound: 7 kernels for 27 segments"))
        # 10 s each: with the default 0.5 s segments the reference's overlap-add pointer overruns the utterance
        # for ~10 % of arbitrary lengths (ValueError there and here), so the batch uses the length cfg #1 names
        wavs = [synth.white_int16(1, u, 160000) for u in range(1000)]
        res.append(run("cfg1 argparse defaults, 1000 utterances of 10 s", DEFAULTS, wavs))
    if "3" in which:
        wavs = cfg3_streams()
        names = ["babble", "buccaneer1", "buccaneer2", "car", "destroyerops", "f16", "factory1", "factory2", "m109",
                 "machinegun", "pink", "street", "volvo", "white"]
        rs = np.random.RandomState(11)
        snr = rs.uniform(0.0, 40.0, len(wavs))
        noise = synth.noise_int16(names[0], 60 * 16000)
        for mode in ("int16_wrap", "float"):
            res.append(run("cfg3 REVERB recipe, 125 x 8 channel streams, noise mix (%s)" % mode, REVERB, wavs,
                           noise=NoiseSpec(noise, snr, rand=np.random.RandomState(5), mode=mode),
                           note="one noise recording per job as in the reference (--add_noise type,snr); per-stream snr"))
        res.append(run("cfg3 REVERB recipe, 125 x 8 channel streams, clean", REVERB, wavs))
        rir = synth.rir_int16("large_room", 16000)[:, 1] / 2.0 ** 15
        res.append(run("cfg3 REVERB recipe, 125 x 8 channel streams, noise mix (float) then --add_reverb (1 s response)",
                       REVERB, wavs, noise=NoiseSpec(noise, snr, rand=np.random.RandomState(5), mode="float"), rir=rir,
                       note="synthetic room response (synth.rir_int16); one response per job as in the reference"))
    if "4" in which:
        wavs = [synth.make_utterance("speechlike" if u % 2 else "white", 9, u, 480000, 30.0) for u in range(256)]
        for fdur, J, p, C in ((1.5, 80, 150, 100), (2.0, 128, 300, 300), (3.0, 128, 300, 300)):
            res.append(run("cfg4 long segments: 256 x 30 s, fduration %.1f" % fdur, long_cfg(fdur, J, p, C), wavs,
                           steps=3, warmup=2))
    if "5" in which:
        lengths = librispeech_lengths(1500)
        wavs = [synth.white_int16(2, u, int(T)) for u, T in enumerate(lengths)]
        o = run("cfg5 LibriSpeech-shaped lengths (lognormal 12.3 s), 1500-utterance batch", WSJ, wavs)
        o["extrapolated_1000h_s_one_gpu"] = round(3.6e6 / o["xRT"], 2)
        print(json.dumps({"workload": "cfg5 extrapolation", "seconds_for_1000h_on_one_gpu":
                          o["extrapolated_1000h_s_one_gpu"]}), flush=True)
    return res


if __name__ == "__main__":
    main()
