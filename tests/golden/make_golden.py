#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference extractor.

Run in the build container only (it needs /root/reference):
    python tests/golden/make_golden.py
The reference modules are imported from /root/reference/src/featgen as they are.  The shims are
exactly the ones SURVEY.md section 8c lists: `dict2Ark` / `get_kaldi_ark` are replaced by an
in-memory capture (Kaldi's copy-feats is not installed), inputs are WAV files written with
scipy.io.wavfile, `random.seed` / `np.random.seed` are set before each call, and the 6-second
`createFbankCochlear` build is memoised.  Nothing from the reference is copied into the repo: only
its OUTPUTS on seeded synthetic inputs are stored.
"""
import argparse
import hashlib
import json
import os
import random
import sys
import tempfile

import numpy as np
import scipy.fftpack
from scipy.io import wavfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/src/featgen")

import features as ref_features  # noqa: E402  (the reference)
import computeFDLPSpectrogram as ref_fdlp  # noqa: E402
import computeMelSpectrum as ref_mel  # noqa: E402

from speech_recognition_tools_b200 import synth  # noqa: E402

_captured = {}


def _capture(feat_dict, outfile, kaldi_cmd="copy-feats"):
    _captured.clear()
    _captured.update({k: np.array(v) for k, v in feat_dict.items()})


ref_fdlp.dict2Ark = _capture
ref_mel.get_kaldi_ark = _capture

_coch_cache = {}
_orig_coch = ref_features.createFbankCochlear


def _memo_coch(*a, **kw):
    key = (a, tuple(sorted(kw.items())))
    if key not in _coch_cache:
        _coch_cache[key] = _orig_coch(*a, **kw)
    return _coch_cache[key]


ref_fdlp.createFbankCochlear = _memo_coch
ref_mel.createFbankCochlear = _memo_coch

FDLP_DEFAULTS = dict(scp_type="wav", nfilters=20, coeff_num=50, coeff_range="1,20", order=50, fduration=0.5,
                     frate=100, overlap_fraction=0.25, kaldi_cmd="copy-feats", add_reverb=None,
                     fbank_type="mel,1", odd_mod_zero=False, gamma_weight="None", lifter_config=None,
                     write_utt2num_frames=True, add_noise=None)
WSJ = dict(nfilters=80, order=150, fduration=1.5, coeff_range="0,100", coeff_num=100, overlap_fraction=0.25,
           fbank_type="cochlear,1,1,1,2.5,1", add_reverb="clean", add_noise="clean")
REVERB = dict(WSJ, coeff_num=450, coeff_range="1,450")
MEL_DEFAULTS = dict(scp_type="wav", spectrum_type="log", nfilters=23, fduration=0.02, frate=100, nfft=1024,
                    add_reverb=None, fbank_type="mel,1", write_utt2num_frames=True, add_noise=None)
MEL_RECIPE = dict(nfilters=80, nfft=2048, fbank_type="mel,1", add_reverb="clean", add_noise="clean")


def make_input(spec):
    x = synth.make_utterance(spec["kind"], spec["seed"], spec["utt"], spec["T"], spec.get("sigma_f", 30.0))
    if spec.get("dtype", "int16") == "float32":
        x = x.astype(np.float32)
    return x


def sha(x):
    return hashlib.sha1(np.ascontiguousarray(x).tobytes()).hexdigest()


def run_reference(kind, params, x, seed, workdir, noise=None):
    """Write the WAV + scp, seed the RNGs, run the reference entry point, return its output array."""
    wav = os.path.join(workdir, "u.wav")
    wavfile.write(wav, 16000, x)
    scp = os.path.join(workdir, "wav.scp")
    with open(scp, "w") as f:
        f.write("utt1 %s\n" % wav)
    if noise is not None:
        os.makedirs(os.path.join(workdir, "noises"), exist_ok=True)
        wavfile.write(os.path.join(workdir, "noises", noise["name"] + ".wav"), 16000, noise["data"])
    cwd = os.getcwd()
    os.chdir(workdir)
    try:
        random.seed(seed)
        np.random.seed(seed)
        if kind == "fdlp":
            args = argparse.Namespace(scp=scp, outfile=os.path.join(workdir, "out"), **dict(FDLP_DEFAULTS, **params))
            ref_fdlp.getFeats(args)
        else:
            args = argparse.Namespace(scp=scp, outfile=os.path.join(workdir, "out"), **dict(MEL_DEFAULTS, **params))
            ref_mel.compute_mel_spectrum(args)
    finally:
        os.chdir(cwd)
    return _captured["utt1"]


def case(name, kind, params, spec, seed=0, noise=None, extra=None, feats_dtype=None):
    """feats_dtype="float32": store the reference output rounded to float32 (long utterances: the product's output
    type, half the fixture size); default float64 as the reference returns it."""
    x = make_input(spec)
    with tempfile.TemporaryDirectory() as d:
        lifter = None
        p = dict(params)
        if p.get("lifter_values") is not None:
            lifter = p.pop("lifter_values")
            lf = os.path.join(d, "lifter_config")
            with open(lf, "w") as f:
                f.write(",".join("%r" % float(v) for v in lifter) + "\n")
            p["lifter_config"] = lf
        noise_obj = None
        if noise is not None:
            noise_obj = dict(name=noise["name"], data=synth.noise_int16(noise["name"], noise["T"]))
        out = run_reference(kind, p, x, seed, d, noise_obj)
    meta = dict(name=name, kind=kind, params=params, input=spec, seed=seed, noise=noise, input_sha1=sha(x))
    arrays = dict(feats=out.astype(feats_dtype) if feats_dtype else out, meta=np.array(json.dumps(meta)))
    if extra:
        arrays.update(extra)
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **arrays)
    print("%-28s %-5s out=%s min=%.5f max=%.5f  -> %s" % (name, kind, out.shape, np.nanmin(out), np.nanmax(out),
                                                        os.path.basename(path)))
    sys.stdout.flush()


def stage_case():
    """Per-stage intermediates from the reference's own helper functions."""
    spec = dict(kind="speechlike", seed=3, utt=5, T=30000, sigma_f=30.0)
    x = make_input(spec)
    fdur, order, C, J = 0.5, 50, 50, 20
    of = 1 - 0.25
    lfr = 1 / (of * fdur)
    frames = np.array(list(ref_features.getFrames(x, 16000, lfr, fdur, np.hamming)))
    cos_trans = scipy.fftpack.dct(frames) / np.sqrt(2 * int(16000 * fdur))
    fbank = ref_features.createFbank(J, int(2 * fdur * 16000), 16000, warp_fact=1.0)
    bands = [0, 7, 19]
    segs = [0, 2, frames.shape[0] - 1]
    r = np.zeros((len(segs), len(bands), order + 2))
    lpc = np.zeros((len(segs), len(bands), order + 1))
    gg = np.zeros((len(segs), len(bands)))
    cep = np.zeros((len(segs), len(bands), C))
    for a, i in enumerate(segs):
        for b, j in enumerate(bands):
            bd = fbank[j, 0:-1] * cos_trans[i, :]
            y = np.real(np.fft.ifft(np.fft.fft(bd, len(bd)) * np.conj(np.fft.fft(bd, len(bd)))))
            r[a, b] = y[:order + 2]
            xl, g = ref_features.computeLpcFast(bd, order)
            lpc[a, b], gg[a, b] = xl, g
            cep[a, b] = ref_features.computeModSpecFromLpc(g, xl.copy(), C)
    meta = dict(name="stages_default", input=spec, bands=bands, segs=segs, fduration=fdur, order=order,
                coeff_num=C, nfilters=J, input_sha1=sha(x))
    np.savez_compressed(os.path.join(HERE, "stages_default.npz"), frames_head=frames[:, :64],
                        frames_tail=frames[:, -64:], frames_shape=np.array(frames.shape),
                        dct_sel=cos_trans[segs][:, ::37], r=r, lpc=lpc, gg=gg, cep=cep,
                        meta=np.array(json.dumps(meta)))
    print("stages_default               frames=%s" % (frames.shape,))


def fbank_case():
    """Filterbank samples (every 53rd column, full first/last columns) from the reference builders."""
    mel = ref_features.createFbank(20, 16000, 16000, warp_fact=1.0)
    mel80 = ref_features.createFbank(80, 2048, 16000, warp_fact=1.0)
    coch = _memo_coch(80, 48000, 16000, om_w=1.0, alp=1.0, fixed=1, bet=2.5, warp_fact=1.0)
    coch_nf = ref_features.createFbankCochlear(12, 4000, 16000, om_w=0.2, alp=2.5, fixed=0, bet=2.5, warp_fact=1.3)
    np.savez_compressed(os.path.join(HERE, "fbanks.npz"), mel20_16000=mel[:, ::53], mel20_rowsum=mel.sum(axis=1),
                        mel80_2048=mel80, coch80_48000=coch[:, ::53], coch80_rowsum=coch.sum(axis=1),
                        coch12_4000_nonfixed=coch_nf)
    print("fbanks                       mel %s coch %s" % (mel.shape, coch.shape))


def main():
    fbank_case()
    stage_case()
    # --- configs[0]: reference defaults, 10 s white (SURVEY.md 8c anchor: shape (1000, 20)) ---
    case("fdlp_cfg1_white_10s", "fdlp", {}, dict(kind="white", seed=0, utt=0, T=160000), seed=0)
    # --- OLA branch coverage with default params: nseg = 1, 2, 3, >=4; Fr < kkb2; T = k*S +- 1 ---
    for T in (100, 3000, 6001, 12000, 12001, 18500, 40001):
        case("fdlp_cfg1_T%d" % T, "fdlp", {}, dict(kind="white", seed=1, utt=T % 97, T=T), seed=T)
    case("fdlp_cfg1_speech_5s", "fdlp", {}, dict(kind="speechlike", seed=2, utt=3, T=80000, sigma_f=30.0), seed=5)
    case("fdlp_cfg1_float32", "fdlp", {}, dict(kind="white", seed=4, utt=1, T=24000, dtype="float32"), seed=9)
    # --- options: odd_mod_zero, gamma weights (needs order == coeff_num), lifter, cochlear non-fixed alpha ---
    case("fdlp_opt_oddzero", "fdlp", dict(odd_mod_zero=True, coeff_range="0,49"),
         dict(kind="white", seed=5, utt=2, T=30000), seed=2)
    case("fdlp_opt_gamma", "fdlp", dict(gamma_weight="2,3,4", coeff_range="0,49"),
         dict(kind="white", seed=6, utt=2, T=30000), seed=3)
    lifter = list(np.round(1.0 + 0.5 * np.sin(np.arange(50) / 7.0), 6))
    case("fdlp_opt_lifter", "fdlp", dict(lifter_values=lifter, coeff_range="1,40"),
         dict(kind="white", seed=7, utt=2, T=30000), seed=4)
    case("fdlp_opt_cochlear_nonfixed", "fdlp", dict(fbank_type="cochlear,0.2,2.5,0,2.5,1.3", nfilters=12),
         dict(kind="speechlike", seed=8, utt=2, T=30000, sigma_f=300.0), seed=6)
    case("fdlp_opt_diff", "fdlp", dict(add_noise="diff"), dict(kind="white", seed=9, utt=4, T=20000), seed=8)
    # --- configs[1]: WSJ recipe params ---
    case("fdlp_wsj_white_4s", "fdlp", WSJ, dict(kind="white", seed=10, utt=0, T=70000), seed=1)
    case("fdlp_wsj_speech30_6s", "fdlp", WSJ, dict(kind="speechlike", seed=11, utt=1, T=100000, sigma_f=30.0), seed=2)
    case("fdlp_wsj_speech300_3s", "fdlp", WSJ, dict(kind="speechlike", seed=12, utt=2, T=52000, sigma_f=300.0), seed=3)
    case("fdlp_wsj_short_1s", "fdlp", WSJ, dict(kind="white", seed=13, utt=3, T=17000), seed=4)
    # --- configs[2]: REVERB recipe params (coeff_num 450 > M = 300) + noise mixing, int16-wrap and float ---
    case("fdlp_reverb_noise_int16", "fdlp", dict(REVERB, add_noise="babble,10"),
         dict(kind="speechlike", seed=14, utt=4, T=60000, sigma_f=300.0), seed=5, noise=dict(name="babble", T=200000))
    case("fdlp_cfg1_noise_int16", "fdlp", dict(add_noise="pink,5"),
         dict(kind="white", seed=15, utt=5, T=40000), seed=6, noise=dict(name="pink", T=100000))
    case("fdlp_cfg1_noise_float32", "fdlp", dict(add_noise="car,20"),
         dict(kind="white", seed=16, utt=6, T=40000, dtype="float32"), seed=7, noise=dict(name="car", T=100000))
    # --- mel baseline ---
    case("mel_default_white", "mel", {}, dict(kind="white", seed=20, utt=0, T=40000), seed=0)
    case("mel_recipe_speech", "mel", MEL_RECIPE, dict(kind="speechlike", seed=21, utt=1, T=40001, sigma_f=30.0), seed=0)
    case("mel_recipe_power", "mel", dict(MEL_RECIPE, spectrum_type="power"), dict(kind="white", seed=22, utt=2, T=8000), seed=0)
    case("mel_cochlear", "mel", dict(fbank_type="cochlear,0.2,2.5,1,2.5,1", nfilters=30),
         dict(kind="white", seed=23, utt=3, T=8001), seed=0)
    case("mel_short", "mel", {}, dict(kind="white", seed=24, utt=4, T=161), seed=0)


if __name__ == "__main__":
    main()
