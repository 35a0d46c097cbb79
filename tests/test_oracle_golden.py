"""Pin the numpy oracle against outputs of the unmodified reference (tests/golden/*.npz).

The reference has no tests or golden vectors of its own, so the fixtures were
produced by running its code in the build container (tests/golden/make_golden.py).  The oracle must
reproduce them to float64 round-off; frame counts and floor frames must match exactly."""
import random

import numpy as np
import pytest

from oracle import fdlp as ofdlp
from oracle import mel as omel
from oracle import reverb as oreverb
from tests.golden_util import FDLP_DEFAULTS, MEL_DEFAULTS, golden_cases, load_case, split_params

FLOOR = np.log(1e-14)


def oracle_fdlp_from_case(c):
    meta = c["meta"]
    p, add_noise, lifter = split_params(meta)
    kw = dict(FDLP_DEFAULTS, **p)
    x = c["x"]
    if add_noise and add_noise != "clean":
        if add_noise == "diff":
            x = ofdlp.diff_filter(x)
        else:
            name, snr = add_noise.split(",")
            rand = np.random.RandomState(meta["seed"]).rand()
            x = ofdlp.add_noise_to_wav(x, c["noise"]["data"], float(snr), rand=rand)
    if c["rir"] is not None:
        x = oreverb.add_reverb(x, c["rir"])
    return ofdlp.fdlp_spectrogram(x, lifter=lifter, rng=random.Random(meta["seed"]), **kw)


@pytest.mark.parametrize("name", golden_cases("fdlp"))
def test_fdlp_oracle_matches_reference(name):
    c = load_case(name)
    got = oracle_fdlp_from_case(c)
    ref = c["feats"]
    assert got.shape == ref.shape
    floor_ref = ref <= FLOOR + 1e-9
    floor_got = got <= FLOOR + 1e-9
    assert np.array_equal(floor_ref, floor_got)
    # the Toeplitz systems are ill-conditioned on speech-like input; float64 round-off differences between
    # batched and per-band FFT calls are amplified but stay far below the 1e-3 product tolerance
    np.testing.assert_allclose(got[~floor_ref], ref[~floor_ref], rtol=0, atol=2e-6)


@pytest.mark.parametrize("name", golden_cases("mel"))
def test_mel_oracle_matches_reference(name):
    c = load_case(name)
    p, add_noise, _ = split_params(c["meta"])
    kw = dict(MEL_DEFAULTS, **p)
    x = c["x"]
    if add_noise == "diff":  # computeMelSpectrum.py:134-139
        x = ofdlp.diff_filter(x)
    elif add_noise and add_noise != "clean":  # :140-141
        rand = np.random.RandomState(c["meta"]["seed"]).rand()
        x = ofdlp.add_noise_to_wav(x, c["noise"]["data"], float(add_noise.split(",")[1]), rand=rand)
    if c["rir"] is not None:
        x = oreverb.add_reverb(x, c["rir"])
    got = omel.mel_spectrogram(x, **kw)
    assert got.shape == c["feats"].shape
    np.testing.assert_allclose(got, c["feats"], rtol=1e-12, atol=1e-12)


def test_filterbanks_match_reference():
    z = np.load(__import__("os").path.join(__import__("tests.golden_util", fromlist=["GOLDEN_DIR"]).GOLDEN_DIR, "fbanks.npz"))
    mel = ofdlp.create_fbank(20, 16000, 16000, 1.0)
    np.testing.assert_allclose(mel[:, ::53], z["mel20_16000"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(mel.sum(axis=1), z["mel20_rowsum"], rtol=1e-13)
    np.testing.assert_allclose(ofdlp.create_fbank(80, 2048, 16000, 1.0), z["mel80_2048"], rtol=0, atol=1e-15)
    coch = ofdlp.create_fbank_cochlear(80, 48000, 16000, om_w=1.0, alp=1.0, fixed=1, bet=2.5, warp_fact=1.0)
    np.testing.assert_allclose(coch[:, ::53], z["coch80_48000"], rtol=1e-13, atol=0)
    np.testing.assert_allclose(coch.sum(axis=1), z["coch80_rowsum"], rtol=1e-13)
    nf = ofdlp.create_fbank_cochlear(12, 4000, 16000, om_w=0.2, alp=2.5, fixed=0, bet=2.5, warp_fact=1.3)
    np.testing.assert_allclose(nf, z["coch12_4000_nonfixed"], rtol=1e-13, atol=0)


def test_stage_intermediates_match_reference():
    import json, os
    from tests.golden_util import GOLDEN_DIR
    from speech_recognition_tools_b200 import synth
    z = np.load(os.path.join(GOLDEN_DIR, "stages_default.npz"))
    meta = json.loads(str(z["meta"]))
    spec = meta["input"]
    x = synth.make_utterance(spec["kind"], spec["seed"], spec["utt"], spec["T"], spec["sigma_f"])
    fb = ofdlp.create_fbank(meta["nfilters"], int(2 * meta["fduration"] * 16000), 16000, 1.0)
    st = ofdlp.fdlp_stages(x, nfilters=meta["nfilters"], order=meta["order"], fduration=meta["fduration"],
                           overlap_fraction=0.25, coeff_num=meta["coeff_num"], fbank=fb)
    assert list(st["frames"].shape) == list(z["frames_shape"])
    np.testing.assert_array_equal(st["frames"][:, :64], z["frames_head"])
    np.testing.assert_array_equal(st["frames"][:, -64:], z["frames_tail"])
    segs, bands = meta["segs"], meta["bands"]
    np.testing.assert_allclose(st["dct"][segs][:, ::37], z["dct_sel"], rtol=1e-12, atol=1e-9)
    for a, i in enumerate(segs):
        for b, j in enumerate(bands):
            scale = abs(z["r"][a, b, 0])
            np.testing.assert_allclose(st["r"][i, j], z["r"][a, b], rtol=0, atol=1e-12 * scale)
            np.testing.assert_allclose(st["cep"][i, j], z["cep"][a, b], rtol=1e-6, atol=1e-6)


def test_nseg_and_frame_count_formulas():
    """nseg = ceil((T-1)/S), Fr = ceil(T*frate/srate) for even N."""
    for fdur in (0.5, 1.5):
        g = ofdlp.derived_sizes(fdur, 100, 0.25)
        for T in (2, 100, g["S"], g["S"] + 1, g["S"] + 2, 3 * g["S"] + 1, 50001):
            fr = ofdlp.get_frames(np.ones(T), 16000, g["lfr"], fdur)
            assert fr.shape[0] == -(-(T - 1) // g["S"])
    assert ofdlp.derived_sizes(0.5, 100, 0.25) == dict(N=8000, S=6000, M=100, kk=50, kkb2=25, kks=38,
                                                      of=0.75, lfr=1 / (0.75 * 0.5))
    g = ofdlp.derived_sizes(1.5, 100, 0.25)
    assert (g["N"], g["S"], g["M"], g["kk"], g["kkb2"], g["kks"]) == (24000, 18000, 300, 150, 75, 112)


@pytest.mark.parametrize("name", golden_cases("modspec"))
def test_modspec_oracle_matches_reference(name):
    """oracle/modspec.py == the unmodified computeModulationSpectrum.py on the committed fixtures."""
    from oracle import modspec as omod
    c = load_case(name)
    p = dict(c["meta"]["params"])
    p.pop("add_reverb", None)
    x = c["x"] if c["rir"] is None else oreverb.add_reverb(c["x"], c["rir"])  # computeModulationSpectrum.py:146-148
    if p.pop("complex_modulation", False):
        got = omod.modulation_spectrum_complex(x, **p)  # :150-152, :164-173
    else:
        got = omod.modulation_spectrum(x, **p)
    ref = c["feats"]
    assert got.shape == ref.shape
    np.testing.assert_allclose(got, ref, rtol=1e-6, atol=1e-9)


def test_modspec_complex_keep_even_needs_absolute_value():
    """The reference sizes a band's slot from coeff_num and fills it with 2*coeff_num values when --keep_even meets
    --complex_modulation without --absolute_value (computeModulationSpectrum.py:64-71 vs :171-185) and stops with a
    ValueError; the oracle and the host mirror refuse the combination up front."""
    from oracle import modspec as omod
    from speech_recognition_tools_b200.modspec import ModSpecConfig, feature_selection
    with pytest.raises(ValueError):
        omod.modulation_spectrum_complex(np.zeros(400, dtype=np.int16), keep_even=True)
    with pytest.raises(ValueError):
        feature_selection(ModSpecConfig(complex_modulation=True, keep_even=True))
    idx, scale = feature_selection(ModSpecConfig(complex_modulation=True, keep_even=True, absolute_value=True,
                                                 compensate_noise=True, coeff_0=6, coeff_n=25))
    assert list(idx) == list(range(5, 25))[1::2]
    np.testing.assert_array_equal(scale, np.linspace(0, 20 / 0.5, 25)[idx])


@pytest.mark.parametrize("name", golden_cases("mfcc"))
def test_mfcc_oracle_matches_reference(name):
    """oracle/mfcc.py == the unmodified computeMfccFeatures.py on the committed fixtures."""
    import json
    import os
    from oracle import mfcc as omfcc
    from speech_recognition_tools_b200 import synth
    from tests.golden_util import GOLDEN_DIR
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    sp = meta["input"]
    x = synth.make_utterance(sp["kind"], sp["seed"], sp["utt"], sp["T"], sp.get("sigma_f", 30.0))
    kw = dict(meta["params"])
    if meta["noise"]:
        nz = meta["noise"]
        kw.update(noise=synth.noise_int16(nz["type"], nz["T"], seed=nz["seed"]), snr=nz["snr"], rand=nz["rand"])
    if meta.get("room"):  # computeMfccFeatures.py:113-120: signal / 2^15, then addReverb
        rir = synth.rir_int16(meta["room"], meta["R"], meta["rir_seed"])[:, 1] / np.power(2, 15)
        kw.update(reverb=rir)
    out = omfcc.mfcc_features(x, **kw)
    assert out.shape == z["feats"].shape
    np.testing.assert_allclose(out, z["feats"], rtol=0, atol=1e-10)


@pytest.mark.parametrize("name", golden_cases("reverb"))
def test_reverb_oracle_matches_reference(name):
    """addReverb (features.py:110-115): alignment index exact, reverberated waveform to float64 round-off."""
    c = load_case(name)
    out, ind = oreverb.add_reverb(c["x"], c["rir"], return_index=True)
    ref = c["z"]["out"]
    assert ind == c["meta"]["ind"]
    assert out.shape == ref.shape
    np.testing.assert_allclose(out, ref, rtol=0, atol=1e-9 * max(1.0, np.abs(ref).max()))


@pytest.mark.parametrize("name", ["modsegs_default", "modsegs_unity_reverb"])
def test_modspec_segments_oracle_matches_reference(name):
    """computeModulationSpectrum_segments.py:24-118 restated with the oracle pieces: cut, / 2^15, addReverb, hanning
    frames, mel bands, the first nmodulations cepstral coefficients (c0 = 0 under --set_unity_gain)."""
    import json
    import os
    from oracle import modspec as omod
    from speech_recognition_tools_b200 import synth
    from tests.golden_util import GOLDEN_DIR
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    prm = dict(dict(nfilters=15, nmodulations=12, order=50, fduration=0.5, frate=100), **meta["params"])
    wavs = {"w%d" % (i + 1): synth.make_utterance(sp["kind"], sp["seed"], sp["utt"], sp["T"], sp.get("sigma_f", 30.0))
            for i, sp in enumerate(meta["inputs"])}
    rir = None
    if prm.get("add_reverb"):
        rir = synth.rir_int16(meta["room"], meta["R"], meta["rir_seed"])[:, 1] / np.power(2, 15)
    for seg_id, wav_id, t0, t1 in meta["segments"]:
        x = wavs[wav_id][int(float(t0) * 16000):int(float(t1) * 16000)] / np.power(2, 15)
        if rir is not None:
            x = oreverb.add_reverb(x, rir)
        got = omod.modulation_spectrum(x, nfilters=prm["nfilters"], coeff_0=1, coeff_n=prm["nmodulations"],
                                       order=prm["order"], fduration=prm["fduration"], frate=prm["frate"])
        if prm.get("set_unity_gain"):
            got = got.reshape(got.shape[0], prm["nfilters"], prm["nmodulations"]).copy()
            got[:, :, 0] = 0.0
            got = got.reshape(got.shape[0], -1)
        assert got.shape == z[seg_id].shape
        np.testing.assert_allclose(got, z[seg_id], rtol=1e-6, atol=1e-8)
