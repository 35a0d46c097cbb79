"""GPU parity tests of the modulation-spectrum extractor against the golden outputs of the
unmodified reference (tests/golden/modspec_*.npz) and the oracle.  The CUDA path computes in float64 and stores
float32: tolerance 2e-6 absolute + 2e-6 relative on cepstral coefficients of magnitude <= ~5."""
import numpy as np
import pytest
import torch

from tests.golden_util import golden_cases, load_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", golden_cases("modspec"))
def test_modspec_matches_reference_golden(name):
    from speech_recognition_tools_b200.modspec import ModSpecConfig, ModSpecExtractor
    c = load_case(name)
    prm = dict(c["meta"]["params"])
    prm.pop("add_reverb", None)  # computeModulationSpectrum.py:146-148: addReverb with the case's ./RIR recording
    ex = ModSpecExtractor(ModSpecConfig(**prm), device="cuda:0")
    outs, st = ex([c["x"]], return_status=True, reverb=c["rir"])
    assert int(st.cpu()[0]) == 0
    got = outs[0].cpu().numpy().astype(np.float64)
    assert got.shape == c["feats"].shape
    np.testing.assert_allclose(got, c["feats"], rtol=2e-6, atol=2e-6)
    ex.close()


def test_modspec_batch_equals_single_and_oracle():
    """Ragged batch (empty list, 2-sample and sub-frame utterances, a longer one) == single utterances, == oracle;
    the batch is split into runs of segments (max_segments) without changing the result."""
    from oracle import modspec as omod
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.modspec import ModSpecConfig, ModSpecExtractor
    kw = dict(nfilters=10, coeff_0=3, coeff_n=20, order=30, fbank_type="cochlear,1,1,1,2.5,1")
    ex = ModSpecExtractor(ModSpecConfig(**kw), device="cuda:0")
    assert ex([]) == []
    wavs = [synth.make_utterance("white" if i % 2 else "speechlike", 7, i, T) for i, T in enumerate([2, 161, 1200, 4000])]
    outs = ex(wavs)
    split = ex(wavs, max_segments=8)
    for w, o, s in zip(wavs, outs, split):
        assert torch.equal(ex([w])[0], o)
        assert torch.equal(s, o)
        ref = omod.modulation_spectrum(w, **kw)
        assert o.shape == ref.shape
        # 2- and 161-sample utterances reflected to 8000 samples are exactly periodic: the Toeplitz systems are
        # singular to working precision and both implementations return round-off there, so only the shape (and
        # the bit-for-bit batch / single agreement above) is checked for them
        if len(w) >= 1200:
            np.testing.assert_allclose(o.cpu().numpy(), ref, rtol=2e-6, atol=2e-6)
    ex.close()


def test_modspec_cli_matches_golden(tmp_path):
    import argparse
    from scipy.io import wavfile
    from speech_recognition_tools_b200.cli import computeModulationSpectrum as cli
    from speech_recognition_tools_b200.kaldi_io import read_ark
    c = load_case("modspec_cochlear_even")
    wav = tmp_path / "u.wav"
    wavfile.write(str(wav), 16000, c["x"])
    scp = tmp_path / "wav.scp"
    scp.write_text("utt1 %s\n" % wav)
    args = cli.get_args([str(scp), str(tmp_path / "out")])
    for k, v in c["meta"]["params"].items():
        setattr(args, k, v)
    cli.getFeats(args)
    got = dict(read_ark(str(tmp_path / "out.ark")))["utt1"]
    np.testing.assert_allclose(got, c["feats"], rtol=2e-6, atol=2e-6)


@pytest.mark.parametrize("name", ["modsegs_default", "modsegs_unity_reverb"])
def test_modspec_segments_cli_matches_reference_golden(name, tmp_path, monkeypatch):
    """cli/computeModulationSpectrum_segments.py == the unmodified computeModulationSpectrum_segments.py on a scp of two
    recordings and a Kaldi segments file (segments of both, out of order; --set_unity_gain; --add_reverb)."""
    import json
    import os
    from scipy.io import wavfile
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.cli import computeModulationSpectrum_segments as cli
    from speech_recognition_tools_b200.kaldi_io import read_ark
    from tests.golden_util import GOLDEN_DIR
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    os.makedirs(tmp_path / "RIR")
    wavfile.write(str(tmp_path / "RIR" / "RIR_LargeRoom1_far_AnglA.wav"), 16000,
                  synth.rir_int16(meta["room"], meta["R"], meta["rir_seed"]))
    with open(tmp_path / "wav.scp", "w") as f:
        for i, sp in enumerate(meta["inputs"]):
            x = synth.make_utterance(sp["kind"], sp["seed"], sp["utt"], sp["T"], sp.get("sigma_f", 30.0))
            wavfile.write(str(tmp_path / ("w%d.wav" % (i + 1))), 16000, x)
            f.write("w%d %s\n" % (i + 1, tmp_path / ("w%d.wav" % (i + 1))))
    with open(tmp_path / "segments", "w") as f:
        for s_ in meta["segments"]:
            f.write("%s %s %g %g\n" % tuple(s_))
    argv = [str(tmp_path / "wav.scp"), str(tmp_path / "segments"), str(tmp_path / "o")]
    for k, v in meta["params"].items():
        argv += ["--%s" % k] if v is True else ["--%s=%s" % (k, v)]
    monkeypatch.chdir(tmp_path)
    cli.getFeats(cli.get_args(argv))
    ark = read_ark(str(tmp_path / "o.ark"))
    assert list(ark) == [s_[0] for s_ in meta["segments"]]
    for seg_id in ark:
        assert ark[seg_id].shape == z[seg_id].shape
        np.testing.assert_allclose(ark[seg_id], z[seg_id], rtol=2e-6, atol=2e-6)


def test_modspec_complex_batch_equals_single_and_oracle(tmp_path):
    """--complex_modulation on a ragged batch (one utterance shorter than a frame; int16 and float32 input; dense
    cochlear bands / mel bands with exact zeros): batch == single bit for bit, == oracle/modspec.py; then the
    same through the CLI shim."""
    from oracle import modspec as omod
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.modspec import ModSpecConfig, ModSpecExtractor
    for kw in (dict(nfilters=9, coeff_0=2, coeff_n=24, order=36, fbank_type="cochlear,0.2,2.5,1,2.5,1", fduration=0.25),
               dict(nfilters=7, coeff_0=1, coeff_n=12, order=16, fbank_type="mel,1", absolute_value=True,
                    compensate_noise=True, fduration=0.1)):
        ex = ModSpecExtractor(ModSpecConfig(complex_modulation=True, **kw), device="cuda:0")
        wavs = [synth.make_utterance("white" if i % 2 else "speechlike", 17, i, T, 40.0) for i, T in enumerate([900, 2500, 5001])]
        if "absolute_value" in kw:   # float32 input in the second configuration (a batch has one sample type)
            wavs = [w.astype(np.float32) for w in wavs]
        outs, st = ex(wavs, return_status=True)
        assert not st.cpu().numpy().any()
        split = ex(wavs, max_segments=10)
        for w, o, s in zip(wavs, outs, split):
            assert torch.equal(ex([w])[0], o) and torch.equal(s, o)
            ref = omod.modulation_spectrum_complex(w, **kw)
            assert o.shape == ref.shape
            np.testing.assert_allclose(o.cpu().numpy(), ref, rtol=2e-6, atol=2e-6)
        ex.close()


def test_modspec_complex_cli_matches_golden(tmp_path):
    from scipy.io import wavfile
    from speech_recognition_tools_b200.cli import computeModulationSpectrum as cli
    from speech_recognition_tools_b200.kaldi_io import read_ark
    c = load_case("modspec_complex_speech_pad")
    wav = tmp_path / "u.wav"
    wavfile.write(str(wav), 16000, c["x"])
    scp = tmp_path / "wav.scp"
    scp.write_text("utt1 %s\n" % wav)
    p = c["meta"]["params"]
    argv = [str(scp), str(tmp_path / "out"), "--complex_modulation", "--order", str(p["order"]), "--coeff_0", str(p["coeff_0"]),
            "--coeff_n", str(p["coeff_n"]), "--fduration", str(p["fduration"]), "--nfilters", str(p["nfilters"])]
    cli.getFeats(cli.get_args(argv))
    got = dict(read_ark(str(tmp_path / "out.ark")))["utt1"]
    assert got.shape == c["feats"].shape
    np.testing.assert_allclose(got, c["feats"], rtol=2e-6, atol=2e-6)
    with pytest.raises(ValueError):
        cli.getFeats(cli.get_args(argv + ["--keep_even"]))


def test_stage_frames_equals_getframes():
    """fdlp_stage_frames (the framing kernel's output as float64) == features.py:118-154 getFrames for hanning and
    square windows, int16 and float32 input, utterances shorter than a frame (multi-bounce reflection)."""
    import ctypes as C
    from oracle import fdlp as ofdlp
    from oracle import modspec as omod
    from speech_recognition_tools_b200 import _lib as L
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.fdlp import FdlpExtractor
    from speech_recognition_tools_b200.modspec import ModSpecConfig, ModSpecExtractor
    for no_window, fdur, cast in ((False, 0.5, None), (True, 0.1, np.float32)):
        ex = ModSpecExtractor(ModSpecConfig(no_window=no_window, fduration=fdur, order=20, coeff_n=10, coeff_0=1), device="cuda:0")
        wavs = [synth.make_utterance("white", 44, i, T) for i, T in enumerate([300, 1601, 9000])]
        if cast is not None:
            wavs = [w.astype(cast) for w in wavs]
        host, lengths = FdlpExtractor.pack(wavs)
        lay = ex.layout(lengths).to(ex.device)
        wav_dev = host.to(ex.device)
        b = L.FdlpBatch()
        d = lay.dev
        b.wav = wav_dev.data_ptr()
        b.wav_dtype = L.FDLP_WAV_INT16 if wav_dev.dtype == torch.int16 else L.FDLP_WAV_FLOAT32
        b.n_utt, b.n_seg, b.n_frames, b.max_utt_frames = lay.n_utt, lay.n_seg, lay.n_frames, lay.max_utt_frames
        b.utt_wav_off, b.utt_seg_off = d["utt_wav_off"].data_ptr(), d["utt_seg_off"].data_ptr()
        b.utt_frame_off, b.seg_utt, b.seg_ptr = d["utt_frame_off"].data_ptr(), d["seg_utt"].data_ptr(), d["seg_ptr"].data_ptr()
        b.noise_mode = L.FDLP_NOISE_NONE
        N = ex.geom["N"]
        frames = torch.zeros((lay.n_seg, N), dtype=torch.float64, device=ex.device)
        L.check(ex._lib.fdlp_stage_frames(ex._plan, C.byref(b), C.c_void_p(frames.data_ptr()),
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        got = frames.cpu().numpy()
        ref = np.concatenate([ofdlp.get_frames(w, 16000, 100, fdur, omod.sq_wind if no_window else np.hanning) for w in wavs])
        assert got.shape == ref.shape
        np.testing.assert_allclose(got, ref, rtol=1e-13, atol=1e-9)
        ex.close()
