"""GPU parity tests of the MFCC path (csrc/mfcc.cuh) against golden outputs of the unmodified reference
(tests/golden/mfcc_*.npz, made by tests/golden/make_golden_mfcc.py) and against oracle/mfcc.py.
Tolerance: 2e-4 absolute on the cepstral coefficients (c0 sums 2 x 30 log10 band energies of a float32 DFT)."""
import json
import os

import numpy as np
import pytest

from tests.golden_util import GOLDEN_DIR, golden_cases

pytestmark = pytest.mark.gpu
ATOL = 2e-4


def load_mfcc_case(name):
    from speech_recognition_tools_b200 import synth
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    sp = meta["input"]
    x = synth.make_utterance(sp["kind"], sp["seed"], sp["utt"], sp["T"], sp.get("sigma_f", 30.0))
    return meta, x, z["feats"]


@pytest.mark.parametrize("name", golden_cases("mfcc"))
def test_mfcc_matches_reference_golden(name):
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.fdlp import NoiseSpec
    from speech_recognition_tools_b200.mfcc import MfccConfig, MfccExtractor
    meta, x, ref = load_mfcc_case(name)
    ex = MfccExtractor(MfccConfig(**meta["params"]), device="cuda:0")
    noise = None
    if meta["noise"]:
        nz = meta["noise"]
        noise = NoiseSpec(synth.noise_int16(nz["type"], nz["T"], seed=nz["seed"]), nz["snr"], rand=[nz["rand"]])
    rir = None
    if meta.get("room"):
        rir = synth.rir_int16(meta["room"], meta["R"], meta["rir_seed"])[:, 1] / 2.0 ** 15
    got = ex([x], noise=noise, reverb=rir)[0].cpu().numpy().astype(np.float64)
    assert got.shape == ref.shape
    np.testing.assert_allclose(got, ref, rtol=0, atol=ATOL)
    ex.close()


def test_mfcc_batch_equals_single_and_oracle():
    import torch
    from oracle import mfcc as omfcc
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.mfcc import MfccConfig, MfccExtractor
    cfg = dict(nfilters=30, nfft=1024, fduration=0.02, frate=100, context=3)
    ex = MfccExtractor(MfccConfig(**cfg), device="cuda:0")
    assert ex([]) == []
    wavs = [synth.make_utterance("speechlike" if i % 2 else "white", 6, i, T) for i, T in
            enumerate((2, 161, 8000, 40001, 322, 5121))]
    wavs.append(synth.make_utterance("white", 6, 9, 3000).astype(np.float32))  # IEEE-float WAV: separate batch
    outs = ex(wavs[:-1]) + ex(wavs[-1:])
    for w, o in zip(wavs, outs):
        ref = omfcc.mfcc_features(w, **cfg)
        assert o.shape == ref.shape
        np.testing.assert_allclose(o.cpu().numpy(), ref, rtol=0, atol=ATOL)
        assert torch.equal(ex([w])[0], o)
    ex.close()


def test_mfcc_cli_shim_writes_kaldi_ark(tmp_path):
    from scipy.io import wavfile
    from oracle import mfcc as omfcc
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.cli import computeMfccFeatures as cli
    from speech_recognition_tools_b200.kaldi_io import read_ark
    scp = tmp_path / "wav.scp"
    lines, wavs = [], {}
    for i, T in enumerate((16000, 9001)):
        x = synth.make_utterance("speechlike", 8, i, T)
        wavfile.write(str(tmp_path / ("u%d.wav" % i)), 16000, x)
        lines.append("utt%d %s" % (i, tmp_path / ("u%d.wav" % i)))
        wavs["utt%d" % i] = x
    scp.write_text("\n".join(lines) + "\n")
    out = str(tmp_path / "mfcc")
    feats = cli.extractMelEnergyFeats(cli.get_args([str(scp), out, "--nfilters", "26"]), return_feats=True)
    ark = read_ark(out + ".ark")
    assert set(ark) == set(wavs) and os.path.exists(out + ".scp")
    for k, x in wavs.items():
        ref = omfcc.mfcc_features(x, nfilters=26)
        np.testing.assert_allclose(feats[k], ref, rtol=0, atol=ATOL)
        np.testing.assert_allclose(ark[k], feats[k], rtol=0, atol=1e-6)


def test_noise_then_reverb_against_oracle():
    """computeMfccFeatures.py:113-120: signal / 2^15, noise mix, then addReverb."""
    from oracle import mfcc as omfcc
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.fdlp import NoiseSpec
    from speech_recognition_tools_b200.mfcc import MfccConfig, MfccExtractor
    x = synth.make_utterance("speechlike", 46, 0, 24000, 30.0)
    nz = synth.noise_int16("factory1", 70000, seed=5)
    rir = synth.rir_int16("small_room", 2500)[:, 1] / 2.0 ** 15
    ex = MfccExtractor(MfccConfig(), device="cuda:0")
    got = ex([x], noise=NoiseSpec(nz, 12.0, rand=[0.37]), reverb=rir)[0].cpu().numpy().astype(np.float64)
    ref = omfcc.mfcc_features(x, noise=nz, snr=12.0, rand=0.37, reverb=rir)
    assert got.shape == ref.shape
    np.testing.assert_allclose(got, ref, rtol=0, atol=ATOL)
    ex.close()


def test_mfcc_tensor_core_kernel_agrees_with_fp32_pipe_kernel():
    """The default configuration runs the 513-point transform as a 3xTF32 GEMM on the tensor cores (csrc/mfcc_tc.cuh);
    the FP32-pipe kernel stays for other shapes.  Both against the oracle on a ragged batch (int16 and float32 input,
    noise mixing, an utterance shorter than a frame, a frame count that is not a multiple of the 128-frame tile), and
    against each other; a configuration the GEMM kernel does not cover reports so."""
    import torch
    from oracle import mfcc as omfcc
    from speech_recognition_tools_b200 import synth
    from speech_recognition_tools_b200.fdlp import NoiseSpec
    from speech_recognition_tools_b200.mfcc import MfccConfig, MfccExtractor
    ex = MfccExtractor(MfccConfig(), device="cuda:0")
    assert ex.set_tensor_cores(True) is True
    lens = [200, 16000, 48001, 7777, 130000]
    wavs = [synth.make_utterance("speechlike" if i % 2 else "white", 91, i, T, 35.0) for i, T in enumerate(lens)]
    for cast in (lambda w: w, lambda w: w.astype(np.float32)):
        ws = [cast(w) for w in wavs]
        tc = [o.cpu().numpy() for o in ex(ws)]
        assert ex.set_tensor_cores(False) is False
        ff = [o.cpu().numpy() for o in ex(ws)]
        assert ex.set_tensor_cores(True) is True
        for w, a, b in zip(ws, tc, ff):
            ref = omfcc.mfcc_features(w)
            assert a.shape == ref.shape == b.shape
            np.testing.assert_allclose(a, ref, rtol=0, atol=ATOL)
            np.testing.assert_allclose(b, ref, rtol=0, atol=ATOL)
            np.testing.assert_allclose(a, b, rtol=0, atol=ATOL)
    nz = synth.noise_int16("babble", 140000, seed=4)
    noise = NoiseSpec(nz, 12.0, rand=[0.1 * i for i in range(len(wavs))])
    tc = [o.cpu().numpy() for o in ex(wavs, noise=noise)]
    ex.set_tensor_cores(False)
    ff = [o.cpu().numpy() for o in ex(wavs, noise=noise)]
    for a, b in zip(tc, ff):
        np.testing.assert_allclose(a, b, rtol=0, atol=ATOL)
    ex.close()
    other = MfccExtractor(MfccConfig(nfft=512, nfilters=23), device="cuda:0")
    assert other.set_tensor_cores(True) is False
    other.close()
    torch.cuda.synchronize()
