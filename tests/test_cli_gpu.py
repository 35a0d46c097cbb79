"""GPU test of the drop-in CLI shims: scp of WAV files in, Kaldi binary ark/scp + .len out, values equal to
the reference's golden output under the same `random.seed` / `np.random.seed`."""
import argparse
import os
import random

import numpy as np
import pytest
from scipy.io import wavfile

from speech_recognition_tools_b200 import kaldi_io, synth
from tests.golden_util import load_case

pytestmark = pytest.mark.gpu


def test_fdlp_cli_matches_reference_golden(tmp_path):
    from speech_recognition_tools_b200.cli import computeFDLPSpectrogram as cf
    names = ["fdlp_cfg1_white_10s", "fdlp_cfg1_T18500"]
    cases = [load_case(n) for n in names]
    scp = tmp_path / "wav.scp"
    with open(scp, "w") as f:
        for n, c in zip(names, cases):
            wav = str(tmp_path / (n + ".wav"))
            wavfile.write(wav, 16000, c["x"])
            f.write("%s %s\n" % (n, wav))
        f.write("missing /nonexistent/file.wav\n")  # unreadable entries are skipped, like the reference
    out = str(tmp_path / "feats")
    args = cf.get_parser().parse_args([str(scp), out, "--write_utt2num_frames"])
    # one process, one RNG stream over the scp, as in the reference: utterance 1 is seeded 0 (its golden seed);
    # utterance 2 continues the stream, so only its shape / finiteness is checked here
    random.seed(cases[0]["meta"]["seed"])
    cf.getFeats(args)
    ark = kaldi_io.read_ark(out + ".ark")
    assert list(ark) == names
    np.testing.assert_allclose(ark[names[0]], cases[0]["feats"], rtol=2e-7, atol=5e-6)
    assert ark[names[1]].shape == cases[1]["feats"].shape and np.isfinite(ark[names[1]]).all()
    lens = dict(l.split() for l in open(out + ".len"))
    assert lens == {n: str(c["feats"].shape[0]) for n, c in zip(names, cases)}
    for line in open(out + ".scp"):
        k, entry = line.split()
        np.testing.assert_array_equal(kaldi_io.read_scp_entry(entry), ark[k])


def test_fdlp_cli_noise_matches_reference_golden(tmp_path, monkeypatch):
    from speech_recognition_tools_b200.cli import computeFDLPSpectrogram as cf
    c = load_case("fdlp_cfg1_noise_int16")
    os.makedirs(tmp_path / "noises")
    wavfile.write(str(tmp_path / "noises" / "pink.wav"), 16000, c["noise"]["data"])
    wavfile.write(str(tmp_path / "u.wav"), 16000, c["x"])
    with open(tmp_path / "wav.scp", "w") as f:
        f.write("utt1 %s\n" % (tmp_path / "u.wav"))
    monkeypatch.chdir(tmp_path)
    args = cf.get_parser().parse_args([str(tmp_path / "wav.scp"), str(tmp_path / "o"), "--add_noise", "pink,5"])
    random.seed(c["meta"]["seed"])
    np.random.seed(c["meta"]["seed"])
    feats = cf.getFeats(args, return_feats=True)
    np.testing.assert_allclose(feats["utt1"], c["feats"], rtol=2e-7, atol=5e-6)


def test_mel_cli_matches_reference_golden(tmp_path):
    from speech_recognition_tools_b200.cli import computeMelSpectrum as cm
    c = load_case("mel_recipe_speech")
    wavfile.write(str(tmp_path / "u.wav"), 16000, c["x"])
    with open(tmp_path / "wav.scp", "w") as f:
        f.write("utt1 %s\n" % (tmp_path / "u.wav"))
    args = cm.get_args([str(tmp_path / "wav.scp"), str(tmp_path / "m"), "--nfilters", "80", "--nfft", "2048",
                        "--write_utt2num_frames"])
    feats = cm.compute_mel_spectrum(args, return_feats=True)
    np.testing.assert_allclose(feats["utt1"], c["feats"], rtol=0, atol=1e-4)
    assert open(str(tmp_path / "m.len")).read() == "utt1 %d\n" % c["feats"].shape[0]


def test_cmvn_apply_kernel():
    import torch
    from oracle import cmvn as ocmvn
    from speech_recognition_tools_b200 import cmvn
    x = torch.randn(1000, 80, device="cuda:0") * 3 + 1
    stats = torch.from_numpy(ocmvn.cmvn_stats([x.cpu().numpy()])).cuda()
    ref = ocmvn.apply_cmvn(x.cpu().numpy(), stats.cpu().numpy())
    got = cmvn.apply(x.clone(), stats).cpu().numpy()
    np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-5)
