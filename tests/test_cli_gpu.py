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


def _write_scp(tmp_path, wavs, extra_lines=()):
    scp = tmp_path / "wav.scp"
    with open(scp, "w") as f:
        for k, x in wavs.items():
            wavfile.write(str(tmp_path / (k + ".wav")), 16000, x)
            f.write("%s %s\n" % (k, tmp_path / (k + ".wav")))
        for line in extra_lines:
            f.write(line + "\n")
    return str(scp)


def test_fdlp_cli_skips_bad_entries_batches_and_pipes(tmp_path, monkeypatch):
    """a corrupt / empty WAV (scipy raises ValueError) or a failing pipe must be skipped, not abort the
    shard; batches are written incrementally; a `cat file |` pipe entry and a float32 WAV go through scipy's reader /
    a batch of their own.  Results equal the extractor called directly with the same per-utterance jitter seeds."""
    from speech_recognition_tools_b200.cli import computeFDLPSpectrogram as cf
    from speech_recognition_tools_b200.cli import pipeline as pl
    from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor
    wavs = {"u%02d" % i: synth.make_utterance("white" if i % 2 else "speechlike", 71, i, 9000 + 3777 * i, 30.0)
            for i in range(9)}
    wavs["f32"] = synth.make_utterance("white", 72, 0, 21000).astype(np.float32)
    open(tmp_path / "empty.wav", "wb").close()
    with open(tmp_path / "trunc.wav", "wb") as f:
        f.write(b"RIFF\x10\x00\x00\x00WAVEfmt ")
    scp = _write_scp(tmp_path, wavs, ["bad1 %s" % (tmp_path / "empty.wav"), "bad2 %s" % (tmp_path / "trunc.wav"),
                                      "bad3 false |", "piped cat %s |" % (tmp_path / "u03.wav")])
    monkeypatch.setenv("FDLP_JITTER_SEED", "5")
    monkeypatch.setattr(cf, "MAX_SEGMENTS_PER_BATCH", 12)  # several batches
    out = str(tmp_path / "o")
    args = cf.get_parser().parse_args([scp, out, "--write_utt2num_frames"])
    kept = cf.getFeats(args, return_feats=True)
    ark = kaldi_io.read_ark(out + ".ark")
    want = list(wavs) + ["piped"]
    assert list(ark) == want == list(kept)
    ex = FdlpExtractor(FdlpConfig())
    for k in want:
        x = wavs["u03"] if k == "piped" else wavs[k]
        ref = ex([x], jitter=[pl.utt_seed(5, k)])[0].cpu().numpy()
        np.testing.assert_array_equal(ark[k], ref)
        np.testing.assert_array_equal(kept[k], ref)
    lens = dict(l.split() for l in open(out + ".len"))
    assert lens == {k: str(ark[k].shape[0]) for k in want}
    for line in open(out + ".scp"):
        k, entry = line.split()
        np.testing.assert_array_equal(kaldi_io.read_scp_entry(entry), ark[k])


def test_fdlp_cli_two_ranks_equal_one_process(tmp_path):
    """torchrun mode: two ranks shard the scp, write `<out>.<JOB>.ark` and a merged scp / len + one all-reduced
    cmvn.ark; with FDLP_JITTER_SEED the features equal the single-process run bit for bit and the CMVN sums agree to
    1e-12 (another summation order)."""
    import subprocess
    import sys

    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cli = os.path.join(root, "speech_recognition_tools_b200", "cli", "computeFDLPSpectrogram.py")
    lengths = synth.wsj_shaped_lengths(24, seed=3)
    wavs = {"utt%02d" % i: synth.make_utterance("white", 73, i, int(T)) for i, T in enumerate(lengths)}
    scp = _write_scp(tmp_path, wavs)
    flags = ["--write_utt2num_frames", "--fbank_type=cochlear,1,1,1,2.5,1", "--coeff_num=100", "--coeff_range=0,100",
             "--order=150", "--nfilters=80", "--fduration=1.5"]
    env = dict(os.environ, FDLP_JITTER_SEED="9", FDLP_WRITE_CMVN="1")
    one, two = str(tmp_path / "one"), str(tmp_path / "two")
    subprocess.run([sys.executable, cli, scp, one] + flags, check=True, env=env, stdout=subprocess.DEVNULL)
    subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                    "--master-addr", "127.0.0.1", "--master-port", "29671", cli, scp, two] + flags,
                   check=True, env=env, stdout=subprocess.DEVNULL)
    a = kaldi_io.read_ark(one + ".ark")
    b = {}
    for line in open(two + ".scp"):
        k, entry = line.split()
        b[k] = kaldi_io.read_scp_entry(entry)
    assert list(b) == list(a) == list(wavs)
    for k in a:
        np.testing.assert_array_equal(a[k], b[k])
    assert open(one + ".len").read() == open(two + ".len").read()
    shards = [kaldi_io.read_ark("%s.%d.ark" % (two, r)) for r in (1, 2)]
    assert all(len(s) > 0 for s in shards) and len(shards[0]) + len(shards[1]) == len(a)

    def stats(path):
        raw = open(path, "rb").read()
        return np.frombuffer(raw[raw.index(b"BDM ") + 4 + 10:], dtype="<f8").reshape(2, 81)
    s1, s2 = stats(one + ".cmvn.ark"), stats(two + ".cmvn.ark")
    np.testing.assert_allclose(s2, s1, rtol=1e-12, atol=1e-9)
    assert s2[0, 80] == sum(v.shape[0] for v in a.values())


def test_mel_cli_add_noise_matches_reference_golden(tmp_path, monkeypatch):
    """computeMelSpectrum.py --add_noise type,snr (:69-72,134-141) through the shim, same np.random.seed as the reference."""
    from speech_recognition_tools_b200.cli import computeMelSpectrum as cm
    c = load_case("mel_noise_babble10")
    os.makedirs(tmp_path / "noises")
    wavfile.write(str(tmp_path / "noises" / "babble.wav"), 16000, c["noise"]["data"])
    wavfile.write(str(tmp_path / "u.wav"), 16000, c["x"])
    with open(tmp_path / "wav.scp", "w") as f:
        f.write("utt1 %s\n" % (tmp_path / "u.wav"))
    monkeypatch.chdir(tmp_path)
    args = cm.get_args([str(tmp_path / "wav.scp"), str(tmp_path / "m"), "--nfilters", "80", "--nfft", "2048",
                        "--add_noise", "babble,10"])
    np.random.seed(c["meta"]["seed"])
    feats = cm.compute_mel_spectrum(args, return_feats=True)
    np.testing.assert_allclose(feats["utt1"], c["feats"], rtol=0, atol=1e-4)
    c2 = load_case("mel_noise_diff")
    wavfile.write(str(tmp_path / "d.wav"), 16000, c2["x"])
    with open(tmp_path / "d.scp", "w") as f:
        f.write("utt1 %s\n" % (tmp_path / "d.wav"))
    feats = cm.compute_mel_spectrum(cm.get_args([str(tmp_path / "d.scp"), str(tmp_path / "d"), "--add_noise", "diff"]),
                                    return_feats=True)
    np.testing.assert_allclose(feats["utt1"], c2["feats"], rtol=0, atol=1e-4)
