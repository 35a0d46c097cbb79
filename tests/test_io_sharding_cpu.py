"""CPU tests: Kaldi ark/scp writer, utterance sharding, CLI argparse surface, and the N > 1 CMVN reduction
on world_size-2 gloo (the host-side logic of the multi-GPU path)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import cmvn as ocmvn
from speech_recognition_tools_b200 import kaldi_io, sharding


def test_ark_scp_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    feats = {"utt_a": rng.standard_normal((7, 5)).astype(np.float32), "utt-b": rng.standard_normal((1, 5)).astype(np.float32),
             "c": np.zeros((0, 5), dtype=np.float32)}
    ark, scp = str(tmp_path / "f.ark"), str(tmp_path / "f.scp")
    kaldi_io.write_ark_scp(feats.items(), ark, scp)
    back = kaldi_io.read_ark(ark)
    assert list(back) == list(feats)
    for k in feats:
        np.testing.assert_array_equal(back[k], feats[k])
    for line in open(scp):
        k, entry = line.split()
        np.testing.assert_array_equal(kaldi_io.read_scp_entry(entry), feats[k])
    # header bytes as Kaldi documents them
    raw = open(ark, "rb").read()
    assert raw.startswith(b"utt_a \0BFM \4\x07\0\0\0\4\x05\0\0\0")
    # the reference's '%.3f' text round trip
    kaldi_io.write_ark_scp(feats.items(), ark, scp, quantize_3dp=True)
    np.testing.assert_allclose(kaldi_io.read_ark(ark)["utt_a"], np.round(feats["utt_a"].astype(np.float64), 3), atol=1e-6)


def test_sharding_balances_segments():
    rng = np.random.default_rng(1)
    nseg = rng.integers(2, 15, size=1000)
    shards = sharding.assign_to_ranks(nseg, 8)
    assert sorted(u for s in shards for u in s) == list(range(1000))
    loads = [int(nseg[s].sum()) for s in shards]
    assert max(loads) - min(loads) <= 14
    batches = sharding.make_batches(shards[0], nseg, 64)
    assert sorted(u for b in batches for u in b) == shards[0]
    assert all(sum(int(nseg[u]) for u in b) <= 64 for b in batches)
    assert sharding.assign_to_ranks([3, 1], 4) == [[0], [1], [], []]


def test_cli_defaults_match_reference():
    from speech_recognition_tools_b200.cli import computeFDLPSpectrogram as cf
    from speech_recognition_tools_b200.cli import computeMelSpectrum as cm
    a = vars(cf.get_parser().parse_args(["in.scp", "out"]))
    # computeFDLPSpectrogram.py:241-262
    assert a == dict(scp="in.scp", outfile="out", scp_type="wav", nfilters=20, coeff_num=50, coeff_range="1,20", order=50,
                     fduration=0.5, frate=100, overlap_fraction=0.25, kaldi_cmd="copy-feats", add_reverb=None,
                     fbank_type="mel,1", odd_mod_zero=False, gamma_weight="None", lifter_config=None,
                     write_utt2num_frames=False, add_noise=None)
    # the wrapper's call, recipes/timit/local_pyspeech/make_FDLPspectrum_feats.sh:143-157
    w = cf.get_parser().parse_args(["s", "o", "--write_utt2num_frames", "--fbank_type=cochlear,1,1,1,2.5,1",
                                    "--gamma_weight=None", "--add_reverb=clean", "--add_noise=clean", "--coeff_num=100",
                                    "--coeff_range=0,100", "--order=150", "--overlap_fraction=0.25", "--nfilters=80",
                                    "--fduration=1.5", "--frate=100"])
    assert w.order == 150 and w.fduration == 1.5 and w.write_utt2num_frames
    m = vars(cm.get_args(["in.scp", "out"]))
    # computeMelSpectrum.py:20-37
    assert m == dict(scp="in.scp", outfile="out", scp_type="wav", spectrum_type="log", nfilters=23, fduration=0.02,
                     frate=100, nfft=1024, add_reverb=None, fbank_type="mel,1", write_utt2num_frames=False, add_noise=None)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _cmvn_worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from speech_recognition_tools_b200 import cmvn
    rng = np.random.default_rng(10)
    nseg = rng.integers(2, 15, size=40)
    mats = [np.random.default_rng(100 + u).standard_normal((int(nseg[u]) * 3, 6)).astype(np.float32) for u in range(40)]
    mine = sharding.assign_to_ranks(nseg, world)[rank]
    part = torch.from_numpy(ocmvn.cmvn_stats([mats[u] for u in mine])) if mine else torch.zeros((2, 7), dtype=torch.float64)
    tot = cmvn.all_reduce(part.clone())
    if rank == 0:
        np.save(os.path.join(tmp, "tot.npy"), tot.numpy())
        np.save(os.path.join(tmp, "ref.npy"), ocmvn.cmvn_stats(mats))
    dist.barrier()
    dist.destroy_process_group()


def test_cmvn_allreduce_gloo_world2(tmp_path):
    port = _free_port()
    mp.spawn(_cmvn_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    tot, ref = np.load(tmp_path / "tot.npy"), np.load(tmp_path / "ref.npy")
    np.testing.assert_allclose(tot, ref, rtol=1e-13)
    assert tot[0, -1] == ref[0, -1]
