#!/usr/bin/env python
"""Drop-in for /root/reference/src/featgen/computeFDLPSpectrogram.py: same positional arguments, flags,
defaults (:241-262) and outputs (<outfile>.ark/.scp Kaldi binary float matrices, <outfile>.len), computed
by libfdlp_b200.so on the GPU.  `getFeats(args)` keeps the reference's signature.

Randomness: the overlap-add jitter is drawn from the module-level `random` generator and the noise offset
from `np.random`, in scp order, exactly like the reference -- seed them the same way to reproduce it."""
import argparse
import os
import random
import sys
import time

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

from speech_recognition_tools_b200 import _lib  # noqa: E402
from speech_recognition_tools_b200.cli.common import load_noise, write_lens  # noqa: E402
from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor, NoiseSpec, segment_count  # noqa: E402
from speech_recognition_tools_b200.reverb import Reverberator, load_rir  # noqa: E402

MAX_SEGMENTS_PER_BATCH = int(os.environ.get("FDLP_CLI_MAX_SEGMENTS", "2048"))  # several batches per shard: the three pipeline stages overlap


_PLANS = {}


class _GlobalRandom:
    """`random.randrange(2)` of computeFDLPSpectrogram.py:225 on the module-level generator."""

    mt_instance = random._inst  # the generator behind random.randrange / random.seed (bulk jitter draws, fdlp.py)

    @staticmethod
    def randrange(n):
        return random.randrange(n)


def getFeats(args, srate=16000, window=np.hamming, return_feats=False):
    """The reference's entry point (computeFDLPSpectrogram.py:29-237) on the GPU.  Beyond the reference:
    * launched under torchrun (RANK / WORLD_SIZE / LOCAL_RANK) every rank takes a balanced share of the scp, writes
      `<outfile>.<JOB>.ark/.scp/.len` (JOB = rank + 1, the wrapper's naming) and rank 0 writes the merged
      `<outfile>.scp/.len` plus the all-reduced `<outfile>.cmvn.ark` (compute-cmvn-stats, e2e/wsj/run_fdlp_e1.sh:280);
    * FDLP_JITTER_SEED=<int> draws the overlap-add jitter per utterance from Random(crc32(uttid) ^ f(seed)) instead of
      the process-wide stream, so that results do not depend on batching or on the rank that owns the utterance;
    * FDLP_WRITE_CMVN=1 writes `<outfile>.cmvn.ark` in a single-process run too."""
    import torch

    from speech_recognition_tools_b200 import cmvn as cmvn_mod
    from speech_recognition_tools_b200.cli import pipeline as pl
    pl.trace("start")
    from speech_recognition_tools_b200.fdlp import batch_layout
    from speech_recognition_tools_b200.kaldi_io import write_cmvn_ark
    if window is not np.hamming:
        raise NotImplementedError("only the Hamming window of the reference recipes is implemented")
    cfg = FdlpConfig.from_args(args)
    add_noise, add_reverb = args.add_noise, args.add_reverb
    # computeFDLPSpectrogram.py:75-91: ./RIR/RIR_<room>_AnglA.wav, second channel, / 2^15; ValueError for an unknown
    # room, the file-not-found error of scipy's reader when the recordings are not in the working directory
    rir = load_rir(add_reverb)
    if add_reverb == "clean":
        print("%s: No reverberation added!" % sys.argv[0])
    reverberator = None
    noise = None
    if add_noise and add_noise not in ("clean",):
        if add_noise == "diff":
            noise = "diff"  # computeFDLPSpectrogram.py:162-164: 13-tap FIR instead of a noise recording
        else:
            noise_info = add_noise.strip().split(",")
            noise = (load_noise(noise_info[0]), float(noise_info[1]))
    if args.odd_mod_zero:
        print("%s: Ignoring odd modulations... " % sys.argv[0])

    rank, world, local_rank = pl.dist_env()
    device = None
    if world > 1:
        import torch.distributed as dist
        device = torch.device("cuda:%d" % local_rank)
        torch.cuda.set_device(device)
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=device)
    # one plan per (configuration, device) and process: a second getFeats call in the same process (tests, bench.py's
    # e2e_cli, a wrapper that loops over shards) reuses filterbank, tables and scratch
    key = (repr(cfg), str(device))
    ex = _PLANS.get(key)
    if ex is None:
        ex = _PLANS[key] = FdlpExtractor(cfg, device=device)
    g, J = ex.geom, cfg.nfilters

    def seg_count(T):
        return segment_count(T, g["N"], g["S"])

    entries = pl.parse_scp(args.scp)
    pl.trace("scp_parsed")
    mine = pl.shard_entries(entries, world, rank, seg_count, args.scp_type) if world > 1 else entries
    outbase = args.outfile if world == 1 else "%s.%d" % (args.outfile, rank + 1)
    seed_env = os.environ.get("FDLP_JITTER_SEED")
    acc = cmvn_mod.new_accumulator(J, ex.device)
    overlapped = rir is None and noise != "diff"  # samples go to the kernels as they are: copies overlap the compute

    def n_frames(T):  # computeFDLPSpectrogram.py:182; reverberation may change the length by a sample: no reservation then
        return -(-T * cfg.frate // cfg.srate)

    with pl.HostPipeline(mine, args.scp_type, srate, seg_count, MAX_SEGMENTS_PER_BATCH, outbase + ".ark",
                         outbase + ".scp", singular_bit=_lib.FDLP_ST_SINGULAR, keep=return_feats,
                         frames_fn=n_frames if overlapped else None, dim=J) as pipe:
        pl.trace("pipe_open")
        for hb in pipe:
            pl.trace("m_got")
            n_utt = len(hb.ids)
            jitter = _GlobalRandom() if seed_env is None else [pl.utt_seed(int(seed_env), k) for k in hb.ids]
            nspec = None
            if noise == "diff":
                nspec = "diff"
            elif noise is not None:
                draws = [np.random.rand() for _ in range(n_utt)]  # features.py:25, one draw per utterance in scp order
                nspec = NoiseSpec(samples=noise[0], snr_db=noise[1], rand=draws, mode="int16_wrap")
            st_host = pipe.status_buffer(hb, n_utt)
            if overlapped:
                lay = batch_layout(hb.lengths, g, cfg.frate, cfg.srate, jitter)
                pl.trace("m_layout")
                nd = ex._noise_to_device(nspec, lay) if nspec is not None else None
                feats_host = pipe.feats_buffer(hb, lay.n_frames, J)
                status = torch.zeros(n_utt, dtype=torch.int32, device=ex.device)
                ex.extract_host(hb.wav, lay, feats_host=feats_host, cmvn=acc, status=status, noise_dev=nd,
                                n_chunks=max(1, min(6, lay.n_seg // 256)))
                frame_off = lay.utt_frame_off
            else:
                if rir is not None and reverberator is None:
                    reverberator = Reverberator(rir, device=ex.device)
                off = np.concatenate([[0], np.cumsum(hb.lengths)])
                wavs = [hb.wav[int(off[u]):int(off[u + 1])] for u in range(n_utt)]
                outs, status = ex(wavs, jitter=jitter, noise=nspec, return_status=True, reverb=reverberator, cmvn=acc)
                frame_off = np.concatenate([[0], np.cumsum([o.shape[0] for o in outs])])  # reverberation may drop a sample
                feats_host = pipe.feats_buffer(hb, int(frame_off[-1]), J)
                feats_host.copy_(torch.cat(outs), non_blocking=True)  # one device-to-host copy per batch
            st_host.copy_(status, non_blocking=True)
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(ex.device))
            pipe.submit(hb, done, feats_host, frame_off, st_host)
            pl.trace("m_submitted")
    all_lens, bad, kept = pipe.lens, pipe.bad, pipe.kept
    pl.trace("pipe_closed")
    if os.environ.get("FDLP_CLI_TRACE"):
        print("%s: stage seconds %s" % (sys.argv[0], {k: round(v, 4) for k, v in pipe.timings().items()}), file=sys.stderr)
    if args.write_utt2num_frames:
        write_lens(outbase, {k: all_lens[k] for k, _ in mine if k in all_lens})
    if world > 1:
        import torch.distributed as dist
        cmvn_mod.all_reduce(acc)
        dist.barrier()
        if rank == 0:
            pl.merge_rank_outputs(args.outfile, entries, world, args.write_utt2num_frames)
    if (world > 1 and rank == 0) or (world == 1 and os.environ.get("FDLP_WRITE_CMVN")):
        # Kaldi's documented stats layout; parity unpinned -- Kaldi is not on this box
        write_cmvn_ark(acc.cpu().numpy(), args.outfile + ".cmvn.ark")
    pl.trace("done")
    if os.environ.get("FDLP_CLI_TRACE"):
        pl.dump_trace(sys.argv[0])
    if bad:
        # the reference raises numpy.linalg.LinAlgError('Singular principal minor') at the first such utterance; here
        # the shard is complete on disk (the singular bands are NaN) and the error names the utterances
        raise np.linalg.LinAlgError("Singular principal minor in utterance(s): %s" % " ".join(bad))
    if return_feats:
        return kept


def get_parser():
    parser = argparse.ArgumentParser("Extract FDLP Spectrogram.")
    parser.add_argument("scp", help='"scp" list')
    parser.add_argument("outfile", help="output file")
    parser.add_argument("--scp_type", default="wav", help="scp type can be 'wav' or 'segment'")
    parser.add_argument("--nfilters", type=int, default=20, help="number of filters (15)")
    parser.add_argument("--coeff_num", type=int, default=50, help="Total Number of coefficients to compute")
    parser.add_argument("--coeff_range", type=str, default="1,20", help="Range of Modulation coefficients to keep")
    parser.add_argument("--order", type=int, default=50, help="LPC filter order (50)")
    parser.add_argument("--fduration", type=float, default=0.5, help="Window length (0.5 sec)")
    parser.add_argument("--frate", type=int, default=100, help="Frame rate (100 Hz)")
    parser.add_argument("--overlap_fraction", type=float, default=0.25, help="Fraction of Overlap for OLA")
    parser.add_argument("--kaldi_cmd", default="copy-feats", help="accepted for compatibility; the ark is written natively")
    parser.add_argument("--add_reverb", help='input "clean" OR "small_room" OR "large_room"')
    parser.add_argument("--fbank_type", type=str, default="mel,1",
                        help="mel,warp_fact OR cochlear,om_w,alpa,fixed,beta,warp_fact")
    parser.add_argument("--odd_mod_zero", action="store_true", help="Ignore the odd modulation coefficients")
    parser.add_argument("--gamma_weight", type=str, default="None", help="Configured as scale,shape,pk")
    parser.add_argument("--lifter_config", type=str, default=None, help="Configuration for general liftering")
    parser.add_argument("--write_utt2num_frames", action="store_true", help="Set to write utt2num_frames")
    parser.add_argument("--add_noise", help='Specify "type of noise, snr"')
    return parser


if __name__ == "__main__":
    args = get_parser().parse_args()
    start_time = time.time()
    print("%s: Extracting features...." % sys.argv[0])
    sys.stdout.flush()
    getFeats(args)
    print("Execution Time: {t:.3f} seconds".format(t=time.time() - start_time))
    sys.stdout.flush()
