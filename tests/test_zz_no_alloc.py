"""cmvn_accumulate_noalloc (include/fdlp_b200.h): the allocation-free variant of the CMVN accumulation, for callers
that cannot call cudaMalloc.  No counterpart in the reference.

Collected last on purpose (a failure here must not hide any other test).  The checks go
through the C ABI and compare with the existing entry points' `cmvn_acc` output on two calls in a row that share the
static scratch array: mel_forward (same kernels, same summation order: bit-identical) and fdlp_forward (other
summation order: the 1e-12 relative tolerance of tests/test_fdlp_gpu.py)."""
import ctypes as C

import numpy as np
import pytest
import torch

from speech_recognition_tools_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def _noalloc(lib, feats, stats):
    assert feats.is_contiguous() and feats.dtype == torch.float32 and stats.dtype == torch.float64
    stream = torch.cuda.current_stream(feats.device).cuda_stream
    with torch.cuda.device(feats.device):
        return lib.cmvn_accumulate_noalloc(C.c_void_p(feats.data_ptr()), feats.shape[0], feats.shape[1],
                                           C.c_void_p(stats.data_ptr()), C.c_void_p(stream))


def test_noalloc_cmvn_matches_the_allocating_entry_points_twice_in_a_row():
    from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor
    from speech_recognition_tools_b200.mel import MelConfig, MelExtractor
    lib = _lib.load()
    dev = torch.device("cuda:0")
    lengths = synth.wsj_shaped_lengths(12, seed=9)
    wavs = [synth.make_utterance("white" if i % 2 else "speechlike", 61, i, int(T), 30.0) for i, T in enumerate(lengths)]

    # existing entry point 1: mel_forward with cmvn_acc (80 bands)
    mel = MelExtractor(MelConfig(nfilters=80, nfft=2048), device=dev)
    cm_mel = torch.zeros((2, 81), dtype=torch.float64, device=dev)
    m_outs = mel(wavs, cmvn=cm_mel)
    m_feats = torch.cat(m_outs).contiguous()
    # existing entry point 2: fdlp_forward with cmvn_acc (reference defaults: 20 bands, another width of the slots)
    ex = FdlpExtractor(FdlpConfig(), device=dev)
    cm_fdlp = torch.zeros((2, 21), dtype=torch.float64, device=dev)
    f_outs = ex(wavs[:5], jitter=list(range(5)), cmvn=cm_fdlp)
    f_feats = torch.cat(f_outs).contiguous()

    got_mel = torch.zeros((2, 81), dtype=torch.float64, device=dev)
    got_fdlp = torch.zeros((2, 21), dtype=torch.float64, device=dev)
    got_mel2 = torch.zeros((2, 81), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    alloc0 = torch.cuda.memory_allocated(dev)
    # two calls in a row on one stream (then a third: back to the first shape over the second call's leftovers)
    assert _noalloc(lib, m_feats, got_mel) == _lib.FDLP_OK
    assert _noalloc(lib, f_feats, got_fdlp) == _lib.FDLP_OK
    assert _noalloc(lib, m_feats, got_mel2) == _lib.FDLP_OK
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 >= free0 and torch.cuda.memory_allocated(dev) == alloc0, "the entry point must not allocate"

    assert torch.equal(got_mel, cm_mel)      # same kernels, same order: bit for bit
    assert torch.equal(got_mel2, cm_mel)     # and again after another call used the array with another slot width
    np.testing.assert_allclose(got_fdlp.cpu().numpy(), cm_fdlp.cpu().numpy(), rtol=1e-12, atol=1e-9)
    assert float(got_fdlp[0, 20]) == f_feats.shape[0] and float(got_mel[0, 80]) == m_feats.shape[0]
    # the stats argument is ADDED to, like cmvn_acc
    assert _noalloc(lib, m_feats, got_mel) == _lib.FDLP_OK
    torch.cuda.synchronize()
    np.testing.assert_allclose(got_mel.cpu().numpy(), 2 * cm_mel.cpu().numpy(), rtol=1e-15, atol=0)


def test_noalloc_cmvn_rejects_sizes_beyond_the_static_array():
    lib = _lib.load()
    dev = torch.device("cuda:0")
    wide = torch.zeros((4, _lib.FDLP_NOALLOC_MAX_DIM + 1), dtype=torch.float32, device=dev)
    stats = torch.zeros((2, _lib.FDLP_NOALLOC_MAX_DIM + 2), dtype=torch.float64, device=dev)
    assert _noalloc(lib, wide, stats) == _lib.FDLP_E_ARG
    assert b"FDLP_NOALLOC_MAX_DIM" in lib.fdlp_last_error()
    assert float(stats.abs().sum()) == 0.0
    # the largest supported width works and agrees with a float64 sum
    x = torch.randn((1000, _lib.FDLP_NOALLOC_MAX_DIM), dtype=torch.float32, device=dev,
                    generator=torch.Generator(device=dev).manual_seed(3))
    st = torch.zeros((2, _lib.FDLP_NOALLOC_MAX_DIM + 1), dtype=torch.float64, device=dev)
    assert _noalloc(lib, x, st) == _lib.FDLP_OK
    xd = x.double()
    np.testing.assert_allclose(st[0, :-1].cpu().numpy(), xd.sum(0).cpu().numpy(), rtol=1e-12, atol=1e-9)
    np.testing.assert_allclose(st[1, :-1].cpu().numpy(), (xd * xd).sum(0).cpu().numpy(), rtol=1e-12, atol=1e-9)
    assert float(st[0, -1]) == 1000.0 and float(st[1, -1]) == 0.0
