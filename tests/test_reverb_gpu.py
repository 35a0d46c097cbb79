"""GPU parity tests of the --add_reverb kernels (csrc/reverb.cuh) through the C ABI (fdlp_reverb_*).

The alignment index indM is an index decision and must be exact.  The reverberated waveform is float64: for int16
input and an int16 / 2^15 impulse response the reference's result is exact, the CUDA path's error is that of its
float64 block FFTs, held here to 1e-11 of the largest output sample (the FDLP features need ~1e-10, NOTES_r01.md)."""
import numpy as np
import pytest
import torch

from oracle import fdlp as ofdlp
from oracle import reverb as oreverb
from tests.golden_util import golden_cases, load_case
from speech_recognition_tools_b200 import synth

pytestmark = pytest.mark.gpu

REL = 1e-11


def reverberator(rir, block=0):
    from speech_recognition_tools_b200.reverb import Reverberator
    return Reverberator(rir, device="cuda:0", block=block)


def assert_wave_close(got, ref, rel=REL):
    got = np.asarray(got, dtype=np.float64)
    assert got.shape == ref.shape
    np.testing.assert_allclose(got, ref, rtol=0, atol=rel * max(1.0, float(np.abs(ref).max(initial=0.0))))


@pytest.mark.parametrize("name", golden_cases("reverb"))
@pytest.mark.parametrize("block", [0, 64, 512])
def test_reverb_matches_reference_golden(name, block):
    c = load_case(name)
    rv = reverberator(c["rir"], block)
    wav = torch.from_numpy(c["x"]).to("cuda:0")
    out, ind, out_len = rv.forward_packed(wav, [len(c["x"])])
    ref = c["z"]["out"]
    assert int(ind.cpu()[0]) == c["meta"]["ind"]
    assert int(out_len.cpu()[0]) == len(ref)
    assert_wave_close(out.cpu().numpy()[:len(ref)], ref)
    assert np.all(out.cpu().numpy()[len(ref):] == 0.0)


@pytest.mark.parametrize("block,R", [(64, 333), (256, 256), (1024, 2500), (4096, 9000)])
def test_ragged_batch_against_oracle(block, R):
    rng = np.random.default_rng(block + R)
    lengths = [2, 63, 64, 65, int(block), int(block) + 1, 3 * int(block), 7001, 12345, R - 1, R, R + 1]
    wavs = [synth.make_utterance("white" if i % 2 else "speechlike", 40, i, int(T)) for i, T in enumerate(lengths)]
    rir = synth.rir_int16("medium_room", R, seed=int(rng.integers(1, 99)))[:, 1] / 2.0 ** 15
    rv = reverberator(rir, block)
    assert rv.block == block and rv.partitions == -(-R // block)
    outs = rv(wavs)
    for x, o in zip(wavs, outs):
        ref, ind = oreverb.add_reverb(x, rir, return_index=True)
        assert o.numel() == len(ref)
        assert_wave_close(o.cpu().numpy(), ref)


def test_float32_input_output_and_noise_mix_before_reverb():
    from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor, NoiseSpec
    T, R = 20000, 1800
    x = synth.make_utterance("speechlike", 41, 0, T, 30.0)
    rir = synth.rir_int16("small_room", R)[:, 1] / 2.0 ** 15
    rv = reverberator(rir)
    # float32 samples in, float32 out (mel / MFCC paths)
    xf = (x.astype(np.float32) * np.float32(0.37))
    out32, ind, _ = rv.forward_packed(torch.from_numpy(xf).to("cuda:0"), [T], out_dtype=torch.float32)
    ref, rind = oreverb.add_reverb(xf.astype(np.float64), rir, return_index=True)
    assert int(ind.cpu()[0]) == rind
    np.testing.assert_allclose(out32.cpu().numpy(), ref.astype(np.float32), rtol=2e-7, atol=1e-7 * np.abs(ref).max())
    # noise first, then reverberation (computeFDLPSpectrogram.py:158-170), through the extractor
    nz = synth.noise_int16("babble", 60000)
    ex = FdlpExtractor(FdlpConfig(), device="cuda:0")
    lay = ex.layout([T], 0)
    wav = torch.from_numpy(x).to("cuda:0")
    nd = ex._noise_to_device(NoiseSpec(samples=nz, snr_db=7.0, rand=3), lay)
    out, n = ex.reverberate(wav, lay, rv, nd)
    rand = np.random.RandomState(3).rand()
    noisy = ofdlp.add_noise_to_wav(x, nz, 7.0, rand=rand)
    assert_wave_close(out.cpu().numpy(), oreverb.add_reverb(noisy, rir))


def test_full_size_batch_delta_response_is_a_shift():
    """Size-independent property at the bench's batch shape (200 utterances of 2-15 s, a 1 s impulse response):
    a unit impulse at delay k gives full = sig shifted by k, the alignment finds d* = k, and the reference's
    indM = d* + 1 returns sig advanced by one sample."""
    lengths = synth.wsj_shaped_lengths(200, seed=77)
    wav_dev, off = synth.white_int16_device(5, np.arange(200), lengths, device="cuda:0")
    R, k = 16000, 100
    rir = np.zeros(R)
    rir[k] = 1.0
    rv = reverberator(rir)
    out, ind, out_len = rv.forward_packed(wav_dev, lengths, off)
    assert np.all(ind.cpu().numpy() == k + 1)
    assert np.array_equal(out_len.cpu().numpy(), lengths)
    sig = wav_dev.double()
    o = off.cpu().numpy()
    worst = 0.0
    for u in (0, 1, 57, 199):
        a, b = int(o[u]), int(o[u + 1])
        worst = max(worst, float((out[a:b - 1] - sig[a + 1:b]).abs().max()), abs(float(out[b - 1])))
    assert worst < 1e-8


def test_cli_add_reverb_reads_rir_directory(tmp_path):
    """The drop-in script with --add_reverb: ./RIR/<recording>.wav relative to the working directory, channel 1 / 2^15."""
    import os
    from scipy.io import wavfile
    from speech_recognition_tools_b200.reverb import RIR_FILES, load_rir
    os.makedirs(tmp_path / "RIR")
    rec = synth.rir_int16("large_room", 3000)
    wavfile.write(str(tmp_path / RIR_FILES["large_room"]), 16000, rec)
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        rir = load_rir("large_room")
    finally:
        os.chdir(cwd)
    np.testing.assert_array_equal(rir, rec[:, 1] / 2.0 ** 15)
    assert load_rir("clean") is None and load_rir(None) is None
    with pytest.raises(ValueError):
        load_rir("cathedral")


def test_one_plan_serves_batches_of_changing_shape():
    """The plan's scratch grows on demand (pairs and utterances independently): a long utterance, many short ones,
    the long one again -- every result equal to the oracle's."""
    rir = synth.rir_int16("small_room", 600)[:, 1] / 2.0 ** 15
    rv = reverberator(rir, 128)
    long_wav = [synth.make_utterance("white", 50, 0, 30000)]
    short = [synth.make_utterance("white", 50, 1 + i, 200 + 7 * i) for i in range(60)]
    for wavs in (long_wav, short, long_wav, short[:3]):
        outs = rv(wavs)
        for x, o in zip(wavs, outs):
            assert_wave_close(o.cpu().numpy(), oreverb.add_reverb(x, rir))
