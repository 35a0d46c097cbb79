"""CPU tests of the host-side logic and of the C-ABI library (no GPU needed): the library loads and exports
every symbol include/fdlp_b200.h declares; the Python bookkeeping (segment / frame counts, overlap-add
pointer trace with jitter, folded weights) agrees with the oracle."""
import os
import random
import re

import numpy as np
import pytest

from oracle import fdlp as ofdlp
from speech_recognition_tools_b200 import _lib, synth
from speech_recognition_tools_b200.fdlp import (FdlpConfig, batch_layout, build_weights, frame_count, ola_pointers,
                                                parse_fbank_type, segment_count)
from speech_recognition_tools_b200.mel import mel_frame_count

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "fdlp_b200.h")).read()
    declared = set(re.findall(r"\b((?:fdlp|mel|mfcc|cmvn)_[a-z0-9_]+)\s*\(", header))
    declared -= {"fdlp_plan", "mel_plan", "mfcc_plan"}
    assert declared, "no declarations parsed"
    bound = {name for name, _, _ in _lib.SYMBOLS}
    assert declared == bound, (declared ^ bound)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.fdlp_version() == _lib.FDLP_ABI_VERSION == int(re.search(r"#define FDLP_ABI_VERSION (\d+)", header).group(1))


HEADER_STRUCTS = {"fdlp_params": "FdlpParams", "fdlp_geometry": "FdlpGeometry", "fdlp_batch": "FdlpBatch",
                  "mel_params": "MelParams", "mel_batch": "MelBatch", "fdlp_reverb_batch": "ReverbBatch",
                  "mfcc_params": "MfccParams", "mfcc_batch": "MfccBatch"}
C_TO_CTYPES = {"uint32_t": "c_uint32", "int32_t": "c_int32", "int64_t": "c_int64", "double": "c_double"}


def _header_fields(header, cname):
    """[(field, ctypes type name)] of `typedef struct { ... } cname;` in include/fdlp_b200.h."""
    body = re.search(r"typedef struct \{((?:(?!typedef struct).)*?)\}\s*%s;" % cname, header, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    out = []
    for decl in body.split(";"):
        decl = " ".join(decl.split())
        if not decl:
            continue
        m = re.match(r"(const )?(\w+)(\s*\*)?\s*(\w+)(\[(\d+)\])?$", decl)
        assert m, decl
        base, ptr, name, arr = m.group(2), m.group(3), m.group(4), m.group(6)
        if ptr:
            t = "pointer"
        elif arr:
            t = "%s*%s" % (C_TO_CTYPES[base], arr)
        else:
            t = C_TO_CTYPES[base]
        out.append((name, t))
    return out


def _binding_fields(cls):
    import ctypes as C
    out = []
    for name, t in cls._fields_:
        if t is C.c_void_p or (isinstance(t, type) and issubclass(t, C._Pointer)):
            out.append((name, "pointer"))
        elif isinstance(t, type) and issubclass(t, C.Array):
            out.append((name, "c_%s*%d" % ({"d": "double"}[t._type_._type_], t._length_)))
        else:
            size, signed = C.sizeof(t), t(-1).value < 0 if t is not C.c_double else True
            out.append((name, "c_double" if t is C.c_double else "c_%sint%d" % ("" if signed else "u", 8 * size)))
    return out


@pytest.mark.parametrize("cname", sorted(HEADER_STRUCTS))
def test_binding_structs_match_the_header_field_by_field(cname):
    """include/fdlp_b200.h is the contract; _lib.py must declare the same fields in the same order with the same
    types, and the library's own sizeof() must equal ctypes.sizeof (VERDICT r1 weak #10 / ADVICE: a stale stub made
    the library read past the end of fdlp_params)."""
    import ctypes as C
    header = open(os.path.join(ROOT, "include", "fdlp_b200.h")).read()
    cls = getattr(_lib, HEADER_STRUCTS[cname])
    assert _header_fields(header, cname) == _binding_fields(cls)
    which = [k for k, v in _lib.STRUCTS.items() if v == HEADER_STRUCTS[cname]][0]
    assert _lib.load().fdlp_sizeof_struct(which) == C.sizeof(cls)
    if cname != "fdlp_geometry":  # every struct the caller fills leads with struct_size
        assert cls._fields_[0][0] == "struct_size" and cls().struct_size == C.sizeof(cls)


def test_integration_md_stub_is_the_generated_one():
    import subprocess
    import sys
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    gen = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_binding_stub.py")], check=True,
                         capture_output=True, text=True).stdout.strip()
    assert gen in doc, "INTEGRATION.md is stale: run python tools/gen_binding_stub.py --write"


def test_wrong_struct_size_is_refused_before_anything_is_read():
    """A binding made from another header version (shorter or longer struct) gets FDLP_E_ARG, no CUDA call needed."""
    import ctypes as C
    lib = _lib.load()
    w = np.ones(50)
    for delta in (-12, 4):
        prm = _lib.FdlpParams()
        prm.srate, prm.nfilters, prm.order, prm.coeff_num, prm.frate = 16000, 20, 50, 50, 100
        prm.fduration, prm.overlap_fraction = 0.5, 0.25
        prm.fbank_args = (C.c_double * 5)(1, 0, 0, 0, 0)
        prm.weights = w.ctypes.data_as(C.POINTER(C.c_double))
        prm.struct_size = C.sizeof(prm) + delta
        plan = C.c_void_p()
        assert lib.fdlp_plan_create(C.byref(prm), C.byref(plan)) == _lib.FDLP_E_ARG
        assert b"struct_size" in lib.fdlp_last_error() and not plan.value
    mp = _lib.MelParams()
    mp.struct_size = 8
    assert lib.mel_plan_create(C.byref(mp), C.byref(plan)) == _lib.FDLP_E_ARG
    fp = _lib.MfccParams()
    fp.struct_size = 0
    assert lib.mfcc_plan_create(C.byref(fp), C.byref(plan)) == _lib.FDLP_E_ARG


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "speech_recognition_tools_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f


@pytest.mark.parametrize("fdur", [0.5, 1.5])
def test_segment_and_frame_counts_match_oracle(fdur):
    g = ofdlp.derived_sizes(fdur, 100, 0.25)
    rng = np.random.default_rng(3)
    Ts = [2, 3, 100, g["S"], g["S"] + 1, g["S"] + 2, 2 * g["S"] + 1] + list(rng.integers(2, 300000, 40))
    for T in Ts:
        fr = ofdlp.get_frames(np.ones(int(T)), 16000, g["lfr"], fdur)
        assert segment_count(int(T), g["N"], g["S"]) == fr.shape[0]
        assert frame_count(int(T), 100, 16000) == int(np.ceil(T * 100 / 16000))


def test_mel_frame_count_matches_oracle():
    for T in (2, 161, 162, 320, 321, 8000, 8001, 40001):
        fr = ofdlp.get_frames(np.ones(T), 16000, 100, 0.02)
        assert mel_frame_count(T, 320, 160) == fr.shape[0]


@pytest.mark.parametrize("fdur,T", [(0.5, 40001), (0.5, 100), (0.5, 3000), (1.5, 100000), (1.5, 17000), (1.5, 160000)])
def test_ola_pointer_trace_matches_oracle(fdur, T):
    g = ofdlp.derived_sizes(fdur, 100, 0.25)
    x = synth.white_int16(0, 0, T)
    params = dict(nfilters=4, order=8, fduration=fdur, coeff_num=10, coeff_range="0,9")
    for seed in (0, 1, 7):
        _, tr = ofdlp.fdlp_spectrogram(x, rng=random.Random(seed), return_trace=True, **params)
        ptrs = ola_pointers(tr["nseg"], tr["Fr"], g["kk"], g["kkb2"], g["kks"], random.Random(seed))
        assert ptrs == tr["ptrs"]


def test_ola_pointer_overflow_raises_like_the_reference():
    # default params drift +1 frame per segment and overflow for long utterances
    g = ofdlp.derived_sizes(0.5, 100, 0.25)
    T = 35 * 16000
    nseg, Fr = segment_count(T, g["N"], g["S"]), frame_count(T, 100, 16000)
    with pytest.raises(ValueError):
        ola_pointers(nseg, Fr, g["kk"], g["kkb2"], g["kks"], random.Random(0))


def test_weights_match_oracle():
    lifter = list(1.0 + 0.1 * np.arange(50))
    for kw in (dict(), dict(coeff_range="0,49", gamma_weight="2,3,4"), dict(lifter=lifter, coeff_range="1,40"),
               dict(odd_mod_zero=True, coeff_range="0,49"), dict(coeff_num=450, coeff_range="1,450", order=150)):
        cfg = FdlpConfig(**kw)
        w = build_weights(cfg)
        ref = ofdlp.build_weights(cfg.coeff_num, cfg.coeff_range, cfg.order, cfg.fduration, cfg.gamma_weight, cfg.lifter)
        if cfg.odd_mod_zero:
            ref = ref.copy()
            ref[1::2] = 0
        np.testing.assert_array_equal(w, ref)
    with pytest.raises(ValueError):
        build_weights(FdlpConfig(gamma_weight="2,3,4", order=50, coeff_num=40))


def test_fbank_type_parsing_errors():
    with pytest.raises(ValueError):
        parse_fbank_type("mel")
    with pytest.raises(ValueError):
        parse_fbank_type("cochlear,1,1")
    with pytest.raises(ValueError):
        parse_fbank_type("gammatone,1")
    assert parse_fbank_type("cochlear,1,1,1,2.5,1")[1] == [1.0, 1.0, 1.0, 2.5, 1.0]


def test_synth_is_deterministic_and_never_silent():
    a = synth.white_int16(3, 5, 5000)
    b = synth.white_int16(3, 5, 5000)
    assert np.array_equal(a, b) and a.dtype == np.int16
    assert 2500 < a.astype(np.float64).std() < 3500
    s = synth.speechlike_int16(3, 5, 48000, 30.0)
    assert s.dtype == np.int16 and np.abs(s).max() > 1000
    for k in range(0, 48000 - 8000, 6000):
        assert np.any(s[k:k + 8000] != 0)


def test_reverb_block_layout_covers_the_full_convolution():
    """Host tables of fdlp_reverb_forward: the pairs of an utterance cover all T + R - 1 samples of the full
    convolution (features.py:111), owners and prefix sums are consistent."""
    from speech_recognition_tools_b200.reverb import block_layout, load_rir
    lengths = np.array([2, 63, 64, 65, 2048, 2049, 40001, 240000])
    for R, B, m in ((1, 64, 1), (3, 64, 1), (700, 128, 2), (16000, 2048, 1), (48000, 4096, 2)):
        pair_off, pair_utt, conv_off, conv_utt = block_layout(lengths, R, B, m)
        D = np.diff(pair_off)
        assert np.all(2 * D * B >= lengths + R - 1) and np.all(2 * (D - 1) * B < lengths + R - 1 + B)
        assert np.array_equal(np.bincount(pair_utt, minlength=len(lengths)), D)
        assert np.array_equal(np.diff(conv_off), -(-D // m))
        assert np.array_equal(np.bincount(conv_utt, minlength=len(lengths)), np.diff(conv_off))
    assert load_rir(None) is None and load_rir("clean") is None
    with pytest.raises(ValueError):  # computeFDLPSpectrogram.py:90-91
        load_rir("cathedral")


@pytest.mark.parametrize("fdur", [0.5, 1.5])
def test_vectorised_batch_layout_equals_the_per_segment_loop(fdur):
    """batch_layout / ola_pointers_batch draw the jitter in the reference's order (computeFDLPSpectrogram.py:225) and
    give the trace, counts and errors of the scalar restatement, for one generator over the batch and for
    per-utterance seeds."""
    g = ofdlp.derived_sizes(fdur, 100, 0.25)
    rng = np.random.default_rng(5)
    hi = 150000 if fdur == 0.5 else 560000  # default parameters can overflow beyond ~10 s
    lengths = [2, 3, 100, g["S"], g["S"] + 1, g["S"] + 2, 2 * g["S"] + 1] + [int(t) for t in rng.integers(2, hi, 200)]

    def scalar(rngs):
        out = []
        for T, r in zip(lengths, rngs):
            out += ola_pointers(segment_count(T, g["N"], g["S"]), frame_count(T, 100, 16000), g["kk"], g["kkb2"],
                                g["kks"], r)
        return out

    one = random.Random(11)
    lay = batch_layout(lengths, g, 100, 16000, jitter=random.Random(11))
    assert list(lay.seg_ptr) == scalar([one] * len(lengths))
    assert lay.n_seg == sum(segment_count(T, g["N"], g["S"]) for T in lengths)
    assert list(np.diff(lay.utt_frame_off)) == [frame_count(T, 100, 16000) for T in lengths]
    seeds = list(range(100, 100 + len(lengths)))
    lay2 = batch_layout(lengths, g, 100, 16000, jitter=seeds)
    assert list(lay2.seg_ptr) == scalar([random.Random(s) for s in seeds])
    # the generator is left in the same state as after the loop
    a, b = random.Random(3), random.Random(3)
    batch_layout(lengths, g, 100, 16000, jitter=a)
    scalar([b] * len(lengths))
    assert a.random() == b.random()
    if fdur == 0.5:  # both raise, or neither, on lengths where the reference's pointer drifts past the utterance
        for T in range(150000, 330000, 3331):
            for seed in (0, 1):
                try:
                    ola_pointers(segment_count(T, g["N"], g["S"]), frame_count(T, 100, 16000), g["kk"], g["kkb2"], g["kks"],
                                 random.Random(seed))
                    raised = False
                except ValueError:
                    raised = True
                if raised:
                    with pytest.raises(ValueError, match="overlap-add pointer ran past"):
                        batch_layout([T], g, 100, 16000, jitter=[seed])
                else:
                    batch_layout([T], g, 100, 16000, jitter=[seed])


def test_make_fbank_entry_point_matches_reference_fixtures():
    """fdlp_make_fbank (host only, no device work): createFbank / createFbankCochlear values of the unmodified
    reference (tests/golden/fbanks.npz, written by tests/golden/make_golden.py)."""
    import ctypes as C
    lib = _lib.load()
    z = np.load(os.path.join(ROOT, "tests", "golden", "fbanks.npz"))

    def build(kind, args, J, nfft):
        out = np.empty((J, nfft // 2 + 1), dtype=np.float64)
        a = np.array(list(args) + [0.0] * (5 - len(args)), dtype=np.float64)
        assert lib.fdlp_make_fbank(kind, a.ctypes.data_as(C.c_void_p), J, nfft, 16000, out.ctypes.data_as(C.c_void_p)) == 0
        return out
    mel = build(_lib.FDLP_FBANK_MEL, [1.0], 20, 16000)
    np.testing.assert_allclose(mel[:, ::53], z["mel20_16000"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(build(_lib.FDLP_FBANK_MEL, [1.0], 80, 2048), z["mel80_2048"], rtol=0, atol=1e-15)
    coch = build(_lib.FDLP_FBANK_COCHLEAR, [1.0, 1.0, 1, 2.5, 1.0], 80, 48000)
    np.testing.assert_allclose(coch[:, ::53], z["coch80_48000"], rtol=1e-12, atol=0)
    nf = build(_lib.FDLP_FBANK_COCHLEAR, [0.2, 2.5, 0, 2.5, 1.3], 12, 4000)
    np.testing.assert_allclose(nf, z["coch12_4000_nonfixed"], rtol=1e-12, atol=0)
    assert lib.fdlp_make_fbank(7, np.zeros(5).ctypes.data_as(C.c_void_p), 4, 64, 16000, nf.ctypes.data_as(C.c_void_p)) == _lib.FDLP_E_ARG


def test_ark_writer_prealloc_is_only_a_hint(tmp_path):
    """ArkWriter.prealloc reserves pages ahead of the batches (fallocate); too much, too little or none at all must give
    the same bytes, and the file ends at its real length."""
    from concurrent.futures import ThreadPoolExecutor
    from speech_recognition_tools_b200.kaldi_io import ArkWriter, read_ark
    rng = np.random.default_rng(5)
    frames = [int(v) for v in rng.integers(1, 400, 40)]
    off = np.concatenate([[0], np.cumsum(frames)])
    mat = rng.standard_normal((int(off[-1]), 7)).astype(np.float32)
    keys = ["utt%03d" % i for i in range(40)]
    exact = sum(ArkWriter.entry_bytes(k, f, 7) for k, f in zip(keys, frames))
    blobs = {}
    with ThreadPoolExecutor(4) as pool:
        for name, hint in (("none", 0), ("exact", exact), ("short", exact // 3), ("long", 2 * exact + 12345)):
            ark, scp = str(tmp_path / (name + ".ark")), str(tmp_path / (name + ".scp"))
            w = ArkWriter(ark, scp, pool=pool, threads=4, scp_ark_name="x.ark")
            w.prealloc(hint)
            w.write_batch(keys[:25], mat[:off[25]], off[:26])
            w.prealloc(hint // 2)
            w.write_batch(keys[25:], mat[off[25]:], off[25:] - off[25])   # second batch: small, takes the pwritev path
            w.close()
            blobs[name] = (open(ark, "rb").read(), open(scp).read())
            assert len(blobs[name][0]) == exact
    assert blobs["exact"] == blobs["none"] == blobs["short"] == blobs["long"]
    got = read_ark(str(tmp_path / "long.ark"))
    for i, k in enumerate(keys):
        np.testing.assert_array_equal(got[k], mat[off[i]:off[i + 1]])


def test_native_wav_probe_and_read_match_scipy_and_the_python_probe(tmp_path):
    """fdlp_wav_probe / fdlp_wav_read (host-only entry points of the CLI reader): same verdicts as pipeline.probe_wav on
    PCM16, float32, stereo, a file with a LIST chunk before the data, a truncated file, a non-WAV and a missing file;
    samples equal scipy.io.wavfile.read."""
    import ctypes as C
    import struct
    from scipy.io import wavfile
    from speech_recognition_tools_b200.cli import pipeline as pl
    lib = _lib.load()
    rng = np.random.default_rng(11)
    a16 = rng.integers(-3000, 3000, 5000).astype(np.int16)
    a32 = rng.standard_normal(3001).astype(np.float32)
    files = {}
    for name, data in (("pcm16", a16), ("f32", a32), ("stereo", np.stack([a16, a16], axis=1)), ("u8", (a16 // 64 + 128).astype(np.uint8))):
        files[name] = str(tmp_path / (name + ".wav"))
        wavfile.write(files[name], 16000, data)
    raw = open(files["pcm16"], "rb").read()
    lst = b"LIST" + struct.pack("<I", 10) + b"INFOabcdef"
    body = raw[12:36] + lst + raw[36:]
    files["list"] = str(tmp_path / "list.wav")
    open(files["list"], "wb").write(b"RIFF" + struct.pack("<I", len(body) + 4) + b"WAVE" + body)
    files["cut"] = str(tmp_path / "cut.wav")
    open(files["cut"], "wb").write(raw[:44 + 2001])        # data chunk says 10000 bytes, 2001 are there
    files["text"] = str(tmp_path / "text.wav")
    open(files["text"], "wb").write(b"not a wav file at all, but longer than forty-four bytes for sure.")
    files["missing"] = str(tmp_path / "missing.wav")
    names = list(files)
    paths = (C.c_char_p * len(names))(*[os.fsencode(files[n]) for n in names])
    info = np.zeros((len(names), 4), dtype=np.int64)
    for threads in (1, 4):
        info[:] = -7
        assert lib.fdlp_wav_probe(paths, len(names), threads, info.ctypes.data) == 0
        for r, n in enumerate(names):
            py = pl.probe_wav(files[n])
            if py is None:
                assert info[r, 1] == -1, n
            else:
                fmt = 0 if py.dtype == np.dtype("<i2") else 1
                assert tuple(info[r]) == (py.srate, fmt, py.data_off, py.n_samples), n
    got = dict(zip(names, info))
    assert got["pcm16"][3] == 5000 and got["f32"][3] == 3001 and got["list"][3] == 5000 and got["cut"][3] == 1000
    assert all(got[n][1] == -1 for n in ("stereo", "u8", "text", "missing"))
    # read the four plain files back to back into one buffer
    take = ["pcm16", "list", "cut", "f32"]
    nbytes = np.array([got[n][3] * (4 if n == "f32" else 2) for n in take], dtype=np.int64)
    dst_off = np.concatenate([[0], np.cumsum(nbytes)[:-1]]).astype(np.int64)
    src_off = np.array([got[n][2] for n in take], dtype=np.int64)
    buf = np.zeros(int(nbytes.sum()), dtype=np.uint8)
    status = np.full(len(take), 9, dtype=np.int32)
    p2 = (C.c_char_p * len(take))(*[os.fsencode(files[n]) for n in take])
    assert lib.fdlp_wav_read(p2, src_off.ctypes.data, nbytes.ctypes.data, dst_off.ctypes.data, buf.ctypes.data, len(take), 3,
                             status.ctypes.data) == 0
    assert not status.any()
    np.testing.assert_array_equal(buf[:10000].view(np.int16), a16)
    np.testing.assert_array_equal(buf[10000:20000].view(np.int16), a16)
    np.testing.assert_array_equal(buf[20000:22000].view(np.int16), a16[:1000])
    np.testing.assert_array_equal(buf[22000:].view(np.float32), wavfile.read(files["f32"])[1])
    # asking for more than the file holds is reported per file
    nbytes[2] += 64
    assert lib.fdlp_wav_read(p2, src_off.ctypes.data, nbytes.ctypes.data, dst_off.ctypes.data, buf.ctypes.data, 3, 2,
                             status.ctypes.data) == 0
    assert list(status[:3]) == [0, 0, 1]


def test_host_threads_shares_the_cores_between_local_ranks(monkeypatch):
    from speech_recognition_tools_b200.fdlp import host_threads
    cores = len(os.sched_getaffinity(0))
    monkeypatch.delenv("LOCAL_WORLD_SIZE", raising=False)
    assert host_threads(8) == max(2, min(8, cores))
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "8")
    assert host_threads(16) == max(2, min(16, cores // 8))
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "1000")
    assert host_threads(16) == 2


def test_ark_three_decimal_mode_matches_the_text_round_trip(tmp_path, monkeypatch):
    """FDLP_ARK_3DP=1: the values the reference's dict2Ark leaves on disk ('%.3f' text, features.py:63-69, parsed back
    to float32 by copy-feats), in both writers; off by default."""
    import io
    from speech_recognition_tools_b200.kaldi_io import ArkWriter, read_ark, write_ark_scp
    rng = np.random.default_rng(3)
    mat = (rng.standard_normal((50, 6)) * 20).astype(np.float32)
    mat[0, 0], mat[1, 1] = -np.inf, 0.0005
    buf = io.StringIO()
    np.savetxt(buf, mat.astype(np.float64), fmt="%.3f")
    text = np.loadtxt(io.StringIO(buf.getvalue()), dtype=np.float64).astype(np.float32)
    off = np.array([0, 20, 50])
    monkeypatch.setenv("FDLP_ARK_3DP", "1")
    w = ArkWriter(str(tmp_path / "a.ark"), str(tmp_path / "a.scp"))
    w.write_batch(["u0", "u1"], mat, off)
    w.close()
    write_ark_scp([("u0", mat[:20]), ("u1", mat[20:])], str(tmp_path / "b.ark"), str(tmp_path / "b.scp"))
    for name in ("a.ark", "b.ark"):
        got = read_ark(str(tmp_path / name))
        np.testing.assert_array_equal(np.concatenate([got["u0"], got["u1"]]), text)
    monkeypatch.delenv("FDLP_ARK_3DP")
    write_ark_scp([("u0", mat)], str(tmp_path / "c.ark"), str(tmp_path / "c.scp"))
    np.testing.assert_array_equal(read_ark(str(tmp_path / "c.ark"))["u0"], mat)
