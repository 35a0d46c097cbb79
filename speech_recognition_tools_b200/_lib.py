"""ctypes binding of libfdlp_b200.so (C ABI declared in include/fdlp_b200.h).

The library is the product: there is no CPU fallback.  Importing this module without the built
shared object raises ImportError with the build command (`make` / `__graft_entry__.build()`)."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# FDLP_B200_LIB: another build of the same library (A/B runs of compile-time kernel variants)
LIB_PATH = os.environ.get("FDLP_B200_LIB") or os.path.join(_HERE, "libfdlp_b200.so")

FDLP_OK = 0
FDLP_E_ARG, FDLP_E_CUDA, FDLP_E_UNSUPPORTED, FDLP_E_NOMEM = -1, -2, -3, -4
FDLP_ABI_VERSION = 200
FDLP_NOALLOC_MAX_DIM = 255
FDLP_ST_SINGULAR, FDLP_ST_NONFINITE, FDLP_ST_ENV_RANGE = 1, 2, 4
FDLP_FBANK_MEL, FDLP_FBANK_COCHLEAR, FDLP_FBANK_CUSTOM = 0, 1, 2
FDLP_WAV_INT16, FDLP_WAV_FLOAT32, FDLP_WAV_FLOAT64 = 0, 1, 2
FDLP_WINDOW_HAMMING, FDLP_WINDOW_HANNING, FDLP_WINDOW_SQUARE = 0, 1, 2
FDLP_NOISE_NONE, FDLP_NOISE_INT16_WRAP, FDLP_NOISE_FLOAT = 0, 1, 2


class _Sized(C.Structure):
    """Every params / batch struct of the ABI leads with `struct_size` = sizeof(the struct as THIS binding declares
    it); the library refuses a size that is not its own (include/fdlp_b200.h, conventions)."""

    def __init__(self, *args, **kw):
        super().__init__(*args, **kw)
        self.struct_size = C.sizeof(type(self))


class FdlpParams(_Sized):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("srate", C.c_int32), ("nfilters", C.c_int32), ("order", C.c_int32), ("coeff_num", C.c_int32),
        ("frate", C.c_int32), ("fduration", C.c_double), ("overlap_fraction", C.c_double),
        ("fbank_kind", C.c_int32), ("fbank_args", C.c_double * 5),
        ("weights", C.POINTER(C.c_double)), ("custom_fbank", C.POINTER(C.c_double)),
        ("dct_fp32", C.c_int32), ("tail_eps", C.c_double),
        ("window_kind", C.c_int32), ("seg_shift", C.c_int32), ("cepstrum_mode", C.c_int32), ("env_f64", C.c_int32),
    ]


class FdlpGeometry(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("N", "S", "M", "kk", "kkb2", "kks", "L", "nlags", "dct_is_fp64")] + \
               [("support_taps", C.c_int64), ("family_taps", C.c_int64), ("direct_taps", C.c_int64),
                ("nfam", C.c_int32), ("block_len", C.c_int32)]


class FdlpBatch(_Sized):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("wav", C.c_void_p), ("wav_dtype", C.c_int32), ("n_utt", C.c_int32), ("n_seg", C.c_int32),
        ("n_frames", C.c_int64), ("max_utt_frames", C.c_int32),
        ("utt_wav_off", C.c_void_p), ("utt_seg_off", C.c_void_p), ("utt_frame_off", C.c_void_p),
        ("seg_utt", C.c_void_p), ("seg_ptr", C.c_void_p),
        ("noise_mode", C.c_int32), ("noise", C.c_void_p), ("utt_noise_off", C.c_void_p), ("utt_snr_db", C.c_void_p),
        ("feats", C.c_void_p), ("cmvn_acc", C.c_void_p), ("status", C.c_void_p),
    ]


class MelParams(_Sized):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("srate", C.c_int32), ("nfilters", C.c_int32), ("nfft", C.c_int32), ("frate", C.c_int32),
        ("fduration", C.c_double), ("fbank_kind", C.c_int32), ("fbank_args", C.c_double * 5),
        ("custom_fbank", C.POINTER(C.c_double)), ("power", C.c_int32),
    ]


class MelBatch(_Sized):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("wav", C.c_void_p), ("wav_dtype", C.c_int32), ("n_utt", C.c_int32), ("n_frames", C.c_int64),
        ("utt_wav_off", C.c_void_p), ("utt_frame_off", C.c_void_p), ("feats", C.c_void_p), ("cmvn_acc", C.c_void_p),
    ]


class MfccParams(_Sized):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("srate", C.c_int32), ("nfilters", C.c_int32), ("nfft", C.c_int32), ("frate", C.c_int32),
        ("fduration", C.c_double), ("ncep", C.c_int32),
    ]


class MfccBatch(_Sized):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("wav", C.c_void_p), ("wav_dtype", C.c_int32), ("n_utt", C.c_int32), ("n_frames", C.c_int64),
        ("utt_wav_off", C.c_void_p), ("utt_frame_off", C.c_void_p), ("feats", C.c_void_p), ("context", C.c_int32),
        ("noise", C.c_void_p), ("utt_noise_off", C.c_void_p), ("utt_snr_db", C.c_void_p),
    ]


class ReverbBatch(_Sized):
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("wav", C.c_void_p), ("wav_dtype", C.c_int32), ("n_utt", C.c_int32), ("utt_wav_off", C.c_void_p),
        ("max_len", C.c_int64), ("n_pair", C.c_int32), ("utt_pair_off", C.c_void_p), ("pair_utt", C.c_void_p),
        ("n_conv", C.c_int32), ("utt_conv_off", C.c_void_p), ("conv_utt", C.c_void_p),
        ("noise_mode", C.c_int32), ("noise", C.c_void_p), ("utt_noise_off", C.c_void_p), ("utt_snr_db", C.c_void_p),
        ("out", C.c_void_p), ("out_dtype", C.c_int32), ("ind", C.c_void_p), ("out_len", C.c_void_p),
    ]


# every symbol include/fdlp_b200.h declares: (name, restype, argtypes)
SYMBOLS = [
    ("fdlp_last_error", C.c_char_p, []),
    ("fdlp_version", C.c_int, []),
    ("fdlp_sizeof_struct", C.c_int64, [C.c_int32]),
    ("fdlp_plan_create", C.c_int, [C.POINTER(FdlpParams), C.POINTER(C.c_void_p)]),
    ("fdlp_plan_destroy", C.c_int, [C.c_void_p]),
    ("fdlp_plan_get_geometry", C.c_int, [C.c_void_p, C.POINTER(FdlpGeometry)]),
    ("fdlp_plan_get_fbank", C.c_int, [C.c_void_p, C.c_void_p]),
    ("fdlp_plan_scratch_bytes", C.c_int64, [C.c_void_p, C.c_int32]),
    ("fdlp_forward", C.c_int, [C.c_void_p, C.POINTER(FdlpBatch), C.c_void_p]),
    ("fdlp_plan_set_profiling", C.c_int, [C.c_void_p, C.c_int32]),
    ("fdlp_plan_set_overlap", C.c_int, [C.c_void_p, C.c_int32]),
    ("fdlp_plan_get_stage_ms", C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    ("fdlp_plan_set_autocorr_mode", C.c_int, [C.c_void_p, C.c_int32]),
    ("fdlp_plan_get_autocorr_ms", C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    ("fdlp_plan_get_kernel_ms", C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    ("fdlp_stage_dct", C.c_int, [C.c_void_p, C.POINTER(FdlpBatch), C.c_void_p, C.c_void_p]),
    ("fdlp_stage_autocorr", C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    ("fdlp_stage_lpc_envelope", C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p]),
    ("mel_plan_create", C.c_int, [C.POINTER(MelParams), C.POINTER(C.c_void_p)]),
    ("mel_plan_destroy", C.c_int, [C.c_void_p]),
    ("mel_plan_get_fbank", C.c_int, [C.c_void_p, C.c_void_p]),
    ("mel_forward", C.c_int, [C.c_void_p, C.POINTER(MelBatch), C.c_void_p]),
    ("mfcc_plan_create", C.c_int, [C.POINTER(MfccParams), C.POINTER(C.c_void_p)]),
    ("mfcc_plan_destroy", C.c_int, [C.c_void_p]),
    ("mfcc_plan_set_tensor_cores", C.c_int, [C.c_void_p, C.c_int32]),
    ("mfcc_forward", C.c_int, [C.c_void_p, C.POINTER(MfccBatch), C.c_void_p]),
    ("cmvn_apply", C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]),
    ("cmvn_accumulate_noalloc", C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    ("fdlp_modspec_forward", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    ("fdlp_make_fbank", C.c_int, [C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    ("fdlp_stage_frames", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("fdlp_modspec_complex_forward", C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                          C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("fdlp_wav_probe", C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    ("fdlp_wav_read", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    ("fdlp_diff_filter", C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]),
    ("fdlp_noise_mix", C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("fdlp_reverb_plan_create", C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    ("fdlp_reverb_plan_destroy", C.c_int, [C.c_void_p]),
    ("fdlp_reverb_plan_get_block", C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    ("fdlp_reverb_plan_set_profiling", C.c_int, [C.c_void_p, C.c_int32]),
    ("fdlp_reverb_plan_get_kernel_ms", C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    ("fdlp_reverb_plan_scratch_bytes", C.c_int64, [C.c_void_p, C.c_int32, C.c_int32]),
    ("fdlp_reverb_forward", C.c_int, [C.c_void_p, C.POINTER(ReverbBatch), C.c_void_p]),
    ("fdlp_synth_white", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_int32,
                                   C.c_void_p]),
    ("fdlp_measure_fp64_peak", C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_double)]),
]

_lib = None


def load():
    """Load the shared library once; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libfdlp_b200.so is not built (%s). Run `make` or `python -c 'import __graft_entry__ as g; g.build()'` "
            "at the repository root. There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, restype, argtypes in SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.fdlp_version() != FDLP_ABI_VERSION:
        raise ImportError("%s implements ABI %d, this binding was written for ABI %d: rebuild the library (`make`)"
                          % (LIB_PATH, lib.fdlp_version(), FDLP_ABI_VERSION))
    _lib = lib
    return lib


# index -> binding class for fdlp_sizeof_struct (include/fdlp_b200.h)
STRUCTS = {0: "FdlpParams", 1: "FdlpGeometry", 2: "FdlpBatch", 3: "MelParams", 4: "MelBatch", 5: "ReverbBatch",
           6: "MfccParams", 7: "MfccBatch"}


class FdlpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libfdlp_b200 error %d: %s" % (code, msg))
        self.code = code


def check_count(rc):
    """For entry points that return a non-negative count / flag, or a negative FDLP_E_* code."""
    if rc < 0:
        check(rc)
    return rc


def check(rc):
    if rc != FDLP_OK:
        msg = load().fdlp_last_error()
        raise FdlpError(rc, msg.decode("utf-8", "replace") if msg else "")
