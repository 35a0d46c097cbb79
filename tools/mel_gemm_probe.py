#!/usr/bin/env python
"""Developer probe (run on the GPU box): would a tensor-core DFT beat the FFT kernel of the mel path?

The mel extractor (computeMelSpectrum.py:147-152) is |FFT_2048(320-sample Hamming frame)|[:1025] @ fbank.T -> log10.
As GEMMs: C = A[F x 320] . B[320 x 2050] (cos | -sin), magnitude, [F x 1025] . [1025 x 80].  This probe times the first
GEMM -- 98 % of the flops -- with cuBLAS standing in for a hand-written tcgen05 kernel (an upper bound on what such a
kernel's mainloop can do; a fused kernel would not write C, so the pure-math bound flops / measured bf16 peak is printed
beside it), in the precisions a tensor-core kernel could use (TF32 x1 / x3, BF16 x1 / x3 / x6 split products, all with
FP32 accumulation), and measures what each costs in accuracy: max |d log10 mel| against a float64 evaluation, on white
and on speech-like utterances (60 dB of spectral tilt), where the reference's parity tolerance on this path is 1e-4.

    python tools/mel_gemm_probe.py [n_utt_timing]        -> prints a JSON object, one entry per variant
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200 import synth  # noqa: E402
from speech_recognition_tools_b200.mel import MelConfig, MelExtractor  # noqa: E402

N, SHIFT, NFFT, J = 320, 160, 2048, 80
DEV = "cuda:0"


def frames_of(wavs):
    """getFrames (features.py:118-154): reflect padding by N/2 - 1, Hamming window, shift 160; float64 [F, 320]."""
    win = torch.from_numpy(np.hamming(N)).to(DEV)
    out = []
    for w in wavs:
        x = torch.from_numpy(np.asarray(w, dtype=np.float64)).to(DEV)
        xp = torch.nn.functional.pad(x[None, None], (N // 2 - 1, N // 2 - 1), mode="reflect")[0, 0]
        nfr = (len(xp) - N) // SHIFT + 1
        out.append(xp.unfold(0, N, SHIFT)[:nfr] * win)
    return torch.cat(out)


def dft_matrix():
    n = torch.arange(N, dtype=torch.float64, device=DEV)[:, None]
    k = torch.arange(NFFT // 2 + 1, dtype=torch.float64, device=DEV)[None, :]
    ang = 2 * np.pi * ((n * k) % NFFT) / NFFT
    return torch.cat([torch.cos(ang), -torch.sin(ang)], dim=1)   # [320, 2050]


def split(x, dtype, parts):
    """x (float32) = p1 + p2 + ... with every part representable in `dtype` (bf16) or TF32 (10-bit mantissa)."""
    out, r = [], x.clone()
    for _ in range(parts):
        if dtype == "tf32":
            p = (r.view(torch.int32) & ~0x1FFF).view(torch.float32)  # truncation to 10 mantissa bits
        else:
            p = r.to(torch.bfloat16).to(torch.float32)
        out.append(p)
        r = r - p
    return out


def gemm_variant(A32, B32, kind):
    """Returns a closure computing C (float32) and the number of tensor-core passes."""
    if kind == "fp32":
        return (lambda: torch.mm(A32, B32)), 1, False
    dtype, parts = kind.split("x")
    parts = int(parts)
    npart = {1: 1, 3: 2, 6: 3}[parts]
    As, Bs = split(A32, dtype, npart), split(B32, dtype, npart)
    pairs = {1: [(0, 0)], 3: [(0, 0), (0, 1), (1, 0)], 6: [(0, 0), (0, 1), (1, 0), (0, 2), (2, 0), (1, 1)]}[parts]
    Ak = torch.cat([As[i] for i, _ in pairs], dim=1).contiguous()     # [F, 320 * passes]: one GEMM, FP32 accumulation
    Bk = torch.cat([Bs[j] for _, j in pairs], dim=0).contiguous()
    if dtype == "bf16":
        Ak, Bk = Ak.to(torch.bfloat16), Bk.to(torch.bfloat16)
        return (lambda: torch.mm(Ak, Bk, out_dtype=torch.float32)), len(pairs), False
    return (lambda: torch.mm(Ak, Bk)), len(pairs), True


def log_mel(C, fb_t):
    re, im = C[:, :NFFT // 2 + 1], C[:, NFFT // 2 + 1:]
    return torch.log10(torch.sqrt(re * re + im * im) @ fb_t)


def main():
    n_time = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    bf16_peak = float(peaks.get("bf16_tflops", 1644.9))
    mel = MelExtractor(MelConfig(nfilters=J, nfft=NFFT, fbank_type="mel,1"), device=DEV)
    fb = torch.from_numpy(mel.filterbank()).to(DEV)                   # [80, 1025] float64
    B64 = dft_matrix()
    B32 = B64.float()

    # accuracy sets
    sets = {"white": [synth.make_utterance("white", 71, i, 48000) for i in range(12)],
            "speechlike": [synth.make_utterance("speechlike", 72, i, 48000, 30.0) for i in range(12)]}
    truth, A_acc = {}, {}
    for name, wavs in sets.items():
        A64 = frames_of(wavs)
        truth[name] = log_mel(A64 @ B64, fb.t())
        A_acc[name] = A64.float()
        ours = torch.cat(mel(wavs)).double()
        print("# %s: %d frames, FFT kernel vs float64: max |d log10| = %.2e" % (name, A64.shape[0], float((ours - truth[name]).abs().max())),
              file=sys.stderr)

    # timing batch: the bench batch shape (white utterances generated on the device)
    lengths = synth.wsj_shaped_lengths(n_time)
    wav, off = synth.white_int16_device(9, np.arange(n_time), lengths, device=DEV)
    lay = mel.layout(lengths).to(mel.device)
    feats = torch.empty((lay.n_frames, J), dtype=torch.float32, device=DEV)
    F = int(lay.n_frames)
    A_time = torch.randn((F, N), dtype=torch.float32, device=DEV) * 3000.0

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    out = {"frames": F, "fft_kernel_ms": timed(lambda: mel.forward_packed(wav, lay, feats=feats), 5),
           "bf16_peak_tflops": bf16_peak, "variants": {}}
    flops1 = 2.0 * F * N * (NFFT + 2)
    for kind in ("fp32", "tf32x1", "tf32x3", "bf16x1", "bf16x3", "bf16x6"):
        torch.backends.cuda.matmul.allow_tf32 = kind.startswith("tf32")
        rec = {}
        for name in sets:
            fn, passes, _ = gemm_variant(A_acc[name], B32, kind)
            got = log_mel(fn().double(), fb.t())
            d = (got - truth[name]).abs()
            d = torch.where(torch.isfinite(d), d, torch.full_like(d, float("inf")))
            rec["max_dlog10_" + name] = float(d.max())
            rec["frac_above_1e-4_" + name] = float((d > 1e-4).double().mean())
        fn, passes, _ = gemm_variant(A_time, B32, kind)
        ms = timed(fn)
        rec.update({"tensor_core_passes": passes, "cublas_ms": ms, "cublas_tflops": passes * flops1 / ms * 1e-9,
                    "math_bound_ms_at_bf16_peak": (passes * flops1 / (bf16_peak * (0.5 if kind.startswith("tf32") else 1.0) * 1e12) * 1e3)
                    if kind != "fp32" else None})
        del fn
        torch.cuda.empty_cache()
        out["variants"][kind] = rec
    torch.backends.cuda.matmul.allow_tf32 = False
    print(json.dumps(out))


if __name__ == "__main__":
    main()
