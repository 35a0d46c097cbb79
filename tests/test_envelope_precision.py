"""North-star question: can the envelope contraction exp(sum_n ms[n] cos(2 pi n t / M)) go on the low
precision tensor paths?  Measured numerically (operands rounded as the tensor core would see them,
float32 accumulation) on real weighted cepstra from the oracle:
  * TF32 (10-bit mantissa operands):        5.4e-3 max-abs, above the 1e-3 target        -> rejected
  * BF16x3 (3-term bf16 split, 6 products): 1.6e-5, same as plain FP32 FMA (1.3e-5)      -> would hold
  * float64 (the DMMA kernel that ships):   exact to round-off
The contraction is 0.4 % of the path's flops and its input (the cepstrum) is already float64 in registers
of the FP64 pipeline, so the kernel that ships uses DMMA.8x8x4: no operand conversion, no split, and
the measured cost (< 1 ms per 8500 audio-seconds) leaves nothing for a bf16 tcgen05 variant to win."""
import numpy as np
import torch

from oracle import fdlp as ofdlp
from speech_recognition_tools_b200 import synth


def _round_mantissa(x, bits):
    """Round float32 values to `bits` explicit mantissa bits (tf32: 10, bf16: 7), nearest-even."""
    i = x.astype(np.float32).view(np.uint32).astype(np.uint64)
    drop = 23 - bits
    i = (i + (1 << (drop - 1)) - 1 + ((i >> drop) & 1)) >> drop << drop
    return i.astype(np.uint32).view(np.float32)


def _bf16_split3(x):
    a = _round_mantissa(x, 7)
    b = _round_mantissa((x - a).astype(np.float32), 7)
    c = _round_mantissa((x - a - b).astype(np.float32), 7)
    return a, b, c


def test_tf32_and_bf16x3_envelope_error():
    kw = dict(nfilters=80, order=150, fduration=1.5, overlap_fraction=0.25, coeff_num=100)
    fb = ofdlp.make_fbank("cochlear,1,1,1,2.5,1", 80, 48000, 16000)
    x = synth.make_utterance("speechlike", 11, 1, 40000, 30.0)
    st = ofdlp.fdlp_stages(x, fbank=fb, **kw)
    ms = st["cep"].reshape(-1, 100)                  # mask 0..100 keeps every coefficient (WSJ recipe)
    M, kk = 300, 150
    n, t = np.arange(100)[:, None], np.arange(kk)[None, :]
    basis = np.cos(2 * np.pi * ((n * t) % M) / M)
    exact = ms @ basis                                # log-envelope before exp (the feature is ~ this value)
    ms32, b32 = ms.astype(np.float32), basis.astype(np.float32)
    tf32 = (_round_mantissa(ms32, 10).astype(np.float32) @ _round_mantissa(b32, 10).astype(np.float32)).astype(np.float64)
    ma, mb = _bf16_split3(ms32), _bf16_split3(b32)
    bf = np.zeros_like(exact, dtype=np.float32)
    for i in range(3):
        for j in range(3 - i):
            bf += ma[i] @ mb[j]
    e_tf32 = np.abs(tf32 - exact).max()
    e_bf3 = np.abs(bf.astype(np.float64) - exact).max()
    e_f32 = np.abs((ms32 @ b32).astype(np.float64) - exact).max()
    print("log-envelope max-abs error: tf32 %.2e  bf16x3 %.2e  fp32 %.2e" % (e_tf32, e_bf3, e_f32))
    assert e_tf32 > 1e-3            # TF32 cannot hold the tolerance (c0 ~ 10, 10-bit mantissa)
    assert e_f32 < 1e-4             # FP32 FMA would hold it; FP64 DMMA is what ships (exact, same pipe as K3)
    assert e_bf3 < 1e-4             # BF16x3 would hold it too (6 MMAs per tile + operand conversion)
