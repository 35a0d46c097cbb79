"""CPU oracle for the FDLP-spectrogram hot path -- TEST INFRASTRUCTURE ONLY.

This package is a numpy/scipy restatement of the reference extractor
(`/root/reference/src/featgen/{computeFDLPSpectrogram,computeMelSpectrum,features}.py`).
It exists so that the CUDA path can be checked on a box where `/root/reference` is absent.

Rules:
  * only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference`
    legs may import it -- always as the checker or the timed CPU baseline, never as the product;
  * the product package `speech_recognition_tools_b200` never imports it and has no CPU fallback.

Parity pinning: the reference repository holds no tests, golden vectors or fixtures for this path
(SURVEY.md section 4).  The oracle is therefore pinned against outputs of the reference code itself,
run in the build container by `tests/golden/make_golden.py` (imports the unmodified reference modules)
and committed as `tests/golden/*.npz`; `tests/test_oracle_golden.py` replays them.
The Kaldi boundary (`compute-cmvn-stats`, binary ark) is un-vendored and un-pinned upstream:
`oracle/cmvn.py` restates Kaldi's documented layout -- parity unpinned for that row.
"""
