#!/usr/bin/env python
"""Developer probe (run on the GPU box): every golden case through the CUDA path, error table printed."""
import os
import random
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor, NoiseSpec  # noqa: E402
from tests.golden_util import FDLP_DEFAULTS, golden_cases, load_case, split_params  # noqa: E402

FLOOR = np.log(1e-14)


def run_case(name, cache):
    c = load_case(name)
    meta = c["meta"]
    p, add_noise, lifter = split_params(meta)
    kw = dict(FDLP_DEFAULTS, **p)
    key = repr(sorted(kw.items())) + repr(lifter)
    if key not in cache:
        cache[key] = FdlpExtractor(FdlpConfig(lifter=lifter, **kw))
    ex = cache[key]
    noise = None
    if add_noise and add_noise not in ("clean",):
        if add_noise == "diff":
            return name, None
        nm, snr = add_noise.split(",")
        noise = NoiseSpec(samples=c["noise"]["data"], snr_db=float(snr), rand=meta["seed"])
    t0 = time.time()
    outs, st = ex([c["x"]], jitter=[meta["seed"]], noise=noise, return_status=True)
    got = outs[0].cpu().numpy().astype(np.float64)
    dt = time.time() - t0
    ref = c["feats"]
    ok_shape = got.shape == ref.shape
    fl_ref = ref <= FLOOR + 1e-6
    fl_got = got <= FLOOR + 1e-3
    m = ~fl_ref
    err = np.abs(got - ref)
    maxabs = float(np.nanmax(err[m])) if m.any() else 0.0
    rel = float(np.nanmax(err[m] / np.maximum(np.abs(ref[m]), 1e-3))) if m.any() else 0.0
    return name, dict(shape=ok_shape, floor=bool(np.array_equal(fl_ref, fl_got)), maxabs=maxabs, rel=rel,
                      status=int(st.cpu()[0]), sec=dt, nan=int(np.isnan(got).sum()))


def main():
    cache = {}
    for name in golden_cases("fdlp"):
        try:
            n, r = run_case(name, cache)
            print("%-30s %s" % (n, r))
        except Exception as e:  # noqa: BLE001
            print("%-30s EXC %r" % (name, e))
        sys.stdout.flush()


if __name__ == "__main__":
    main()
