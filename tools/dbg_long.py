import os, sys, random
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from speech_recognition_tools_b200 import synth
from speech_recognition_tools_b200.fdlp import FdlpConfig, FdlpExtractor
from oracle import fdlp as ofdlp
for fdur, order, kind in ((2.0, 300, "speechlike"), (1.5, 300, "speechlike"), (2.0, 150, "speechlike"), (2.0, 300, "white")):
    kw = dict(nfilters=12, order=order, fduration=fdur, frate=100, overlap_fraction=0.25, coeff_num=300,
              coeff_range="0,300", fbank_type="cochlear,1,1,1,2.5,1")
    ex = FdlpExtractor(FdlpConfig(**kw))
    x = synth.make_utterance(kind, 41, 0, 90000, 300.0)
    fb = ofdlp.make_fbank(kw["fbank_type"], 12, int(2*fdur*16000), 16000)
    st = ofdlp.fdlp_stages(x, nfilters=12, order=order, fduration=fdur, overlap_fraction=0.25, coeff_num=300, fbank=fb)
    dct, lay = ex.stage_dct([x])
    d = dct.cpu().numpy()
    print(fdur, order, kind, "geom", ex.geom)
    print(" dct relerr", np.abs(d - st["dct"]).max() / np.abs(st["dct"]).max())
    r = ex.stage_autocorr(torch.from_numpy(st["dct"]).cuda()).cpu().numpy()
    print(" r relerr (oracle dct in)", (np.abs(r - st["r"]) / np.abs(st["r"][:, :, :1])).max())
    out = ex.stage_lpc_envelope(torch.from_numpy(st["r"]).cuda())
    print(" gain relerr", (np.abs(out["gain"].cpu().numpy() - st["gg"]) / np.abs(st["gg"])).max(), "status", out["status"].cpu().numpy())
    print(" cep err", np.abs(out["cep"].cpu().numpy() - st["cep"][:, :, :ex.geom["L"]]).max())
    got = ex([x], jitter=[3])[0].cpu().numpy().astype(np.float64)
    ref = ofdlp.fdlp_spectrogram(x, rng=random.Random(3), **kw)
    fl = ref <= np.log(1e-14) + 1e-6
    print(" final maxabs", np.nanmax(np.abs(got - ref)[~fl]), "nan", np.isnan(got).sum(), "floor eq", np.array_equal(fl, got <= np.log(1e-14) + 1e-3))
    ex.close()
