import sys, numpy as np, torch
sys.path.insert(0, ".")
from speech_recognition_tools_b200 import synth
from speech_recognition_tools_b200.mel import MelConfig, MelExtractor
n=1000
lengths = synth.wsj_shaped_lengths(n)
wav, off = synth.white_int16_device(9, np.arange(n), lengths, device="cuda:0")
mel = MelExtractor(MelConfig(nfilters=80, nfft=2048, fbank_type="mel,1"), device="cuda:0")
lay = mel.layout(lengths).to(mel.device)
f = torch.empty((lay.n_frames, 80), dtype=torch.float32, device="cuda:0")
for _ in range(3): mel.forward_packed(wav, lay, feats=f)
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): mel.forward_packed(wav, lay, feats=f)
e1.record(); torch.cuda.synchronize()
print("mel ms", e0.elapsed_time(e1)/5, "frames", lay.n_frames)
