set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_modspec_gpu.py -x -q -m gpu > gpurun_out/modspec_cplx.log 2>&1
tail -30 gpurun_out/modspec_cplx.log
timeout 300 python - <<'PY' > gpurun_out/modspec_cplx_time.log 2>&1
import time, torch, numpy as np
from speech_recognition_tools_b200 import synth
from speech_recognition_tools_b200.modspec import ModSpecConfig, ModSpecExtractor
wavs=[synth.make_utterance("speechlike",5,i,160000,30.0) for i in range(8)]
for cplx in (False, True):
    ex=ModSpecExtractor(ModSpecConfig(complex_modulation=cplx),device="cuda:0")
    ex(wavs); torch.cuda.synchronize()
    t=time.perf_counter(); ex(wavs); torch.cuda.synchronize(); dt=time.perf_counter()-t
    print("complex" if cplx else "real", "80 s audio in %.1f ms -> %.0f xRT"%(dt*1e3, 80/dt))
    ex.close()
PY
cat gpurun_out/modspec_cplx_time.log
