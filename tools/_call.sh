mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/gputest.log 2>&1; tail -3 gpurun_out/gputest.log
timeout 600 python bench.py > gpurun_out/bench_v4.json 2> gpurun_out/bench_v4.err; tail -2 gpurun_out/bench_v4.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_v4.json 2> gpurun_out/bench_ref_v4.err; tail -c 600 gpurun_out/bench_ref_v4.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_v4.json').read().strip().splitlines()[-1])
for k in ['value','ms_per_step','e2e','e2e_cli','mfcc','mel_baseline','gpu_launches']:
    v=d.get(k)
    if isinstance(v,dict): v={kk:vv for kk,vv in v.items() if not isinstance(vv,str) or len(vv)<40}
    print(k, v)
print(d['roofline']['frac'], d['roofline']['step_frac'])
PY
