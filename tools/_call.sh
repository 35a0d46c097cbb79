mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/gputest.log 2>&1; tail -3 gpurun_out/gputest.log
timeout 600 python bench.py > gpurun_out/bench_v5.json 2> gpurun_out/bench_v5.err; tail -2 gpurun_out/bench_v5.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_v5.json').read().strip().splitlines()[-1])
print(round(d['value']), d['ms_per_step'], round(d['e2e']['value']), d['e2e_cli']['wall_s'], d['mfcc']['tensor_core']['ms_per_step'], d['roofline']['frac'], d['clocks'])
PY
