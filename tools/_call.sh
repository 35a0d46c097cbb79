mkdir -p gpurun_out
N=4
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
tail -2 gpurun_out/bench_${N}gpu.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_4gpu.json').read().strip().splitlines()[-1])
print(d['n_gpus'], round(d['value']), d['ms_per_step'], round(d['e2e']['value']), d['e2e']['ms_per_step'], d['e2e_cli']['wall_s'], d['multi_gpu_check'])
PY
