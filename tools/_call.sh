set -x
mkdir -p gpurun_out
N=${NG:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err
tail -3 gpurun_out/bench_${N}gpu.err
python - <<'PY'
import json,glob
for f in glob.glob('gpurun_out/bench_*gpu.json'):
    d=json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e'].get('packed_input_ms_per_step'), d['e2e_cli']['wall_s'], d['multi_gpu_check'])
PY
