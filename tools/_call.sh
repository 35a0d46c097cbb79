mkdir -p gpurun_out
timeout 400 python tools/perf_cli.py 1000 2048,1536,3072 > gpurun_out/perf_cli.log 2>&1
grep -v "Computing Features" gpurun_out/perf_cli.log | grep "wall" | grep -v "rep=0"
