mkdir -p gpurun_out
timeout 400 ncu --set full --import-source on --clock-control none -k regex:dct_segments -c 1 -o /tmp/dct python tools/perf_step.py 300 > gpurun_out/dct_ncu.log 2>&1
ncu -i /tmp/dct.ncu-rep --page source --csv --print-source sass > gpurun_out/dct_src.csv 2>/dev/null
ls -la gpurun_out/dct_src.csv; head -c 1500 gpurun_out/dct_src.csv
