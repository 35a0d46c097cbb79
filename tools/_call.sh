cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2_t16.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_r02_v2.json 2> gpurun_out/bench_r02_v2.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_r02_v2.json 2> gpurun_out/bench_ref_r02_v2.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r02_v2.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:'dct_segments|family_block|family_scan|autocorr_ranges|lpc_cepstrum|envelope_dmma|ola_log|cmvn_reduce' --launch-skip 24 --launch-count 8 -f -o /tmp/r2_full_v2 python tools/perf_step.py 1000 > gpurun_out/ncu_full.log 2>&1
ncu -i /tmp/r2_full_v2.ncu-rep --page raw --csv > gpurun_out/r2_full_v2.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none -k regex:'lpc_cepstrum' --launch-skip 3 --launch-count 1 -f -o /tmp/r2_lpc_L300 python tools/perf_step.py 1000 coeff_num=450 coeff_range=1,450 > gpurun_out/ncu_lpc.log 2>&1
ncu -i /tmp/r2_lpc_L300.ncu-rep --page raw --csv > gpurun_out/r2_lpc_L300.csv 2>/dev/null
du -sh gpurun_out
