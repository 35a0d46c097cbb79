cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q -k "fdlp" 2>&1 | tail -5 > gpurun_out/r2_t3.log
python tools/perf_step.py 1000 > gpurun_out/r2_step3.log 2>&1
FDLP_B200_LIB=build_variants/lib_noflush.so python tools/perf_step.py 1000 > gpurun_out/r2_step3_noflush.log 2>&1
