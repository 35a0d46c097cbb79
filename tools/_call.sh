cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests -m gpu -x -q -k "mel" 2>&1 | tail -5 > gpurun_out/r2_t14.log
timeout 120 python tools/perf_mel.py > gpurun_out/r2_mel14.log 2>&1
