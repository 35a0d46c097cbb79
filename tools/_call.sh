cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_t2.log
python tools/perf_step.py 1000 > gpurun_out/r2_step2.log 2>&1
