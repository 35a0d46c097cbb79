cd $GRAFT_REPO_ROOT
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_na_alone.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_na_all.log
compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r2_na_memcheck.log
