timeout 300 python -m pytest tests/test_mfcc_gpu.py -x -q -m gpu 2>&1 | tail -2
timeout 120 python tools/perf_mfcc.py 2>&1 | tail -3
