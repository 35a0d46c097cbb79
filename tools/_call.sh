mkdir -p gpurun_out
timeout 300 python tools/stress_mfcc_tc.py 200 2>&1 | tail -5
