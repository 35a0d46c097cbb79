for v in tchint tchintcl; do
FDLP_B200_LIB=build_variants/lib_$v.so timeout 300 python -m pytest tests/test_mfcc_gpu.py -x -q -m gpu 2>&1 | tail -1
FDLP_B200_LIB=build_variants/lib_$v.so timeout 120 python tools/perf_mfcc.py 2>&1 | head -1
done
timeout 120 python tools/perf_mfcc.py 2>&1 | head -1
