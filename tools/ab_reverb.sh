#!/bin/bash
# A/B runs of compile-time variants of the reverberation transforms: build each with e.g.
#   nvcc ... -DFDLP_RV_THREADS=256 -DFDLP_RV_MINB=2 -shared -o build_variants/<name>.so speech_recognition_tools_b200/csrc/fdlp_b200.cu
# and point FDLP_B200_LIB at it.  Prints the parity-test verdict and the per-kernel times of every variant found.
for lib in build_variants/*.so speech_recognition_tools_b200/libfdlp_b200.so; do
  [ -f "$lib" ] || continue
  echo "== $lib"
  FDLP_B200_LIB=$PWD/$lib timeout 200 python -m pytest tests/test_reverb_gpu.py -x -q 2>&1 | tail -1
  FDLP_B200_LIB=$PWD/$lib timeout 200 python tools/perf_reverb.py 1000 16000 2048,4096 2>&1 | cut -c100-480
done
