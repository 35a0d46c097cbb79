#!/bin/bash
# build_variant.sh <name> [-DFLAG ...]: a compile-time variant of the library for A/B runs (FDLP_B200_LIB=build_variants/lib_<name>.so)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
mkdir -p build_variants
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr "$@" \
  -shared -o build_variants/lib_$name.so speech_recognition_tools_b200/csrc/fdlp_b200.cu
echo built build_variants/lib_$name.so
