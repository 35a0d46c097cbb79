# Builds the C-ABI CUDA library in-tree (sm_100a only) and the oracle helpers.
NVCC ?= nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := -O3 -std=c++17 $(ARCH) -lineinfo -Xcompiler -fPIC -Xptxas -v --expt-relaxed-constexpr
PKG := speech_recognition_tools_b200
SRC := $(PKG)/csrc/fdlp_b200.cu
HDR := $(wildcard $(PKG)/csrc/*.cuh) include/fdlp_b200.h
LIB := $(PKG)/libfdlp_b200.so

all: $(LIB) tools/fp64_peak

$(LIB): $(SRC) $(HDR)
	$(NVCC) $(NVFLAGS) -shared -o $@ $(SRC) 2> build_ptxas.log || (cat build_ptxas.log; exit 1)

tools/fp64_peak: tools/fp64_peak.cu
	$(NVCC) -O3 $(ARCH) -lineinfo -o $@ $<

clean:
	rm -f $(LIB) tools/fp64_peak build_ptxas.log
