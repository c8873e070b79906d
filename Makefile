# Builds libspm_b200/lib/libspmb.so (hand-written sm_100a kernels + the C ABI of include/spmb.h)
# and the CPU oracle (oracle/liborc.so, test infrastructure).
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-Wall -shared
CSRC := libspm_b200/csrc
LIB := libspm_b200/lib/libspmb.so
SYNTH := libspm_b200/lib/libspmsynth.so
HOSTCXX := $(shell test -x /usr/bin/g++ && echo /usr/bin/g++ || echo g++)
HDRS := $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.hpp) include/spmb.h

all: $(LIB) $(SYNTH) oracle

$(LIB): $(CSRC)/spmb_capi.cu $(CSRC)/ingest.cpp $(CSRC)/hostpack.cpp $(CSRC)/multi.cpp $(HDRS)
	@mkdir -p libspm_b200/lib
	$(NVCC) $(NVFLAGS) -o $@ $(CSRC)/spmb_capi.cu $(CSRC)/ingest.cpp $(CSRC)/hostpack.cpp $(CSRC)/multi.cpp -lz -lpthread

# the synthetic-input generator shared by the oracle side and the GPU harness (harness code, no CUDA)
$(SYNTH): $(CSRC)/synth.cpp include/spmb_synth.h
	@mkdir -p libspm_b200/lib
	$(HOSTCXX) -O3 -march=x86-64-v3 -ffp-contract=off -fopenmp -fPIC -Wall -Wextra -std=c++17 -shared -o $@ $(CSRC)/synth.cpp

oracle:
	$(MAKE) -C oracle

clean:
	rm -f $(LIB) $(SYNTH)
	$(MAKE) -C oracle clean

.PHONY: all oracle clean
