# Builds libspm_b200/lib/libspmb.so (hand-written sm_100a kernels + the C ABI of include/spmb.h)
# and the CPU oracle (oracle/liborc.so, test infrastructure).
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-Wall -shared
CSRC := libspm_b200/csrc
LIB := libspm_b200/lib/libspmb.so
HDRS := $(wildcard $(CSRC)/*.cuh) include/spmb.h

all: $(LIB) oracle

$(LIB): $(CSRC)/spmb_capi.cu $(CSRC)/ingest.cpp $(CSRC)/hostpack.cpp $(HDRS)
	@mkdir -p libspm_b200/lib
	$(NVCC) $(NVFLAGS) -o $@ $(CSRC)/spmb_capi.cu $(CSRC)/ingest.cpp $(CSRC)/hostpack.cpp -lz -lpthread

oracle:
	$(MAKE) -C oracle

clean:
	rm -f $(LIB)
	$(MAKE) -C oracle clean

.PHONY: all oracle clean
