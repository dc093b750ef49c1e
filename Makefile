# Builds libdafs_cuda.so (sm_100a only) in-tree, plus the test oracles.
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC,-O2,-pthread,-ffp-contract=off -Xptxas -v --fmad=false
CSRC := dafs_b200/csrc
LIB := dafs_b200/libdafs_cuda.so
OBJS := $(CSRC)/common.o $(CSRC)/pairhmm.o $(CSRC)/dd.o

all: $(LIB) oracle

$(CSRC)/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh) include/dafs_cuda.h
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $@.ptxas.log || (cat $@.ptxas.log; false)

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart_static -lpthread -ldl -lrt

oracle:
	$(MAKE) -C oracle

clean:
	rm -f $(OBJS) $(LIB) $(CSRC)/*.ptxas.log
	$(MAKE) -C oracle clean
.PHONY: all oracle clean
