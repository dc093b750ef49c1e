# Builds libdafs_cuda.so (sm_100a only) in-tree, plus the test oracles.
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC,-O2,-pthread,-ffp-contract=off -Xptxas -v --fmad=false
CSRC := dafs_b200/csrc
LIB := dafs_b200/libdafs_cuda.so
OBJS := $(CSRC)/common.o $(CSRC)/pairhmm.o $(CSRC)/dd.o $(CSRC)/mccaskill.o $(CSRC)/pct.o $(CSRC)/prog.o $(CSRC)/comm.o

all: $(LIB) oracle

$(CSRC)/blstar37.inc: dafs_b200/data/blstar37.i32 tools/bin2inc.py
	python3 tools/bin2inc.py $< $@

$(CSRC)/mccaskill.o: $(CSRC)/blstar37.inc

$(CSRC)/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh) include/dafs_cuda.h
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $@.ptxas.log || (cat $@.ptxas.log; false)

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lcudart_static -lpthread -ldl -lrt

oracle: $(LIB)
	$(MAKE) -C oracle

clean:
	rm -f $(OBJS) $(LIB) $(CSRC)/*.ptxas.log $(CSRC)/blstar37.inc
	$(MAKE) -C oracle clean
.PHONY: all oracle clean
