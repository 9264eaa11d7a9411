# Builds libshifu_b200.so (sm_100a only), the CPU oracle and the bare-C-ABI host program.  nvcc cross-compiles
# without a GPU.  One object per source so that a change rebuilds one file; `make -j` compiles them side by side.
NVCC ?= /usr/local/cuda/bin/nvcc
NVCCFLAGS = -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC,-fvisibility=hidden
CSRC = shifu_b200/csrc
SRCS = api.cu rowlist.cu hist.cu rows.cu split.cu noalloc.cu prep.cu score.cu p2p.cu host_master.cpp wire.cpp
OBJS = $(addprefix build/,$(addsuffix .o,$(basename $(SRCS))))
HDRS = $(CSRC)/common.cuh $(CSRC)/kernels.cuh include/shifu_b200.h

all: shifu_b200/libshifu_b200.so oracle tools/train_gbt

build/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -c -o $@ $<

build/%.o: $(CSRC)/%.cpp $(HDRS)
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -c -o $@ $<

shifu_b200/libshifu_b200.so: $(OBJS)
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -shared -o $@ $(OBJS)

oracle:
	$(MAKE) -s -C oracle

tools/train_gbt: tools/train_gbt.cpp shifu_b200/libshifu_b200.so include/shifu_b200.h
	g++ -O2 -std=c++17 -o $@ tools/train_gbt.cpp -Lshifu_b200 -lshifu_b200 -Wl,-rpath,'$$ORIGIN/../shifu_b200'

clean:
	rm -rf build shifu_b200/libshifu_b200.so tools/train_gbt

.PHONY: all oracle clean
