# Builds the B200-native chunk engine (C ABI in include/b2g.h) for sm_100a, in-tree.
NVCC  ?= nvcc
CSRC  := compressed_image_b200/csrc
OUT   := compressed_image_b200/libb2g.so
FLAGS := -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC,-fvisibility=hidden -shared

all: $(OUT)

$(OUT): $(CSRC)/b2g_api.cu $(wildcard $(CSRC)/*.cuh) $(CSRC)/b2g_plan.h include/b2g.h
	$(NVCC) $(FLAGS) -o $@ $(CSRC)/b2g_api.cu

ptxas: 
	$(NVCC) $(FLAGS) -Xptxas -v -o /tmp/b2g_ptxas.so $(CSRC)/b2g_api.cu

clean:
	rm -f $(OUT)
.PHONY: all clean ptxas
