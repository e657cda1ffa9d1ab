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
.PHONY: all clean ptxas cpptest

# C++20 host layer check: the reference's OIIO-free test cases on include/compressed/*.h + libb2g.so
CPPTEST := tests/cpp/test_compressed
cpptest: $(CPPTEST)
# nlohmann/json.hpp (image<T>::metadata) from wherever this image has one; without it the accessors are left out
JSON_INC := $(firstword $(wildcard /opt/prime-rl/.venv/lib/python3.12/site-packages/include/cudnn_frontend/thirdparty /root/reference/thirdparty/json/include))
$(CPPTEST): tests/cpp/test_compressed.cpp $(wildcard include/compressed/*.h include/compressed/*/*.h) include/b2g.h $(OUT)
	g++ -std=c++20 -O2 -Wall -Wextra -Iinclude $(if $(JSON_INC),-I$(JSON_INC)) -o $@ tests/cpp/test_compressed.cpp -Lcompressed_image_b200 -lb2g -Wl,-rpath,'$$ORIGIN/../../compressed_image_b200'
