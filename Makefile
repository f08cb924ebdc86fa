# Builds libcudapt.so (the product: sm_100a CUDA + C ABI) and, via oracle/Makefile, the CPU oracles (tests only).
NVCC     ?= /usr/local/cuda/bin/nvcc
PKG      := software-path-tracer_b200
CSRC     := $(PKG)/csrc
# -fmad=false / IEEE div+sqrt / no FTZ: the numerics contract of csrc/common.cuh (bit parity with the CPU path).
NVFLAGS  := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
            -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
            -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v
CU       := $(CSRC)/api.cu $(CSRC)/spheres.cu $(CSRC)/mesh_host.cu $(CSRC)/trace_bvh.cu $(CSRC)/wavefront.cu
HDR      := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh) include/cudapt.h
OBJ      := $(CU:.cu=.o) $(CSRC)/bvh_build.o

all: $(PKG)/libcudapt.so oracle

$(CSRC)/%.o: $(CSRC)/%.cu $(HDR)
	$(NVCC) $(NVFLAGS) -dc -o $@ $< 2> $@.ptxas.log || (cat $@.ptxas.log; false)

$(CSRC)/bvh_build.o: $(CSRC)/bvh_build.cpp $(CSRC)/bvh8.h
	$(CXX) -std=c++17 -O2 -fPIC -ffp-contract=off -pthread -c -o $@ $<

$(PKG)/libcudapt.so: $(OBJ)
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -shared -o $@ $(OBJ) -lpthread

oracle:
	$(MAKE) -C oracle all

clean:
	rm -f $(CSRC)/*.o $(CSRC)/*.ptxas.log $(PKG)/libcudapt.so
	$(MAKE) -C oracle clean
.PHONY: all oracle clean
