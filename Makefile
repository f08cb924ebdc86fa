# Builds libcudapt.so (the product: sm_100a CUDA + C ABI) and, via oracle/Makefile, the CPU oracles (tests only).
NVCC     ?= /usr/local/cuda/bin/nvcc
PKG      := software-path-tracer_b200
CSRC     := $(PKG)/csrc
# -fmad=false / IEEE div+sqrt / no FTZ: the numerics contract of csrc/common.cuh (bit parity with the CPU path).
NVFLAGS  := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
            -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
            -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v
CU       := $(CSRC)/api.cu $(CSRC)/spheres.cu $(CSRC)/mesh_host.cu $(CSRC)/trace_bvh.cu $(CSRC)/wavefront.cu $(CSRC)/post.cu $(CSRC)/refit.cu $(CSRC)/primary_packet.cu $(CSRC)/lbvh_build.cu
HDR      := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh) include/cudapt.h
OBJ      := $(CU:.cu=.o) $(CSRC)/bvh_build.o $(CSRC)/env_load.o

all: $(PKG)/libcudapt.so oracle host build/cudapt_cli

$(CSRC)/%.o: $(CSRC)/%.cu $(HDR)
	$(NVCC) $(NVFLAGS) -dc -o $@ $< 2> $@.ptxas.log || (cat $@.ptxas.log; false)

$(CSRC)/bvh_build.o: $(CSRC)/bvh_build.cpp $(CSRC)/bvh8.h $(CSRC)/bvh8_refit.h
	$(CXX) -std=c++17 -O2 -fPIC -ffp-contract=off -pthread -c -o $@ $<

$(CSRC)/env_load.o: $(CSRC)/env_load.cpp $(CSRC)/env_load.h
	$(CXX) -std=c++17 -O2 -fPIC -ffp-contract=off -c -o $@ $<

$(PKG)/libcudapt.so: $(OBJ)
	$(NVCC) -gencode arch=compute_100a,code=sm_100a -shared -o $@ $(OBJ) -lpthread

oracle:
	$(MAKE) -C oracle all

# C++ host side: render::PathTracer adapter (product, compiled against the reference's public headers) and the
# headless two-backend driver (test scaffolding; links the Tier-1 oracle for the reference's own TUs).  Needs the
# reference tree, so it is only (re)built where /root/reference is mounted; the binary travels to the GPU box.
REF      ?= /root/reference
REFINC   := -I$(REF)/libs/render/include -I$(REF)/libs/render/src -I$(REF)/libs/render/vendor/glm
host: $(PKG)/libcudapt.so oracle
	@if [ -d $(REF)/libs/render/src ]; then $(MAKE) build/headless_compare; else echo "host: $(REF) absent, keeping prebuilt build/"; fi

build/headless_compare: $(PKG)/host/CUDAPathTracer.cpp $(PKG)/host/CUDAPathTracer.h tests/cpp/headless_compare.cpp include/cudapt.h
	mkdir -p build
	$(CXX) -std=gnu++23 -O2 -w $(REFINC) -Iinclude -I$(PKG)/host -o $@ $(PKG)/host/CUDAPathTracer.cpp tests/cpp/headless_compare.cpp \
	    -L$(PKG) -lcudapt -Loracle/_ref -lref_render -Wl,-rpath,'$$ORIGIN/../$(PKG):$$ORIGIN/../oracle/_ref'

# headless command-line front door, plain C over the C ABI
build/cudapt_cli: tools/cudapt_cli.c include/cudapt.h $(PKG)/libcudapt.so
	mkdir -p build
	$(CC) -std=c99 -O2 -Wall -Iinclude -o $@ tools/cudapt_cli.c -L$(PKG) -lcudapt -Wl,-rpath,'$$ORIGIN/../$(PKG)'

clean:
	rm -f $(CSRC)/*.o $(CSRC)/*.ptxas.log $(PKG)/libcudapt.so
	$(MAKE) -C oracle clean
.PHONY: all oracle host clean
