# Builds libcudapt.so (the product: sm_100a CUDA + C ABI) and, via oracle/Makefile, the CPU oracles (tests only).
NVCC     ?= /usr/local/cuda/bin/nvcc
PKG      := software-path-tracer_b200
CSRC     := $(PKG)/csrc
# -fmad=false / IEEE div+sqrt / no FTZ: the numerics contract of csrc/common.cuh (bit parity with the CPU path).
NVFLAGS  := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
            -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
            -Xcompiler -fPIC,-O2,-ffp-contract=off -Xptxas -v
CU       := $(CSRC)/api.cu $(CSRC)/spheres.cu $(CSRC)/mesh_host.cu $(CSRC)/trace_bvh.cu $(CSRC)/wavefront.cu $(CSRC)/post.cu $(CSRC)/refit.cu $(CSRC)/primary_packet.cu $(CSRC)/lbvh_build.cu $(CSRC)/static_scratch.cu
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

# The registered factory: copies of the reference's PathTracer.h / PathTracer.cpp with the GPU_CUDA
# enumerator and case, made by applying host/reference_glue/register_gpu_cuda.patch; the adapter is copied to where
# INTEGRATION.md puts it in the reference tree.  build/ is git-ignored: the repo carries the patch, not the copies.
GLUE     := build/reference_glue
$(GLUE)/PathTracer.cpp: $(PKG)/host/reference_glue/register_gpu_cuda.patch $(PKG)/host/CUDAPathTracer.h $(PKG)/host/CUDAPathTracer.cpp
	rm -rf $(GLUE) && mkdir -p $(GLUE)/render $(GLUE)/engines/pathtracer/backends/cuda $(GLUE)/tree/libs/render/include/render $(GLUE)/tree/libs/render/src/engines/pathtracer
	cp $(REF)/libs/render/include/render/PathTracer.h $(GLUE)/tree/libs/render/include/render/
	cp $(REF)/libs/render/src/engines/pathtracer/PathTracer.cpp $(GLUE)/tree/libs/render/src/engines/pathtracer/
	cd $(GLUE)/tree && patch -p1 --no-backup-if-mismatch < ../../../$(PKG)/host/reference_glue/register_gpu_cuda.patch
	mv $(GLUE)/tree/libs/render/include/render/PathTracer.h $(GLUE)/render/PathTracer.h
	mv $(GLUE)/tree/libs/render/src/engines/pathtracer/PathTracer.cpp $(GLUE)/PathTracer.cpp
	rm -rf $(GLUE)/tree
	cp $(PKG)/host/CUDAPathTracer.h $(PKG)/host/CUDAPathTracer.cpp $(GLUE)/engines/pathtracer/backends/cuda/

# -I$(GLUE) comes FIRST: "render/PathTracer.h" resolves to the patched copy for every TU, the reference's own headers
# included; the executable's create_path_tracer (patched) takes precedence over the one inside libref_render.so.
build/headless_compare: $(GLUE)/PathTracer.cpp tests/cpp/headless_compare.cpp include/cudapt.h
	mkdir -p build
	$(CXX) -std=gnu++23 -O2 -w -I$(GLUE) $(REFINC) -Iinclude -o $@ $(GLUE)/PathTracer.cpp \
	    $(GLUE)/engines/pathtracer/backends/cuda/CUDAPathTracer.cpp tests/cpp/headless_compare.cpp \
	    -L$(PKG) -lcudapt -Loracle/_ref -lref_render -Wl,-rpath,'$$ORIGIN/../$(PKG):$$ORIGIN/../oracle/_ref'

# headless command-line front door, plain C over the C ABI
build/cudapt_cli: tools/cudapt_cli.c include/cudapt.h $(PKG)/libcudapt.so
	mkdir -p build
	$(CC) -std=c99 -O2 -Wall -Iinclude -o $@ tools/cudapt_cli.c -L$(PKG) -lcudapt -Wl,-rpath,'$$ORIGIN/../$(PKG)'

clean:
	rm -f $(CSRC)/*.o $(CSRC)/*.ptxas.log $(PKG)/libcudapt.so
	$(MAKE) -C oracle clean
.PHONY: all oracle host clean
