# Top-level build: libbre.so (CUDA, sm_100a only) + the oracle (test infrastructure).
NVCC ?= /usr/local/cuda/bin/nvcc
PKG := beam-radiance-estimate-pbrt_b200
CSRC := $(PKG)/csrc
OBJDIR ?= build/obj
LIBOUT ?= $(PKG)/libbre.so
NVEXTRA ?=
# -fmad=false: the disk predicate / Morton code / slab test must round exactly like the CPU oracle.
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false \
           -Xcompiler -fPIC,-ffp-contract=off -Iinclude $(NVEXTRA)
SRCS := bre_api.cu bre_build.cu bre_query.cu bre_film.cu bre_trace.cu bre_camera.cu bre_io.cu
OBJS := $(patsubst %.cu,$(OBJDIR)/%.o,$(SRCS)) $(OBJDIR)/bre_nccl.o

.PHONY: all lib oracle ref clean
all: lib oracle
lib: $(LIBOUT)

$(OBJDIR)/%.o: $(CSRC)/%.cu $(CSRC)/bre_internal.cuh $(CSRC)/bre_query_warp.cuh $(CSRC)/bre_query_tile.cuh $(CSRC)/bre_surface.cuh include/bre.h
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; false)

$(OBJDIR)/bre_nccl.o: $(CSRC)/bre_nccl.cpp $(CSRC)/bre_internal.cuh include/bre.h
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

$(LIBOUT): $(OBJS)
	$(NVCC) -shared -o $@ $(OBJS) -lcudart -ldl

oracle:
	$(MAKE) -C oracle oracle
ref:
	$(MAKE) -C oracle ref
clean:
	rm -rf build $(PKG)/libbre.so
	$(MAKE) -C oracle clean
