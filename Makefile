# Builds libnbody_b200.so (sm_100a only), the host drop-in and the test-only oracle libraries.
NVCC   ?= /usr/local/cuda/bin/nvcc
PKG    := n-body-simulation_b200
CSRC   := $(PKG)/csrc
LIB    := $(PKG)/libnbody_b200.so
ARCH   := -gencode arch=compute_100a,code=sm_100a
NVFLAGS:= $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Iinclude -I$(CSRC) --expt-relaxed-constexpr

CU_SRCS := $(CSRC)/nb_force.cu $(CSRC)/nb_integrate.cu $(CSRC)/nb_sort.cu $(CSRC)/nb_context.cu $(CSRC)/nb_noalloc.cu $(CSRC)/nb_render.cu
CU_OBJS := $(CU_SRCS:.cu=.o)

.PHONY: all lib oracle ref refgpu refcuda tools host clean
all: lib oracle ref tools host refgpu refcuda

lib: $(LIB)
$(CSRC)/%.o: $(CSRC)/%.cu $(CSRC)/nb_common.cuh include/nbody_b200.h
	$(NVCC) $(NVFLAGS) -Xptxas -v -c -o $@ $< 2> $@.ptxas.log || (cat $@.ptxas.log; exit 1)
$(LIB): $(CU_OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(CU_OBJS) -ldl

oracle:
	$(MAKE) -C oracle oracle
ref:
	$(MAKE) -C oracle ref
refgpu: lib
	$(MAKE) -C oracle refgpu
refcuda:
	$(MAKE) -C oracle refcuda

tools: tools/pipe_bench
tools/pipe_bench: tools/pipe_bench.cu
	$(NVCC) $(ARCH) -O3 -std=c++17 -lineinfo -o $@ $<

host: lib
	$(MAKE) -C $(PKG)/host

clean:
	rm -f $(CSRC)/*.o $(CSRC)/*.ptxas.log $(LIB) tools/pipe_bench
	$(MAKE) -C oracle clean
