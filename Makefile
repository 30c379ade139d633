# Build everything in-tree.  Artefacts (.so, binaries) are git-ignored but
# travel to the GPU box with the gpurun snapshot.
#
#   make            -> data-path library, preload shim, validator, oracle
#   make ref        -> oracle/_ref harness (mock driver + runner for the
#                      reference libvgpu.so; only meaningful where
#                      /root/reference exists)
PKG      := 4paradigm-k8s-device-plugin_b200
CSRC     := $(PKG)/csrc
NVCC     ?= nvcc
CC       ?= gcc
CXX      ?= g++
ARCH     := -gencode arch=compute_100a,code=sm_100a
NVFLAGS  := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall -Iinclude -I$(CSRC)
CFLAGS   := -O2 -g -fPIC -Wall -Wextra -Iinclude

DP_SO    := $(PKG)/libvgpu_dp.so
SHIM_SO  := $(PKG)/libvgpu.so
SHIM_SRC := $(CSRC)/libvgpu/shim.c $(CSRC)/libvgpu/region.c
ORACLE   := oracle/liboracle.so

VALIDATOR := $(PKG)/vgpuvalidator
all: $(DP_SO) $(SHIM_SO) $(VALIDATOR) $(ORACLE)

$(DP_SO): $(CSRC)/vgpu_datapath.cu $(CSRC)/vgpu_noalloc.cu $(CSRC)/vgpu_kernels.cuh $(CSRC)/vgpu_plugin.cpp $(CSRC)/vgpu_greenctx.cpp $(CSRC)/vgpu_vmem.cpp \
          $(CSRC)/vgpu_internal.h include/vgpu.h
	$(NVCC) $(NVFLAGS) -shared -o $@ $(CSRC)/vgpu_datapath.cu $(CSRC)/vgpu_noalloc.cu $(CSRC)/vgpu_plugin.cpp $(CSRC)/vgpu_greenctx.cpp $(CSRC)/vgpu_vmem.cpp -ldl

# the LD_PRELOAD shim: plain C, no CUDA toolkit needed at run time (driver reached by dlopen)
$(SHIM_SO): $(SHIM_SRC) $(CSRC)/libvgpu/region.h $(CSRC)/libvgpu/shimlog.h
	$(CC) -O2 -g -fPIC -shared -fvisibility=hidden -Wall -Wextra -Wl,-soname,libvgpu.so -o $@ $(SHIM_SRC) -ldl -lpthread

$(VALIDATOR): $(CSRC)/validator/vgpuvalidator.cpp $(DP_SO) include/vgpu.h
	$(CXX) -O2 -g -Wall -Wextra -Iinclude -o $@ $(CSRC)/validator/vgpuvalidator.cpp -L$(PKG) -lvgpu_dp -Wl,-rpath,'$$ORIGIN'

$(ORACLE): oracle/vgpu_oracle.c oracle/vgpu_oracle.h
	$(CC) -O3 -fPIC -shared -Wall -Wextra -o $@ oracle/vgpu_oracle.c -lpthread

clean:
	rm -f $(DP_SO) $(SHIM_SO) $(VALIDATOR) $(ORACLE)

.PHONY: all clean

# oracle/_ref: mock driver + trace runner that execute the *reference* libvgpu.so
# (binary, read from /root/reference at run time; never copied).  See oracle/ref_harness/.
ref:
	@if [ -f oracle/ref_harness/Makefile ]; then $(MAKE) -C oracle/ref_harness; else echo "oracle/ref_harness not present yet"; fi
.PHONY: ref
