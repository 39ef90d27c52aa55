# Builds the CUDA library (C ABI), the host C++ binaries and the oracle.  nvcc cross-compiles for
# sm_100a without a GPU.  `make` = everything; `make lib` = libdrm_cuda.so only.
NVCC      ?= /usr/local/cuda/bin/nvcc
CXX       ?= /usr/bin/g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fopenmp,-O3 --expt-relaxed-constexpr
CSRC      := deepreadmapper_b200/csrc
HOST      := deepreadmapper_b200/host
LIBDIR    := deepreadmapper_b200/lib
BINDIR    := build/bin
OBJDIR    := build/obj

LIB       := $(LIBDIR)/libdrm_cuda.so
CU_SRCS   := $(CSRC)/api.cu $(CSRC)/search_kernel.cu $(CSRC)/search_exact.cu $(CSRC)/sw_kernel.cu $(CSRC)/merge_kernel.cu $(CSRC)/shards.cu $(CSRC)/index_image.cu $(CSRC)/build_gpu.cu $(CSRC)/input_kernels.cu $(CSRC)/align_kernel.cu $(CSRC)/embed_kernel.cu $(CSRC)/l2_kernel.cu
CPP_SRCS  := $(CSRC)/index_io.cpp $(CSRC)/builder.cpp $(CSRC)/ir_model.cpp $(CSRC)/l2_plan.cpp
CU_OBJS   := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(CU_SRCS))
CPP_OBJS  := $(patsubst $(CSRC)/%.cpp,$(OBJDIR)/%.o,$(CPP_SRCS))
EXTRA_OBJS :=
-include $(CSRC)/extra.mk

all: lib host oracle

lib: $(LIB)

$(OBJDIR)/%.o: $(CSRC)/%.cu $(CSRC)/common.h include/drm_b200.h
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; false)

$(OBJDIR)/%.o: $(CSRC)/%.cpp $(CSRC)/common.h include/drm_b200.h
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(CU_OBJS) $(CPP_OBJS) $(EXTRA_OBJS)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(ARCH) -shared -o $@ $^ -Xcompiler -fopenmp -lcudart_static -lpthread -ldl -lrt -lgomp

host: lib
	@if [ -f $(HOST)/Makefile ]; then $(MAKE) -C $(HOST); fi

oracle:
	$(MAKE) -s -f oracle/Makefile all

clean:
	rm -rf build $(LIBDIR)/*.so
.PHONY: all lib host oracle clean
