# Builds libb200img.so (C ABI: host reference + sm_100a kernels + multi-GPU runtime) and the
# b200img command-line binary. sm_100a only: no multi-arch fat binaries, no fallbacks.
#
#   make            library + CLI
#   make baseline   shim-built copy of the unmodified reference (baseline/build_ref.sh)
#   make sass       SASS listings of the hot kernels into profiles/sass/
NVCC      ?= /usr/local/cuda/bin/nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
CXXFLAGS  := -O3 -std=c++17 -fPIC -Wall -Wextra -Iinclude
NVCCFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -Iinclude -Xcompiler -fPIC,-Wall,-Wextra \
             --expt-relaxed-constexpr -Xptxas -v
PKG       := mpi_cuda_imagemanipulation_b200
LIB       := $(PKG)/lib/libb200img.so
CLI       := bin/b200img
OBJDIR    := build/obj

CPP_SRCS := $(wildcard csrc/core/*.cpp csrc/host_ref/*.cpp csrc/runtime/*.cpp)
CU_SRCS  := $(wildcard csrc/kernels/*.cu csrc/runtime/*.cu)
CPP_OBJS := $(patsubst csrc/%.cpp,$(OBJDIR)/%.o,$(CPP_SRCS))
CU_OBJS  := $(patsubst csrc/%.cu,$(OBJDIR)/%.cu.o,$(CU_SRCS))
HDRS     := $(wildcard include/*.h csrc/*/*.hpp csrc/*/*.cuh)

all: $(LIB) $(CLI)

$(OBJDIR)/%.o: csrc/%.cpp $(HDRS)
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -I/usr/local/cuda/include -c $< -o $@

$(OBJDIR)/%.cu.o: csrc/%.cu $(HDRS)
	@mkdir -p $(dir $@)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $@.ptxas.log || (cat $@.ptxas.log; false)
	@grep -E "error|warning|spill|registers" $@.ptxas.log | grep -v "0 bytes spill" | head -60 || true

$(LIB): $(CPP_OBJS) $(CU_OBJS)
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -shared -o $@ $^ -cudart static -lpthread -ldl -lrt

$(CLI): csrc/cli/main.cpp $(LIB) $(HDRS)
	@mkdir -p $(dir $@)
	$(CXX) $(CXXFLAGS) -o $@ csrc/cli/main.cpp -L$(PKG)/lib -lb200img \
	    -Wl,-rpath,'$$ORIGIN/../$(PKG)/lib' -lpthread

baseline:
	bash baseline/build_ref.sh

sass: $(LIB)
	@mkdir -p profiles/sass
	cuobjdump -sass $(LIB) > profiles/sass/libb200img.sass

clean:
	rm -rf build bin $(PKG)/lib/*.so

.PHONY: all baseline sass clean
