# Builds the product library (hand-written sm_100a CUDA + host C++) in-tree:
#   lightgbmv1_b200/lib_lightgbm_b200.so    the drop-in C-ABI library (LGBM_* subset + lgbcu_* layer)
# and, as separate targets, the test-only oracle (see oracle/Makefile).
NVCC     ?= /usr/local/cuda/bin/nvcc
CXX      := $(shell command -v /usr/bin/g++ || echo g++)
ARCH     := -gencode arch=compute_100a,code=sm_100a
NVFLAGS  := $(ARCH) -lineinfo -O3 --fmad=false -std=c++17 -Xcompiler -fPIC -Xcompiler -fopenmp -Iinclude -ccbin $(CXX)
CXXFLAGS := -std=c++17 -O2 -fPIC -fvisibility=hidden -fopenmp -Wall -Wextra -Wno-unused-parameter -Iinclude
PKG      := lightgbmv1_b200
CUDA_SRC := $(PKG)/csrc/cuda/lgbcu.cu
CUDA_DEP := $(wildcard $(PKG)/csrc/cuda/*.cuh) include/lgbcu.h
HOST_SRC := $(wildcard $(PKG)/csrc/host/*.cpp)
HOST_DEP := $(wildcard $(PKG)/csrc/host/*.h) include/lgbcu.h include/lgbm_b200.h
HOST_OBJ := $(patsubst $(PKG)/csrc/host/%.cpp,build/host_%.o,$(HOST_SRC))
LIB      := $(PKG)/lib_lightgbm_b200.so

.PHONY: all lib oracle ref clean
all: lib oracle

lib: $(LIB)

build/lgbcu.o: $(CUDA_SRC) $(CUDA_DEP)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -c $(CUDA_SRC) -o $@

build/host_%.o: $(PKG)/csrc/host/%.cpp $(HOST_DEP)
	@mkdir -p build
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(LIB): build/lgbcu.o $(HOST_OBJ)
	$(NVCC) $(ARCH) -shared -ccbin $(CXX) -Xcompiler -fopenmp -o $@ build/lgbcu.o $(HOST_OBJ) -ldl

oracle:
	$(MAKE) -C oracle port

ref:
	$(MAKE) -C oracle ref

clean:
	rm -rf build $(LIB)
	$(MAKE) -C oracle clean
