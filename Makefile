# Builds libocc_b200.so (sm_100a only) in-tree; the .so is git-ignored but travels to the GPU box.
NVCC ?= /usr/local/cuda/bin/nvcc
ARCH := -gencode arch=compute_100a,code=sm_100a
CFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC -Xcompiler -Wall -Xcompiler -Wno-unused-function
SRC := $(wildcard occm_b200/csrc/*.cu)
OBJ := $(patsubst occm_b200/csrc/%.cu,build/%.o,$(SRC))
LIB := occm_b200/libocc_b200.so

all: $(LIB)

build/%.o: occm_b200/csrc/%.cu $(wildcard occm_b200/csrc/*.h occm_b200/csrc/*.cuh include/*.h)
	@mkdir -p build
	$(NVCC) $(CFLAGS) -c $< -o $@

$(LIB): $(OBJ)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJ) -lcudart -ldl

clean:
	rm -rf build $(LIB)
.PHONY: all clean
