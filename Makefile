# Builds libeagle_b200.so (sm_100a only) in-tree; `make oracle` builds the CPU checker.
NVCC  ?= nvcc
PKG   := eagle-in-llama.cpp_b200
SRC   := $(wildcard $(PKG)/csrc/*.cu)
OBJ   := $(patsubst $(PKG)/csrc/%.cu,build/%.o,$(SRC))
FLAGS := -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -ccbin /usr/bin/g++
LIB   := $(PKG)/libeagle_b200.so

all: $(LIB)
build/%.o: $(PKG)/csrc/%.cu $(wildcard $(PKG)/csrc/*.cuh) include/eagle_b200.h
	@mkdir -p build
	$(NVCC) $(FLAGS) -c $< -o $@
$(LIB): $(OBJ)
	$(NVCC) -shared -o $@ $^ -cudart shared -ccbin /usr/bin/g++
oracle:
	$(MAKE) -C oracle all
clean:
	rm -rf build $(LIB)
.PHONY: all oracle clean
