# nbody-b200 -- builds the product library in-tree (the .so travels to the GPU box with gpurun).
#
#   make            nbody_b200/lib/libnbody.so      World API + MakeGalaxies + CUDA pipeline (sm_100a)
#   make bench      nbody_b200/bin/nbody-bench      C benchmark CLI (the reference's src/bench.c, extended)
#   make oracle     oracle/_build, oracle/_ref      the checkers (test infrastructure)
#   make tune       tools/bin/tune                  kernel-variant / pipe micro-benchmarks
#
# sm_100a only: no other -gencode, no PTX for other architectures, no CPU fallback.

NVCC    := /usr/local/cuda/bin/nvcc
# the environment may export CC=/opt/gcc/bin/gcc (a wrapper without libgomp); use the system compilers
CC      := /usr/bin/gcc
CXX     := /usr/bin/g++

ARCH    := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -lineinfo -std=c++17 -ccbin $(CXX) -Xcompiler -fPIC -Xptxas -v -Iinclude -Inbody_b200/csrc
# -mavx like the reference's default SIMD_SET=AVX; -ffp-contract=off keeps host float math FMA-free;
# -fno-math-errno lets gcc vectorise sqrtf (vsqrtps) in UpdateWorld_CPU without changing any rounding
CFLAGS  := -O3 -std=gnu11 -fPIC -fopenmp -mavx -ffp-contract=off -fno-fast-math -fno-math-errno -Wall -Wextra -Iinclude

SRC     := nbody_b200/csrc
OBJ     := build
LIB     := nbody_b200/lib/libnbody.so
CUH     := $(wildcard $(SRC)/*.cuh) $(wildcard $(SRC)/*.h) $(wildcard include/*.h)

.PHONY: all bench oracle tune clean

all: $(LIB)

$(OBJ)/sim_cuda.o: $(SRC)/sim_cuda.cu $(CUH)
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(OBJ)/sim_cuda.ptxas.log || (cat $(OBJ)/sim_cuda.ptxas.log; false)

$(OBJ)/noalloc.o: $(SRC)/noalloc.cu $(SRC)/force_kernel.cuh $(wildcard include/*.h)
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $(OBJ)/noalloc.ptxas.log || (cat $(OBJ)/noalloc.ptxas.log; false)

$(OBJ)/%.o: $(SRC)/%.c $(wildcard include/*.h)
	@mkdir -p $(OBJ)
	$(CC) $(CFLAGS) -c $< -o $@

$(LIB): $(OBJ)/sim_cuda.o $(OBJ)/noalloc.o $(OBJ)/world.o $(OBJ)/galaxy.o
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -ccbin $(CXX) -shared -cudart static -o $@ $^ -Xlinker -Bsymbolic -lgomp -ldl -lm

bench: nbody_b200/bin/nbody-bench

nbody_b200/bin/nbody-bench: $(SRC)/bench_cli.c $(LIB)
	@mkdir -p $(dir $@)
	$(CC) $(CFLAGS) -o $@ $(SRC)/bench_cli.c -Lnbody_b200/lib -lnbody -Wl,-rpath,'$$ORIGIN/../lib' -lm

oracle: $(LIB)
	$(MAKE) -C oracle all
	@if [ -d /root/reference/src/lib ]; then $(MAKE) -C oracle ref; fi

tune: tools/bin/tune

tools/bin/tune: tools/tune.cu $(CUH)
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -O3 -lineinfo -std=c++17 -ccbin $(CXX) -Iinclude -I$(SRC) -o $@ tools/tune.cu

clean:
	rm -rf $(OBJ) nbody_b200/lib nbody_b200/bin tools/bin
