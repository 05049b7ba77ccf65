# Build of the B200-native solver. Same verbs as the reference makefile (makefile:1-15:
# build / clean / run / runOn2) but: optimised host code, explicit sm_100a device code with
# line info, no MPI, no hard-coded CUDA 11.0 path.
CUDA_HOME ?= /usr/local/cuda
NVCC      ?= $(CUDA_HOME)/bin/nvcc
CXX       ?= g++
PKG       := mpi_openmp_cuda_b200
LIBDIR    := $(PKG)/_lib
GENCODE   := -gencode arch=compute_100a,code=[sm_100a,compute_100a]
CXXFLAGS  := -O3 -std=c++17 -fPIC -Wall -Wextra -Iinclude -pthread
NVCCFLAGS := -O3 -std=c++17 $(GENCODE) -lineinfo -Iinclude -Xcompiler -fPIC -Xptxas -v
INPUT     ?= data/input1.txt
GPUS      ?= 1

CORE_SRC := csrc/core/problem.cpp csrc/core/oracle.cpp csrc/core/plan.cpp
CUDA_SRC := csrc/cuda/align_kernel.cu csrc/cuda/fabric_bench.cu csrc/cuda/native_fabric.cu
CORE_LIB := $(LIBDIR)/libseqalign_core.so
CUDA_LIB := $(LIBDIR)/libseqalign_cuda.so

.PHONY: build core cuda clean run runOn2 test sass asan tsan microbench tc_scan startup_probe cluster_probe

build: core cuda final

core: $(CORE_LIB)
cuda: $(CUDA_LIB)

$(CORE_LIB): $(CORE_SRC) csrc/core/core.hpp include/seqalign.h
	@mkdir -p $(LIBDIR)
	$(CXX) $(CXXFLAGS) -shared -o $@ $(CORE_SRC)

$(CUDA_LIB): $(CUDA_SRC) $(wildcard csrc/cuda/*.cuh) csrc/host/native_wire.h include/seqalign.h $(CORE_LIB)
	@mkdir -p $(LIBDIR)
	$(NVCC) $(NVCCFLAGS) -shared -o $@ $(CUDA_SRC) -L$(LIBDIR) -lseqalign_core -Xlinker -rpath -Xlinker '$$ORIGIN'

# `final` keeps the reference's binary name and stdin->stdout contract (makefile:4,11).
final: csrc/host/final_main.cpp csrc/host/native_launch.cpp csrc/host/native_wire.h $(CORE_LIB) $(CUDA_LIB)
	$(CXX) $(CXXFLAGS) -o $@ csrc/host/final_main.cpp csrc/host/native_launch.cpp -L$(LIBDIR) -lseqalign_core -ldl \
	    -Wl,-rpath,'$$ORIGIN/$(LIBDIR)'

# Pipe-rate micro-benchmarks (LDS / IADD3 / IMAD / VIMNMX / tcgen05.ld): profiles/MEASURED_PIPES.json
microbench: build/microbench
build/microbench: csrc/cuda/microbench.cu
	@mkdir -p build
	$(NVCC) -O3 -std=c++17 $(GENCODE) -lineinfo -o $@ $<

# tcgen05 / TMEM variant of the scan as a stand-alone measured experiment: profiles/tc_scan.json
tc_scan: build/tc_scan
build/tc_scan: csrc/cuda/tc_scan.cu csrc/cuda/ptx_helpers.cuh
	@mkdir -p build
	$(NVCC) -O3 -std=c++17 $(GENCODE) -lineinfo -Xptxas -v -o $@ $<

# Cluster launch + multicast bulk copy for the TILE start-up, measured: profiles/cluster_probe.json
cluster_probe: build/cluster_probe
build/cluster_probe: csrc/cuda/cluster_probe.cu csrc/cuda/ptx_helpers.cuh
	@mkdir -p build
	$(NVCC) -O3 -std=c++17 $(GENCODE) -lineinfo -o $@ $<

# Cold-start breakdown of a CUDA process (driver init / context / module load): profiles/cold_start.txt
startup_probe: build/startup_probe
build/startup_probe: csrc/cuda/startup_probe.cu
	@mkdir -p build
	$(NVCC) -O3 -std=c++17 $(GENCODE) -o $@ $<

clean:
	rm -f final $(LIBDIR)/*.so csrc/**/*.o

run: build
	./final --gpus $(GPUS) < $(INPUT)

# The reference's runOn2 target spreads 2 ranks over 2 hosts (makefile:15); a single
# 8xB200 box is the deployment unit here, so the analogue is 2 GPUs of one node.
runOn2: build
	./final --gpus 2 < $(INPUT)

test: build
	python -m pytest tests -x -q -m "not gpu"

sass: cuda
	$(CUDA_HOME)/bin/cuobjdump -sass $(CUDA_LIB) > profiles/align_kernel.sass

# Host-side sanitizer runs: the whole core library + tests/native/core_selftest.cpp
# under ASan+UBSan, and under TSan for the multi-threaded packer and oracle.
SAN_CXX ?= $(shell test -x /usr/bin/g++ && echo /usr/bin/g++ || echo $(CXX))  # a compiler that ships the sanitizer runtimes
SELFTEST_INPUTS := data/input1.txt data/input2.txt data/input5.txt data/input6.txt data/input4.txt
asan:
	@mkdir -p build
	$(SAN_CXX) -O1 -g -std=c++17 -Iinclude -pthread -fsanitize=address,undefined -fno-omit-frame-pointer -fno-sanitize-recover=undefined \
	    -o build/core_selftest_asan tests/native/core_selftest.cpp $(CORE_SRC)
	ASAN_OPTIONS=detect_leaks=1 build/core_selftest_asan $(SELFTEST_INPUTS)
tsan:
	@mkdir -p build
	$(SAN_CXX) -O1 -g -std=c++17 -Iinclude -pthread -fsanitize=thread -o build/core_selftest_tsan tests/native/core_selftest.cpp $(CORE_SRC)
	build/core_selftest_tsan data/input1.txt data/input6.txt
