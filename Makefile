# Build libbibfs_b200.so (sm_100a only) and the native `bibfs` CLI, in-tree.
#   make            -> bidirectional_bfs_b200/_native/libbibfs_b200.so + bin/bibfs
#   make sass       -> docs/sass/*.sass listings (cuobjdump)
#   make run / make test -> smoke queries on the 1k preset / the CPU test-suite
#   make clean
#   make EXTRA=-DBIBFS_FINE_TRACE BUILD=build_ft OUTDIR=build_ft/lib   -> instrumented copy (load it with BIBFS_B200_LIB=...)
# Counterpart of the reference's per-variant Makefiles (v1..v4/Makefile); the reference's SMS list
# (sm_50..sm_90, v3/Makefile:16-18) is replaced by the single Blackwell target.
NVCC      ?= /usr/local/cuda/bin/nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := $(EXTRA) $(ARCH) -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-O3,-Wall,-Wno-unknown-pragmas \
             -Xptxas -v --expt-relaxed-constexpr -Iinclude
CXXFLAGS  := -O3 -std=c++17 -fPIC -Wall -Wno-unknown-pragmas -ffp-contract=off -Iinclude -pthread
BUILD     ?= build
OUTDIR    ?= bidirectional_bfs_b200/_native
LIB       := $(OUTDIR)/libbibfs_b200.so
CLI       := bin/bibfs

CU_SRCS   := csrc/gen_dev.cu csrc/csr_dev.cu csrc/bfs_dev.cu csrc/ctx.cu csrc/naive_dev.cu csrc/msbfs_dev.cu csrc/scan_noalloc.cu csrc/vrec_dev.cu
CPP_SRCS  := csrc/error.cpp csrc/host_io.cpp csrc/host_graph.cpp csrc/cpu_serial.cpp csrc/cpu_ranks.cpp
CU_OBJS   := $(patsubst csrc/%.cu,$(BUILD)/%.o,$(CU_SRCS))
CPP_OBJS  := $(patsubst csrc/%.cpp,$(BUILD)/%.o,$(CPP_SRCS))
HDRS      := $(wildcard csrc/*.h csrc/*.cuh include/*.h)

PREP      := bin/bench_prep

all: $(LIB) $(CLI) $(PREP)

# stand-alone helper of bench.py (graph file, query list, v1-schedule numerator) — links nothing of the library
$(PREP): baseline/tools/bench_prep.cpp csrc/gen.h csrc/rng.h
	@mkdir -p bin
	$(CXX) -O3 -std=c++17 -pthread -Iinclude $< -o $@

$(BUILD)/%.o: csrc/%.cu $(HDRS)
	@mkdir -p $(BUILD)
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> $(BUILD)/$*.ptxas.log || (cat $(BUILD)/$*.ptxas.log; exit 1)

$(BUILD)/%.o: csrc/%.cpp $(HDRS)
	@mkdir -p $(BUILD)
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(LIB): $(CU_OBJS) $(CPP_OBJS)
	@mkdir -p $(OUTDIR)
	$(NVCC) $(ARCH) -shared -o $@ $^ -lcudart -lpthread

$(CLI): csrc/cli_main.cpp $(LIB) $(HDRS)
	@mkdir -p bin
	$(CXX) $(CXXFLAGS) -I/usr/local/cuda/include csrc/cli_main.cpp -o $@ -L$(OUTDIR) -lbibfs_b200 -Wl,-rpath,'$$ORIGIN/../$(OUTDIR)' -L/usr/local/cuda/lib64 -Wl,-rpath,/usr/local/cuda/lib64 -lcudart

sass: $(LIB)
	@mkdir -p docs/sass
	/usr/local/cuda/bin/cuobjdump -sass $(LIB) > docs/sass/libbibfs_b200.sass

# smoke targets like the reference's `make run` (v2/Makefile:31-32, v3/Makefile:46-47, v4/Makefile:63-64): the 1k preset, query 0 -> 999
graphs/1k.bin:
	python -m bidirectional_bfs_b200 make-graphs --dir graphs

run: all graphs/1k.bin
	$(CLI) graphs/1k.bin 0 999 --variant serial
	$(CLI) graphs/1k.bin 0 999 --variant cpu-ranks --ranks 4
	@if nvidia-smi -L > /dev/null 2>&1; then $(CLI) graphs/1k.bin 0 999 --variant gpu; else echo "(no GPU: gpu variant skipped)"; fi

test: all
	python -m pytest tests -q -m "not gpu"

clean:
	rm -rf $(BUILD) $(LIB) $(CLI)

.PHONY: all clean sass run test
