# Builds the engine IN-TREE: gffx_b200/lib/libgffx_b200.so (C ABI of include/gffx_b200.h) and the
# gffx_b200/bin/gffx-b200 CLI, for sm_100a only; plus the CPU oracle (oracle/_build/).
NVCC ?= /usr/local/cuda/bin/nvcc
HOSTCXX := $(if $(wildcard /usr/bin/g++),/usr/bin/g++,g++)
ARCH := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -ccbin $(HOSTCXX) -Xcompiler -fPIC,-Wall,-Wno-unused-parameter --expt-relaxed-constexpr $(NVEXTRA)
CXXFLAGS := -O3 -std=c++17 -fPIC -Wall -Wextra -Wno-unused-parameter -pthread
SRC := gffx_b200/csrc
OBJ := build/obj
LIB := gffx_b200/lib/libgffx_b200.so
BIN := gffx_b200/bin/gffx-b200
SYNTH_LIB := gffx_b200/lib/libgffx_synth.so

CU_SRCS := $(SRC)/engine.cu $(SRC)/counts.cu $(SRC)/bedparse.cu $(SRC)/linescan.cu
CPP_SRCS := $(SRC)/util.cpp $(SRC)/winjoin_build.cpp $(SRC)/counts_index.cpp $(SRC)/host_index.cpp $(SRC)/linecache.cpp $(SRC)/reads.cpp $(SRC)/index_builder.cpp $(SRC)/synth.cpp $(SRC)/capi.cpp
CU_OBJS := $(patsubst $(SRC)/%.cu,$(OBJ)/%.cu.o,$(CU_SRCS))
CPP_OBJS := $(patsubst $(SRC)/%.cpp,$(OBJ)/%.o,$(CPP_SRCS))
HDRS := $(wildcard $(SRC)/*.h $(SRC)/*.cuh) include/gffx_b200.h

all: $(LIB) $(BIN) $(SYNTH_LIB) oracle

$(OBJ)/%.cu.o: $(SRC)/%.cu $(HDRS)
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> $(OBJ)/$*.ptxas.log || (cat $(OBJ)/$*.ptxas.log; exit 1)

$(OBJ)/%.o: $(SRC)/%.cpp $(HDRS)
	@mkdir -p $(OBJ)
	$(HOSTCXX) $(CXXFLAGS) -c $< -o $@

$(LIB): $(CU_OBJS) $(CPP_OBJS)
	@mkdir -p gffx_b200/lib
	$(NVCC) $(ARCH) -shared -ccbin $(HOSTCXX) -o $@ $^ -lpthread

$(BIN): $(SRC)/cli_main.cpp $(LIB)
	@mkdir -p gffx_b200/bin
	$(HOSTCXX) $(CXXFLAGS) -o $@ $(SRC)/cli_main.cpp -Lgffx_b200/lib -lgffx_b200 -Wl,-rpath,'$$ORIGIN/../lib'

# the workload text generator alone (bench.py --impl reference must not map the engine)
$(SYNTH_LIB): $(OBJ)/synth.o $(OBJ)/synth_capi.o
	@mkdir -p gffx_b200/lib
	$(HOSTCXX) -shared -o $@ $^ -Wl,--no-undefined -lm

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf build gffx_b200/lib gffx_b200/bin
	$(MAKE) -C oracle clean
.PHONY: all oracle clean
