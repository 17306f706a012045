# Build of the B200-native ORFanage hot path.
#   make host     host front-end objects (C++17, no CUDA)
#   make cuda     orfanage_b200/liborfx.so (sm_100a) + orfanage_b200/bin/orfanage
#   make oracle   test infrastructure: oracle/liborfx_oracle.so, oracle/orfanage_oracle_cli, oracle/_ref/*
#   make tools    tools/orfx_synth
CXX := g++
NVCC ?= nvcc
CXXFLAGS := -O2 -g -std=c++17 -fPIC -Wall -Wextra -Wno-unused-parameter -Iinclude
NVFLAGS := -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Iinclude \
           -Xptxas -v --expt-relaxed-constexpr
HOST_SRC := $(wildcard orfanage_b200/csrc/host/*.cpp)
HOST_OBJ := $(patsubst orfanage_b200/csrc/host/%.cpp,build/host/%.o,$(filter-out %/main.cpp %/orfcompare_main.cpp,$(HOST_SRC)))
CU_SRC := $(wildcard orfanage_b200/csrc/*.cu)
CU_HDR := $(wildcard orfanage_b200/csrc/*.cuh) include/orfx.h include/orfx_codon_table.h

all: tools host oracle cuda

host: $(HOST_OBJ)
build/host/%.o: orfanage_b200/csrc/host/%.cpp orfanage_b200/csrc/host/orfx_host.h include/orfx.h include/orfx_pack.h
	@mkdir -p build/host
	$(CXX) $(CXXFLAGS) -c $< -o $@

cuda: orfanage_b200/liborfx.so orfanage_b200/bin/orfanage orfanage_b200/bin/orfcompare
host: orfanage_b200/liborfx_host.so
orfanage_b200/liborfx_host.so: $(HOST_OBJ)
	$(CXX) -shared -o $@ $(HOST_OBJ) -lpthread
orfanage_b200/liborfx.so: $(CU_SRC) $(CU_HDR)
	$(NVCC) $(NVFLAGS) -shared $(CU_SRC) -o $@ -ldl 2> build/ptxas.log || (cat build/ptxas.log; false)
	@grep -E "registers|spill|error" build/ptxas.log | sort | uniq -c | sort -rn | head -40 || true
orfanage_b200/bin/orfanage: $(HOST_OBJ) build/host/main.o orfanage_b200/liborfx.so
	@mkdir -p orfanage_b200/bin
	$(CXX) -o $@ $(HOST_OBJ) build/host/main.o -Lorfanage_b200 -lorfx -Wl,-rpath,'$$ORIGIN/..' -lpthread
orfanage_b200/bin/orfcompare: $(HOST_OBJ) build/host/orfcompare_main.o orfanage_b200/liborfx.so
	@mkdir -p orfanage_b200/bin
	$(CXX) -o $@ $(HOST_OBJ) build/host/orfcompare_main.o -Lorfanage_b200 -lorfx -Wl,-rpath,'$$ORIGIN/..' -lpthread
build/host/main.o: orfanage_b200/csrc/host/main.cpp orfanage_b200/csrc/host/orfx_host.h include/orfx.h
	@mkdir -p build/host
	$(CXX) $(CXXFLAGS) -c $< -o $@

oracle: oracle/liborfx_oracle.so oracle/orfanage_oracle_cli
	$(MAKE) -C oracle ref
oracle/liborfx_oracle.so: oracle/orfx_oracle.cpp include/orfx.h include/orfx_codon_table.h
	$(CXX) $(CXXFLAGS) -fopenmp -shared $< -o $@
oracle/orfanage_oracle_cli: oracle/oracle_cli.cpp oracle/orfx_oracle.cpp $(HOST_OBJ)
	$(CXX) $(CXXFLAGS) -fopenmp -o $@ oracle/oracle_cli.cpp oracle/orfx_oracle.cpp $(HOST_OBJ) -lpthread

tools: tools/orfx_synth
tools/orfx_synth: tools/orfx_synth.cpp
	$(CXX) -O2 -std=c++17 -o $@ $<

clean:
	rm -rf build orfanage_b200/liborfx.so orfanage_b200/liborfx_host.so orfanage_b200/bin oracle/liborfx_oracle.so oracle/orfanage_oracle_cli tools/orfx_synth
.PHONY: all host cuda oracle tools clean
