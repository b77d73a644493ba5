# Build of the B200 log-scan engine.  `make` builds everything that travels to the GPU box:
#   operator_b200/libpodmortem_scan.so   product: compiler + runtime + sm_100a kernels (C ABI)
#   tests/support/libpm_testsupport.so   CPU unit-test harness for the pattern compiler
#   oracle/_build/liboracle_pcre2.so     test oracle / CPU baseline (PCRE2 via dlopen)
#   operator_b200/libpm_synth.so         synthetic corpus generator, host and device (bench + tests input)
NVCC ?= /usr/local/cuda/bin/nvcc
CXX ?= g++
ARCH := -gencode arch=compute_100a,code=sm_100a
CXXFLAGS := -O2 -std=c++17 -fPIC -Wall -Wno-unused-function -Iinclude -I/usr/local/cuda/include
NVFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Iinclude -Xptxas -v --fmad=false

CSRC := operator_b200/csrc
HOST_SRCS := $(CSRC)/pm_json.cc $(CSRC)/pm_regex.cc $(CSRC)/pm_automata.cc $(CSRC)/pm_literal.cc $(CSRC)/pm_library.cc $(CSRC)/pm_split.cc $(CSRC)/pm_service.cc
HOST_OBJS := $(HOST_SRCS:.cc=.o)
CU_SRCS := $(wildcard $(CSRC)/*.cu)
CU_OBJS := $(CU_SRCS:.cu=.o)
HDRS := $(wildcard $(CSRC)/*.h) $(wildcard $(CSRC)/*.cuh) include/podmortem_scan.h

all: operator_b200/libpodmortem_scan.so tests/support/libpm_testsupport.so oracle/_build/liboracle_pcre2.so operator_b200/libpm_synth.so

$(CSRC)/%.o: $(CSRC)/%.cc $(HDRS)
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(CSRC)/%.o: $(CSRC)/%.cu $(HDRS)
	$(NVCC) $(NVFLAGS) -c $< -o $@ 2> $@.ptxas.log || (cat $@.ptxas.log; false)

operator_b200/libpodmortem_scan.so: $(HOST_OBJS) $(CU_OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $^ -lcudart

tests/support/libpm_testsupport.so: tests/support/pm_testsupport.cc $(filter-out $(CSRC)/pm_service.o,$(HOST_OBJS))
	$(CXX) $(CXXFLAGS) -shared -o $@ $^

oracle/_build/liboracle_pcre2.so: oracle/scan_pcre2.c
	mkdir -p oracle/_build
	gcc -O2 -fPIC -shared -pthread -o $@ $< -ldl

operator_b200/libpm_synth.so: operator_b200/csrc/synth/pm_synth.cc operator_b200/csrc/synth/pm_synth_gpu.cu operator_b200/csrc/synth/pm_synth_core.h
	$(NVCC) $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-pthread -shared -o $@ operator_b200/csrc/synth/pm_synth.cc operator_b200/csrc/synth/pm_synth_gpu.cu -lcudart

# Measurement build with warp time stamps in the post-scan kernels (tools_warp_trace.py); not part of `all`.
trace: operator_b200/libpodmortem_scan_trace.so
$(CSRC)/pm_kernels_trace.o: $(CSRC)/pm_kernels.cu $(HDRS)
	$(NVCC) $(NVFLAGS) -DPM_TRACE -c $< -o $@ 2> $@.ptxas.log || (cat $@.ptxas.log; false)
operator_b200/libpodmortem_scan_trace.so: $(HOST_OBJS) $(filter-out $(CSRC)/pm_kernels.o $(CSRC)/pm_kernels_trace.o,$(CU_OBJS)) $(CSRC)/pm_kernels_trace.o
	$(NVCC) $(ARCH) -shared -o $@ $^ -lcudart

clean:
	rm -f $(CSRC)/*.o $(CSRC)/*.log operator_b200/*.so tests/support/*.so oracle/_build/*.so

.PHONY: all clean trace
