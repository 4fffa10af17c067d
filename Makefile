# Top-level build: the product library (CUDA, sm_100a only), the peak probe, the C harness and the checker.
#   make            -> libdgemm_b200.so + peak_probe + dgemm_b200 harness + oracle
#   make lib        -> optimizing-dgemm-on-intel-cpus-with-avx512f_b200/libdgemm_b200.so
# Mirrors the reference's Makefile:1-15 (one `build` target producing dgemm_x86) with an nvcc target added.
PKG      := optimizing-dgemm-on-intel-cpus-with-avx512f_b200
CSRC     := $(PKG)/csrc
NVCC     ?= nvcc
# the image exports CC/CXX=/opt/gcc/bin/..., wrappers without libgomp: pin the system compilers
HOSTCC   := /usr/bin/gcc
HOSTCXX  := /usr/bin/g++
ARCH     := -gencode arch=compute_100a,code=sm_100a
NVFLAGS  := $(ARCH) -O3 -lineinfo -std=c++17 -ccbin $(HOSTCXX) -Xcompiler -fPIC -Xcompiler -Wall
LIB      := $(PKG)/libdgemm_b200.so
LIB_SRCS := $(CSRC)/dgemm_kernels.cu $(CSRC)/dgemm_rank.cu $(CSRC)/dgemm_rankplan.cpp $(CSRC)/dgemm_kernels_small.cu $(CSRC)/dgemm_kernels_small2.cu $(CSRC)/dgemm_kernels_mid.cu $(CSRC)/dgemm_kernels_tiny.cu $(CSRC)/dgemm_kernels_cluster.cu $(CSRC)/dgemm_ladder.cu $(CSRC)/dgemm_api.cu $(CSRC)/dgemm_multi.cu \
            $(CSRC)/dgemm_hostplan.cpp $(CSRC)/dgemm_staging.cu $(CSRC)/dgemm_static_scratch.cu
LIB_HDRS := $(CSRC)/dgemm_device.cuh $(CSRC)/dgemm_launch.h $(CSRC)/dgemm_params.h $(CSRC)/dgemm_tma_kernel.cuh $(CSRC)/dgemm_small_tile.cuh $(CSRC)/dgemm_host.h include/dgemm_b200.h

all: lib probe harness oracle
build: all

lib: $(LIB)
$(LIB): $(LIB_SRCS) $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) -shared -o $@ $(LIB_SRCS) -Xlinker --no-undefined -ldl -lpthread

# experimental build: 16 consumer warps (4 per SM sub-partition), 32x32 warp tiles; load with DGEMM_B200_LIB=...
lib16: $(PKG)/libdgemm_b200_w16.so
$(PKG)/libdgemm_b200_w16.so: $(LIB_SRCS) $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) -DDG_WARPS_M=4 -DDG_WARPS_N=4 -shared -o $@ $(LIB_SRCS) -Xlinker --no-undefined -ldl -lpthread

# measurement build: the producer hands every stage over WITHOUT loading it (results are garbage, timing is the consumers'
# loop alone); load with DGEMM_B200_LIB=.../libdgemm_b200_dbg_notma.so (tools/history/r02_call19.sh)
libdbg: $(PKG)/libdgemm_b200_dbg_notma.so
$(PKG)/libdgemm_b200_dbg_notma.so: $(LIB_SRCS) $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) -DDG_DEBUG_NO_TMA -shared -o $@ $(LIB_SRCS) -Xlinker --no-undefined -ldl -lpthread

# measurement builds: timeline stamps in SM cycles (clock64) instead of nanoseconds, with and without TMA; tools/timeline.py
# then prints kilo-cycles where it says us (per-CTA differences only: the counters of different SMs are not aligned)
libclk: $(PKG)/libdgemm_b200_clk.so $(PKG)/libdgemm_b200_clk_notma.so
$(PKG)/libdgemm_b200_clk.so: $(LIB_SRCS) $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) -DDG_STAMP_CLOCK64 -shared -o $@ $(LIB_SRCS) -Xlinker --no-undefined -ldl -lpthread
$(PKG)/libdgemm_b200_clk_notma.so: $(LIB_SRCS) $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) -DDG_STAMP_CLOCK64 -DDG_DEBUG_NO_TMA -shared -o $@ $(LIB_SRCS) -Xlinker --no-undefined -ldl -lpthread

# measurement builds: the two consumer warps of a sub-partition skewed by n/64 of a stage (DG_SKEW_64THS)
SKEWS ?= 2 4 8 12 16
libskew: $(foreach n,$(SKEWS),$(PKG)/libdgemm_b200_skew$(n).so)
$(PKG)/libdgemm_b200_skew%.so: $(LIB_SRCS) $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) -DDG_SKEW_64THS=$* -shared -o $@ $(LIB_SRCS) -Xlinker --no-undefined -ldl -lpthread

# measurement builds: unroll factor of the consumers' k-tile loop (DG_KT_UNROLL; ptxas unrolls by 2 on its own)
libunroll: $(PKG)/libdgemm_b200_unroll2.so $(PKG)/libdgemm_b200_unroll4.so
$(PKG)/libdgemm_b200_unroll%.so: $(LIB_SRCS) $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) -DDG_KT_UNROLL=$* -shared -o $@ $(LIB_SRCS) -Xlinker --no-undefined -ldl -lpthread

# measurement builds: two k-tiles per ring slot in EVERY tile configuration (DG_KT_PER_STAGE=2, half as many slots)
libkps: $(PKG)/libdgemm_b200_kps2.so
$(PKG)/libdgemm_b200_kps2.so: $(LIB_SRCS) $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) -DDG_KT_PER_STAGE=2 -shared -o $@ $(LIB_SRCS) -Xlinker --no-undefined -ldl -lpthread

# measurement builds of the 64 x 32 tile alone (every other object is shared): libdgemm_b200_tiny_<name>.so
#   kps2: two k-tiles per ring slot; nodual: one accumulator set; latedesc: descriptor prefetch after the barrier init
OBJDIR   := build/obj
LIB_OBJS := $(patsubst $(CSRC)/%,$(OBJDIR)/%.o,$(filter-out $(CSRC)/dgemm_kernels_tiny.cu,$(LIB_SRCS)))
$(OBJDIR)/%.o: $(CSRC)/% $(LIB_HDRS)
	mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -c -o $@ $<
TINY_FLAGS_kps2     := -DDG_TINY_KPS=2
TINY_FLAGS_nodual   := -DDG_TINY_DUAL_ACC=0
TINY_FLAGS_latedesc := -DDG_TINY_EARLY_DESC=0
TINY_FLAGS_plain    := -DDG_TINY_DUAL_ACC=0 -DDG_TINY_EARLY_DESC=0
TINY_FLAGS_w16      := -DDG_TINY_WARPS16=1
TINY_FLAGS_w4       := -DDG_TINY_WARPS4=1
TINY_FLAGS_clk      := -DDG_STAMP_CLOCK64
TINY_FLAGS_skew8    := -DDG_STAMP_CLOCK64 -DDG_SKEW_64THS=8
TINY_FLAGS_skew16   := -DDG_STAMP_CLOCK64 -DDG_SKEW_64THS=16
TINY_FLAGS_skew32   := -DDG_STAMP_CLOCK64 -DDG_SKEW_64THS=32
TINY_FLAGS_regs96   := -DDG_STAMP_CLOCK64 -DDG_MIN_CTAS=2 -DDG_PRODUCER_REGS=40 -DDG_CONSUMER_REGS=96
TINY_FLAGS_st8      := -DDG_STAMP_CLOCK64 -DDG_TINY_STAGES=8
TINY_FLAGS_notma    := -DDG_DEBUG_NO_TMA
TINY_FLAGS_w4notma  := -DDG_TINY_WARPS4=1 -DDG_DEBUG_NO_TMA
libtiny: $(foreach n,kps2 nodual latedesc plain w16 w4,$(PKG)/libdgemm_b200_tiny_$(n).so)
$(PKG)/libdgemm_b200_tiny_%.so: $(LIB_OBJS) $(CSRC)/dgemm_kernels_tiny.cu $(LIB_HDRS)
	$(NVCC) $(NVFLAGS) $(TINY_FLAGS_$*) -c -o $(OBJDIR)/tiny_$*.o $(CSRC)/dgemm_kernels_tiny.cu
	$(NVCC) $(ARCH) -shared -o $@ $(LIB_OBJS) $(OBJDIR)/tiny_$*.o -Xlinker --no-undefined -ldl -lpthread

probe: $(PKG)/bin/peak_probe
$(PKG)/bin/peak_probe: $(CSRC)/peak_probe.cu
	mkdir -p $(PKG)/bin
	$(NVCC) $(ARCH) -O3 -lineinfo -ccbin $(HOSTCXX) -o $@ $<

# measurement aid for the pageable-memory staging (tools/stage_probe.cu)
stage_probe: $(PKG)/bin/stage_probe
$(PKG)/bin/stage_probe: tools/stage_probe.cu
	mkdir -p $(PKG)/bin
	$(NVCC) $(ARCH) -O3 -std=c++17 -ccbin $(HOSTCXX) -o $@ $< -lpthread

harness: $(PKG)/bin/dgemm_b200
$(PKG)/bin/dgemm_b200: $(PKG)/harness/test.c $(PKG)/harness/utils.c $(PKG)/harness/utils.h include/kernel20.h include/dgemm_b200.h $(LIB)
	mkdir -p $(PKG)/bin
	$(HOSTCC) -O2 -Wall -Iinclude -I$(PKG)/harness -o $@ $(PKG)/harness/test.c $(PKG)/harness/utils.c \
	    -L$(PKG) -ldgemm_b200 -Wl,-rpath,'$$ORIGIN/..' -lm

oracle:
	$(MAKE) -C oracle all

sass: $(LIB)
	cuobjdump -sass $(LIB) | grep -E "DMMA|UTMALDG|LDGSTS|SYNCS" | sed 's/\/\*[0-9a-f]*\*\///g' | awk '{print $$1}' | sort | uniq -c

clean:
	rm -f $(LIB) $(PKG)/bin/peak_probe $(PKG)/bin/dgemm_b200
	$(MAKE) -C oracle clean
.PHONY: all build lib lib16 libdbg libkps libtiny libclk libskew libunroll probe stage_probe harness oracle sass clean
