# Builds libcoherent_cuda.so (sm_100a only) in-tree and the oracle checker.
NVCC    ?= nvcc
PKG     := coherent-rtlsdr_b200
CSRC    := $(PKG)/csrc
LIB     := $(PKG)/lib/libcoherent_cuda.so
ARCH    := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -Iinclude -I$(CSRC) --fmad=true
OBJS    := $(CSRC)/api.o $(CSRC)/fft_kernels.o $(CSRC)/fft_big.o $(CSRC)/apply_kernel.o $(CSRC)/cov_kernel.o $(CSRC)/frac_kernel.o $(CSRC)/prims.o $(CSRC)/lag_static.o

all: $(LIB) oracle

$(CSRC)/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh $(CSRC)/*.inc) include/coherent_cuda.h
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(LIB): $(OBJS)
	mkdir -p $(PKG)/lib
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -cudart static

oracle:
	$(MAKE) -s -C oracle all

clean:
	rm -f $(OBJS) $(LIB)
.PHONY: all oracle clean
