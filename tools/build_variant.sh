#!/bin/bash
# tools/build_variant.sh NAME "-DFLAG ..." : builds coherent-rtlsdr_b200/lib/libcoherent_cuda_NAME.so with extra
# nvcc flags (kernel A/B experiments; select at run time with COHCUDA_LIB=<path>).  Not part of the product build.
set -e
cd "$(dirname "$0")/.."
NAME=$1; FLAGS=$2
CSRC=coherent-rtlsdr_b200/csrc
OUT=build/variant_$NAME
mkdir -p $OUT coherent-rtlsdr_b200/lib
for f in api fft_kernels fft_big apply_kernel cov_kernel frac_kernel prims lag_static; do
  nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Xcompiler -fPIC -Iinclude -I$CSRC --fmad=true $FLAGS -c $CSRC/$f.cu -o $OUT/$f.o &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o coherent-rtlsdr_b200/lib/libcoherent_cuda_$NAME.so $OUT/*.o -cudart static
echo built coherent-rtlsdr_b200/lib/libcoherent_cuda_$NAME.so
