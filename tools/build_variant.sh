#!/bin/bash
# A/B builds of libssb200.so with extra -D flags:  tools/build_variant.sh NAME "-DSSB_GRAM_MINB=6 ..."
# -> build/variants/NAME/libssb200.so (select with SSB200_LIB=...); contract_tc.o is shared with the product build.
set -e
cd "$(dirname "$0")/../sparse-solvers_b200/csrc"
out=../../build/variants/$1
mkdir -p "$out"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $2 -c capi.cu -o "$out/capi.o" 2>/dev/null
nvcc -shared -o "$out/libssb200.so" "$out/capi.o" contract_tc.o noalloc.o
rm -f "$out/capi.o"
echo "$out/libssb200.so"
