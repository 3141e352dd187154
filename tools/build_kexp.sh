#!/bin/bash
# builds tools/kexp (experiment binary): nvcc sm_100a, the library's CUDA sources + the host generator object
set -e
cd "$(dirname "$0")/.."
gcc -O2 -fopenmp -ffp-contract=off -c -o /tmp/b200amg.o spmv_research_b200/csrc/amg.c
C=spmv_research_b200/csrc
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fopenmp -o tools/kexp tools/kexp.cu $C/capi.cu $C/csr_kernel.cu $C/sell_kernel.cu $C/convert.cu $C/partition.cu $C/pcg.cu $C/panels.cu /tmp/b200amg.o -lm
