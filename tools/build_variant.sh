#!/bin/bash
# tools/build_variant.sh <name> <nvcc -D flags...>: experiment build of libgm_cuda with different pr_sweep constants
# -> green-marl_b200/lib/libgm_cuda_<name>.so (git-ignored); load it with GMC_LIB=<path> in the tools/ drivers.
set -e
cd "$(dirname "$0")/../green-marl_b200"
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -ccbin g++ -Xcompiler -fPIC,-fopenmp -I../include -I../include/gm -Icsrc \
  --expt-relaxed-constexpr "$@" -c csrc/gmc_pagerank.cu -o build/csrc/gmc_pagerank_$name.o
objs=$(ls build/csrc/*.o build/host/*.o | grep -v "gmc_pagerank")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -ccbin g++ -Xcompiler -fopenmp -o lib/libgm_cuda_$name.so build/csrc/gmc_pagerank_$name.o $objs -lcudart_static -lgomp -lpthread -ldl -lrt
echo built lib/libgm_cuda_$name.so
