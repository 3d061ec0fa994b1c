#!/bin/bash
# Builds an experimental copy of the library: tools/_bin/libepik_b200_<name>.so with extra nvcc flags
# (selected at run time with EPIK_B200_LIB=...). Usage: tools/build_variant.sh <name> [nvcc flags...]
set -e
NAME=$1; shift
cd "$(dirname "$0")/.."
mkdir -p tools/_bin
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared "$@" \
  -o tools/_bin/libepik_b200_$NAME.so epik_b200/csrc/db.cu epik_b200/csrc/placer.cu epik_b200/csrc/microbench.cu epik_b200/csrc/pool.cu
echo built tools/_bin/libepik_b200_$NAME.so
