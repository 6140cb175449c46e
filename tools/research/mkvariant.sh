#!/bin/bash
# Developer aid (not product): build a variant of kernels.cu into asscpp_b200/lib/variants/lib_<name>.so so that
# several variants can be compared in ONE gpurun call (tools/research/variants.sh; ASSCPP_LIB selects the library).
# usage: tools/research/mkvariant.sh <name> <kernels.cu variant> [nvcc flags...]   (run `make` in csrc first)
set -e
name=$1; src=$(readlink -f "$2"); shift; shift
cd "$(dirname "$0")/../../asscpp_b200/csrc"
mkdir -p ../lib/variants build
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC,-ffp-contract=off \
  -I. -I../../include "$@" -c "$src" -o build/kernels_$name.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../lib/variants/lib_$name.so build/kernels_$name.o \
  build/regions.o build/mesh.o build/abi.o build/multi.o build/preprocess.o build/synth.o build/hostloop.o -lpthread -ldl
echo built asscpp_b200/lib/variants/lib_$name.so
