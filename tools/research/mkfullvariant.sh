#!/bin/bash
# Developer aid (not product): build the WHOLE library with extra -D flags into asscpp_b200/lib/variants/lib_<name>.so
# (for constants shared by host preprocessing and kernels).  usage: tools/research/mkfullvariant.sh <name> -DX=Y ...
set -e
name=$1; shift
cd "$(dirname "$0")/../../asscpp_b200/csrc"
B=build_$name
mkdir -p ../lib/variants $B
for f in kernels regions mesh abi; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -Xcompiler -fPIC,-ffp-contract=off "$@" -c $f.cu -o $B/$f.o &
done
for f in multi preprocess synth hostloop; do
  g++ -O3 -std=c++17 -fPIC -ffp-contract=off -fno-fast-math -pthread -I/usr/local/cuda/include "$@" -c $f.cpp -o $B/$f.o &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../lib/variants/lib_$name.so $B/kernels.o $B/regions.o $B/mesh.o $B/abi.o $B/multi.o $B/preprocess.o $B/synth.o $B/hostloop.o -lpthread -ldl
rm -rf $B
echo built asscpp_b200/lib/variants/lib_$name.so
