#!/bin/bash
# Tuning build: recompile ONE source with extra defines and link it with the objects of the regular build.
# usage: bash profiles/build_variant.sh <name> <file.cu> "<-Dflags>"  ->  litap_b200/variants/liblitap_<name>.so
set -e
cd "$(dirname "$0")/../litap_b200/csrc"
name=$1; src=$2; flags=$3
mkdir -p ../variants build_var
obj=build_var/${name}_$(basename $src .cu).o
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false \
  -Xcompiler -fPIC,-O2,-ffp-contract=off -ccbin /usr/bin/g++ -DLITAP_TILE=32 $flags -c $src -o $obj
objs=$(ls build/*.o | grep -v "build/$(basename $src .cu).o")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/liblitap_$name.so $objs $obj -lcudart -ldl
echo built ../variants/liblitap_$name.so
