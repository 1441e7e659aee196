#!/bin/bash
# usage: tools/build_variant.sh <name> <source.cu> [-DFLAG ...]   -> tools/_libmde_<name>.so
# One source recompiled with extra macros, linked against the objects of the regular build (A/B timing of kernel variants in ONE gpurun call).
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; shift 2
P=monodepth-estimation_b200
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -c $P/csrc/$src -o /tmp/_variant_$name.o
base=${src%.cu}; base=${base#_}; base=${base%%_v_*}   # _conv_halo_v_desc.cu replaces conv_halo.o
objs=$(ls $P/build/*.o | grep -v "/${base}.o")
nvcc -shared -o tools/_libmde_$name.so $objs /tmp/_variant_$name.o -gencode arch=compute_100a,code=sm_100a
echo tools/_libmde_$name.so
