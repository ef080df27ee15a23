#!/bin/bash
# scripts/build_variant.sh <name> <source.cu> <extra nvcc flags...>: the library with ONE source recompiled
# under extra flags -> scripts/variants/lib_<name>.so (load it with B200_LIB=...; kernel experiments only)
set -e
cd "$(dirname "$0")/.."
name=$1; src=$2; shift 2
mkdir -p scripts/variants
obj=scripts/variants/${name}_$(basename ${src%.cu}).o
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -DB200_ARCH=100 \
  --expt-relaxed-constexpr "$@" -c wenet_celoss_b200/csrc/$src -o $obj
others=$(ls wenet_celoss_b200/_obj/*.o | grep -v "/$(basename ${src%.cu}).o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o scripts/variants/lib_${name}.so $obj $others -lcudart
echo scripts/variants/lib_${name}.so
