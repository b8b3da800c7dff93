#!/bin/bash
# A/B build: tools/ab_build.sh <tag> <source-without-.cu> "<extra nvcc flags>"  ->  featurematching_b200/ab/libfm_<tag>.so
# (all other objects come from featurematching_b200/build/, so run the normal build first); select with FM_B200_LIB=...
set -e
cd "$(dirname "$0")/../featurematching_b200"
mkdir -p ab
tag=$1; src=$2; extra=$3
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden $extra -c csrc/$src.cu -o ab/${src}_$tag.o
objs=$(ls build/*.o | grep -v "/$src.o")
nvcc -shared -o ab/libfm_$tag.so $objs ab/${src}_$tag.o -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -cudart shared
echo "built ab/libfm_$tag.so"
