#!/bin/bash
# debug build of the fine transformer with barrier-timeout printf (FM_DEBUG_WAIT)
set -e
cd "$(dirname "$0")/../featurematching_b200/csrc"
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden $FM_EXTRA -c fine_transformer.cu -o ../build/fine_transformer.o
cd .. && nvcc -shared -o libfm_b200.so build/*.o -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -cudart shared
