#!/bin/bash
# debug build of one kernel file with extra defines: FM_SRC=coarse_transformer FM_EXTRA="-DCT_EPI_NOP" tools/dbg_build.sh
# (default FM_SRC=fine_transformer; FM_EXTRA=-DFM_DEBUG_WAIT prints barrier timeouts)
set -e
cd "$(dirname "$0")/../featurematching_b200/csrc"
FM_SRC=${FM_SRC:-fine_transformer}
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden $FM_EXTRA -c $FM_SRC.cu -o ../build/$FM_SRC.o
cd .. && nvcc -shared -o libfm_b200.so build/*.o -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -cudart shared
