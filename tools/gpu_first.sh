#!/bin/bash
# First GPU session: run the parity groups in separate processes (a trapped kernel poisons its
# CUDA context), each under its own timeout; logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.log 2>&1
run() { name=$1; shift; timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "$1" -p no:cacheprovider > gpurun_out/pytest_$name.log 2>&1; echo "$name exit $?"; tail -n 25 gpurun_out/pytest_$name.log; }
run fine "fine"
run simt "coarse and simt"
run tc "coarse and tcgen05"
run rest "not fine and not simt and not tcgen05"
