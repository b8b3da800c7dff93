#!/bin/bash
# usage: tools/gpu_ncu_full.sh <kernel-regex> <tag>   -- one `ncu --set full` capture of one launch
# (after the plain command exited 0), report lands in gpurun_out/<tag>.ncu-rep + raw csv.
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --e2e-steps 0 --no-cpu-baseline > gpurun_out/plain_$2.log 2>&1 || { echo "plain run failed"; tail gpurun_out/plain_$2.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$1" --launch-skip 3 -c 1 -f -o gpurun_out/$2 \
    python bench.py --steps 2 --warmup 3 --e2e-steps 0 --no-cpu-baseline > gpurun_out/ncu_$2.log 2>&1
echo "ncu full $2 exit $?"
ncu -i gpurun_out/$2.ncu-rep --page raw --csv > gpurun_out/$2_raw.csv 2>/dev/null
