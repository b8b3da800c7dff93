#!/bin/bash
# Round-2 closing session (1 GPU): full GPU suite, smoke, default bench + reference arm, launch list.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest exit $?"; tail -n 3 gpurun_out/pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
timeout 600 python bench.py > gpurun_out/r02_bench_d.json 2> gpurun_out/bench_d.err; echo "bench exit $?"; head -c 400 gpurun_out/r02_bench_d.json; echo
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_ref_d.json 2>/dev/null; echo "ref exit $?"
bash tools/gpu_launches.sh r02e > gpurun_out/launches_e.txt 2>&1; tail -n 28 gpurun_out/launches_e.txt
