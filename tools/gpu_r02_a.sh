#!/bin/bash
# Round-2 session A: full GPU test suite, default bench line + reference arm, ncu launch list of one step,
# ncu --set full of the shipped fine kernels, compute-sanitizer over the small-shape tests.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.log 2>&1
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -n 8 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/r02_bench_b.json 2> gpurun_out/bench_b.err; echo "bench exit $?"
tail -n 5 gpurun_out/bench_b.err
head -c 600 gpurun_out/r02_bench_b.json; echo
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_ref_b.json 2> gpurun_out/bench_ref_b.err; echo "ref exit $?"
BARGS="--steps 2 --warmup 3 --e2e-steps 0 --no-cpu-baseline --no-gpu-eager-baseline --no-from-images --no-coarse-transformer"
timeout 300 python bench.py $BARGS > gpurun_out/plain.log 2>&1; echo "plain exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/r02_launches_b.csv python bench.py $BARGS > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
for spec in "window_merge_kernel:k3" "fine_match_kernel:k4" "tf_layer_kernel:tf" "coarse_stats_tc_kernel:k1"; do
    k=${spec%%:*}; tag=${spec##*:}
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$k" --launch-skip 3 -c 1 -f -o gpurun_out/r02_$tag \
        python bench.py $BARGS > gpurun_out/ncu_$tag.log 2>&1; echo "ncu full $tag exit $?"
    ncu -i gpurun_out/r02_$tag.ncu-rep --page raw --csv > gpurun_out/r02_${tag}_raw.csv 2>/dev/null
done
# sanitizer: the small-shape parity tests (one process per tool)
SEL="golden or small_windows or empty or special or ties or zero_count or noise_free or without_merge"
for tool in memcheck racecheck synccheck; do
    timeout 900 compute-sanitizer --tool $tool --print-limit 20 python -m pytest tests/test_gpu_parity.py tests/test_fine_transformer.py tests/test_pose.py tests/test_coarse_transformer.py \
        -q -m gpu -k "$SEL" -p no:cacheprovider -x > gpurun_out/r02_sanitizer_$tool.log 2>&1; echo "$tool exit $?"
    grep -E "ERROR SUMMARY|passed|failed|RACECHECK SUMMARY" gpurun_out/r02_sanitizer_$tool.log | tail -n 4
done
