#!/bin/bash
# Round-2 session B (1 GPU): other BASELINE configs at N = 1, launch list of one step (repo kernels only), K1 full capture.
mkdir -p gpurun_out
BARGS="--steps 2 --warmup 3 --e2e-steps 0 --no-cpu-baseline --no-gpu-eager-baseline --no-from-images --no-coarse-transformer"
KRE='regex:blocksum_scan|cand_best|cand_select|cand_redo|cast_bf16|coarse_prep|coarse_stats|emit_kernel|fine_gather|fine_match|finish_lse|fold_kernel|pack_matches|row_blocksum|tf_layer|tf_pack|to_nhwc|window_merge|ctx_gemm|tie_|redo'
timeout 300 python bench.py $BARGS > gpurun_out/plain.log 2>&1; echo "plain exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" -c 400 --csv \
    --log-file gpurun_out/r02_launches_b.csv python bench.py $BARGS > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches exit $?"
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k "$KRE" -c 400 --csv \
    --log-file gpurun_out/r02_launches_b_dram.csv python bench.py $BARGS > gpurun_out/ncu_launches2.log 2>&1; echo "ncu dram exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:coarse_stats_tc_kernel --launch-skip 4 -c 1 -f -o gpurun_out/r02_k1 \
    python bench.py $BARGS > gpurun_out/ncu_k1.log 2>&1; echo "ncu full k1 exit $?"
ncu -i gpurun_out/r02_k1.ncu-rep --page raw --csv > gpurun_out/r02_k1_raw.csv 2>/dev/null
for cfg in 1280x1024 seq320 seq10k; do
    timeout 900 python bench.py --config $cfg --steps 20 --warmup 3 > gpurun_out/r02_bench_${cfg}_n1.json 2> gpurun_out/bench_${cfg}.err; echo "bench $cfg exit $?"
    head -c 400 gpurun_out/r02_bench_${cfg}_n1.json; echo; tail -n 3 gpurun_out/bench_${cfg}.err
done
