#!/bin/bash
# ncu launch list (durations) of one bench step's repo kernels -> gpurun_out/<tag>_launches.csv ; usage: tools/gpu_launches.sh tag [extra bench args]
tag=$1; shift
BARGS="--steps 2 --warmup 3 --e2e-steps 0 --no-cpu-baseline --no-gpu-eager-baseline --no-from-images --no-coarse-transformer $@"
KRE='regex:blocksum_scan|cand_best|cand_select|cast_bf16|coarse_prep|coarse_stats|coarse_resolve|emit_kernel|fine_gather|fine_match|finish_lse|fold_kernel|pack_matches|row_blocksum|tf_layer|tf_pack|to_nhwc|window_merge|ctx_gemm|pose_'
timeout 300 python bench.py $BARGS > gpurun_out/plain_$tag.log 2>&1 || { echo "plain failed"; tail gpurun_out/plain_$tag.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" -c 600 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py $BARGS > gpurun_out/ncu_$tag.log 2>&1
echo "ncu exit $?"
python tools/launch_summary.py gpurun_out/${tag}_launches.csv
