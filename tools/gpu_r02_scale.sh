#!/bin/bash
# Round-2 multi-GPU lines: usage tools/gpu_r02_scale.sh N "cfg1 cfg2 ..."   (N ranks of one box, torchrun, NCCL)
N=$1; CFGS=${2:-"640x512 1280x1024 seq10k"}
mkdir -p gpurun_out
for cfg in $CFGS; do
    extra="--steps 20 --warmup 3 --no-cpu-baseline --no-gpu-eager-baseline --no-from-images"
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
        bench.py --gpus $N --config $cfg $extra > gpurun_out/r02_bench_${cfg}_n$N.json 2> gpurun_out/bench_${cfg}_n$N.err
    echo "bench $cfg N=$N exit $?"
    head -c 300 gpurun_out/r02_bench_${cfg}_n$N.json; echo; grep -v "^W\|^$\|\*\*\*" gpurun_out/bench_${cfg}_n$N.err | tail -n 3
done
