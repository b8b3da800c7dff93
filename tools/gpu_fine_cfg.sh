#!/bin/bash
# fine_match ring-shape sweep (FM_FINE_CFG = <warps><stages>)
mkdir -p gpurun_out
for cfg in 42 81 33 24 22; do
  FM_FINE_CFG=$cfg python bench.py --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 0 > gpurun_out/bench_cfg$cfg.json 2> gpurun_out/bench_cfg$cfg.err
  python -c "
import json; d=json.load(open('gpurun_out/bench_cfg$cfg.json')); print('cfg $cfg', round(d['ms_per_step'],3), {k:(round(v['ms'],3), round(v['frac'],3)) for k,v in d['kernels'].items()})"
done
