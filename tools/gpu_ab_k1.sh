#!/bin/bash
# A/B of K1 builds: parity tests of the coarse stage + the bench's K1 / coarse-stage time.  usage: tools/gpu_ab_k1.sh tag...
for tag in "$@"; do
    if [ "$tag" = "cur" ]; then unset FM_B200_LIB; else export FM_B200_LIB=$PWD/featurematching_b200/ab/libfm_$tag.so; fi
    echo "== $tag"
    timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "coarse" -p no:cacheprovider -x 2>&1 | tail -n 2
    timeout 300 python bench.py --steps 30 --warmup 5 --e2e-steps 0 --no-cpu-baseline --no-gpu-eager-baseline --no-from-images --no-coarse-transformer --without-fine-transformer 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1])
k=d['kernels']; print('K1 ms %.4f frac_burst %.3f | coarse_stage ms %.4f frac_burst %.3f | step %.3f' % (k['coarse_stats_tc']['ms'], k['coarse_stats_tc']['frac'], k['coarse_stage']['ms'], k['coarse_stage']['frac'], d['ms_per_step']))"
done
