#!/bin/bash
# A/B of fine-transformer builds: correctness (vs the oracle) + time per variant.  usage: tools/gpu_ab_tf.sh tag1 tag2 ...
for tag in "$@"; do
    if [ "$tag" = "cur" ]; then unset FM_B200_LIB; else export FM_B200_LIB=$PWD/featurematching_b200/ab/libfm_$tag.so; fi
    echo "== $tag"
    python tools/dbg_tf.py self,cross 301 2>&1 | tail -n 1
    python tools/bench_tf.py 124877 self 2>&1 | tail -n 1
    python tools/bench_tf.py 124877 cross 2>&1 | tail -n 1
done
