#!/bin/bash
# usage: tools/gpu_tests.sh "<pytest -k expression>"   (separate process, own timeout)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -k "$1" -p no:cacheprovider -x > gpurun_out/pytest_sel.log 2>&1; echo "exit $?"
tail -n 40 gpurun_out/pytest_sel.log
