#!/bin/bash
# Trace parity tests against each experiment build: scripts/r02_variants_parity.sh <tag> <variants...>
tag=$1; shift
for v in "$@"; do
  IGB200_LIB=$PWD/build/variants/lib$v.so python -m pytest tests/test_trace_gpu.py -x -q -m gpu > gpurun_out/${tag}_parity_$v.log 2>&1
  echo "$v parity rc $? $(tail -1 gpurun_out/${tag}_parity_$v.log)"
done
