#!/bin/bash
# Round-2 quick loop on one GPU: trace parity tests, C3 kernel bench for each library variant given, optional full ncu
# capture of the closest-hit kernel with the per-line / per-phase summary.   usage: scripts/r02_quick.sh <tag> [profile] [variants...]
tag=$1; prof=$2; shift 2
mkdir -p gpurun_out
python -m pytest tests/test_trace_gpu.py -x -q -m gpu > gpurun_out/${tag}_pytest_trace.log 2>&1; echo "pytest rc $?" >> gpurun_out/${tag}_pytest_trace.log
tail -3 gpurun_out/${tag}_pytest_trace.log
bash scripts/ab_variants.sh default "$@" 2>&1 | tee gpurun_out/${tag}_ab.txt
if [ "$prof" = profile ]; then
  python scripts/ncu_trace_target.py 2 > /dev/null || exit 1
  ncu --set full --clock-control none --import-source on -k regex:trace_kernel -o gpurun_out/${tag}_trace_c3 -f \
      python scripts/ncu_trace_target.py 2 > gpurun_out/${tag}_ncu_trace.log 2>&1
  python scripts/ncu_summary.py gpurun_out/${tag}_trace_c3.ncu-rep gpurun_out/${tag}_trace_c3_full_summary.csv
fi
