#!/bin/bash
# A/B of library variants (build/variants/lib<name>.so; "default" = the in-tree library): C3 closest-hit kernel and the
# C4 diamond path tracer (64 iterations in one wavefront), each repeated so run-to-run noise is visible.
for rep in 1 2; do
for v in "$@"; do
  if [ "$v" = default ]; then unset IGB200_LIB; else export IGB200_LIB=$PWD/build/variants/lib$v.so; fi
  python bench.py --steps 30 --pt-iterations 0 --c5-spp 0 --office-spp 0 --cpu-rays 100000 --variant-steps 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', 'trace', round(d['value'],1), 'Mrays/s  kernel_ms', round(d['roofline']['kernel_ms'],3))"
  echo -n "$v diamond "; FIXTURE=diamond BATCH=32 python scripts/gpu_render_perf.py 1920 4 2 | sed 's/.*: //'
done
done
