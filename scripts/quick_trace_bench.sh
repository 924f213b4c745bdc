#!/bin/bash
# runs the C3 closest-hit bench (kernel only) for each library variant given; prints Mrays/s
for v in "$@"; do
  if [ "$v" = default ]; then unset IGB200_LIB; else export IGB200_LIB=$PWD/build/variants/lib$v.so; fi
  python bench.py --steps 30 --pt-iterations 0 --c5-spp 0 --office-spp 0 --variant-steps 0 --cpu-rays 100000 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['value'],1), 'Mrays/s  kernel_ms', round(d['roofline']['kernel_ms'],3))"
done
