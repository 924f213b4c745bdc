#!/bin/bash
# e2e Mrays/s of igb200_trace_rays (host buffers) for several pipeline chunk schedules: "<bulk> <first/last>"
for c in "1048576 1048576" "1048576 131072" "1048576 65536" "2097152 131072" "524288 65536" "1048576 262144"; do
  set -- $c
  IGB200_TRACE_CHUNK=$1 IGB200_TRACE_CHUNK_MIN=$2 python bench.py --steps 5 --pt-iterations 0 --c5-spp 0 --office-spp 0 --cpu-rays 100000 --variant-steps 0 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('bulk $1 ramp-from $2', 'e2e', round(d['e2e']['value'],1), 'Mrays/s  kernel', round(d['value'],1))"
done
