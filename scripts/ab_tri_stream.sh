#!/bin/bash
# A/B of the triangle cache policy (DeviceScene::tri_stream): IGB200_TRI_STREAM_BYTES = 0 (records always bypass L1 allocation),
# unset (default threshold, 256 KiB of triangle data), 1000000000000 (never) on the trace workloads and the three path-tracer scenes.
for rep in 1 2; do
for mode in always default never; do
  case $mode in always) export IGB200_TRI_STREAM_BYTES=0;; default) unset IGB200_TRI_STREAM_BYTES;; never) export IGB200_TRI_STREAM_BYTES=1000000000000;; esac
  python bench.py --steps 20 --pt-iterations 0 --c5-spp 0 --office-spp 0 --cpu-rays 100000 --variant-steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); v=d['trace_variants']
print('$mode', 'c3', round(d['value']), 'any', round(v['c3-instanced any-hit']['value']), 'flat', round(v['c3-flat closest-hit']['value']), round(v['c3-flat any-hit']['value']), 'dense', round(v['c3-dense closest-hit']['value']), round(v['c3-dense any-hit']['value']), 'dense-flat', round(v['c3-dense-flat closest-hit']['value']), round(v['c3-dense-flat any-hit']['value']))"
  for fx in diamond many-point-lights office; do
    case $fx in diamond) res=1920; spi=4; batch=32;; many-point-lights) res=3840; spi=1; batch=32;; office) res=256; spi=64; batch=64;; esac
    echo -n "$mode $fx: "; FIXTURE=$fx BATCH=$batch python scripts/gpu_render_perf.py $res $spi 2 | sed 's/.*: //'
  done
done
done
