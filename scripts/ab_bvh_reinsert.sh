#!/bin/bash
# A/B of the reinsertion pass of the binary builder (IGB200_BVH_REINSERT = passes) on C3 (closest / any, visit counts), its dense
# variant, and the three path-tracer scenes.  The flat 1 M-triangle mesh is above the pass's size limit and is unaffected.
for rep in 1 2; do
for passes in 0 1 2; do
  export IGB200_BVH_REINSERT=$passes
  python bench.py --steps 20 --pt-iterations 0 --c5-spp 0 --office-spp 0 --cpu-rays 100000 --variant-steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); v=d['trace_variants']; p=d['roofline']['per_ray']
print('reinsert $passes', 'c3', round(d['value']), '(nodes %.2f tris %.2f)' % (p['nodes'], p['tris']), 'any', round(v['c3-instanced any-hit']['value']), 'dense', round(v['c3-dense closest-hit']['value']), '(nodes %.2f tris %.2f)' % (v['c3-dense closest-hit']['per_ray']['nodes'], v['c3-dense closest-hit']['per_ray']['tris']), round(v['c3-dense any-hit']['value']), 'build_s %.3f' % d['config']['bvh']['build_s'])"
  for fx in diamond many-point-lights office; do
    case $fx in diamond) res=1920; spi=4; batch=32;; many-point-lights) res=3840; spi=1; batch=32;; office) res=256; spi=64; batch=64;; esac
    echo -n "reinsert $passes $fx: "; FIXTURE=$fx BATCH=$batch python scripts/gpu_render_perf.py $res $spi 2 | sed 's/.*: //'
  done
done
done
