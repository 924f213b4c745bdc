#!/bin/bash
# Row a25 A/B: builder variant -> Mrays/s and per-ray visit counts on C3 / C3-flat / c3-dense (closest- and any-hit), plus C4.
run() {
  python bench.py --steps 10 --pt-iterations 0 --c5-spp 0 --office-spp 0 --cpu-rays 100000 --variant-steps 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); v=d['trace_variants']
pr=d['roofline']['per_ray']
print('  c3 closest %.0f (nodes %.2f tris %.2f) build %.3fs nodes %.1f KB' % (d['value'], pr['nodes'], pr['tris'], d['config']['bvh']['build_s'], d['config']['bvh']['node_bytes']/1e3))
print('  c3-flat closest %.0f any %.0f  build %.2fs nodes %.1f MB' % (v['c3-flat closest-hit']['value'], v['c3-flat any-hit']['value'], v['c3-flat bvh']['build_s'], v['c3-flat bvh']['node_bytes']/1e6))
for k in ('c3-dense closest-hit','c3-dense any-hit','c3-dense-flat closest-hit','c3-dense-flat any-hit'):
    r=v[k]; print('  %-26s %.0f (nodes %.2f tris %.2f)' % (k, r['value'], r['per_ray']['nodes'], r['per_ray']['tris']))
"
  echo -n "  diamond "; FIXTURE=diamond BATCH=32 python scripts/gpu_render_perf.py 1920 4 2 | sed 's/.*: //'
}
echo "== greedy"; IGB200_BVH_COLLAPSE=greedy run
for cp in 0.2 0.4 0.7 1.0; do echo "== dp c_prim=$cp"; IGB200_BVH_COLLAPSE=dp IGB200_BVH_CPRIM=$cp run; done
