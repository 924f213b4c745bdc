#!/bin/bash
# A/B of library variants (build/variants/lib<name>.so; "default" = the in-tree library) on the three path-tracer workloads:
# C4 diamond, C5 many-point-lights, office.   usage: scripts/ab_wave_variants.sh default v1 v2 ...
for rep in 1 2; do
for v in "$@"; do
  if [ "$v" = default ]; then unset IGB200_LIB; else export IGB200_LIB=$PWD/build/variants/lib$v.so; fi
  for fx in diamond many-point-lights office; do
    case $fx in diamond) res=1920; spi=4; batch=32;; many-point-lights) res=3840; spi=1; batch=32;; office) res=256; spi=64; batch=64;; esac
    echo -n "$v $fx: "; FIXTURE=$fx BATCH=$batch python scripts/gpu_render_perf.py $res $spi 2 | sed 's/.*: //'
  done
done
done
