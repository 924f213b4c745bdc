#!/bin/bash
# A/B of the shade kernel's gather order: IGB200_WAVE_SORT = 0 (no sort) / 1 (global index sort by
# material) / 2 (tile-local index sort: the gather stays inside a 2048-ray tile, i.e. in L2) on C4, C5 and the office scene.
for rep in 1 2; do
for sort in 0 1 2; do
  for fx in diamond many-point-lights office; do
    case $fx in diamond) res=1920; spi=4; batch=32;; many-point-lights) res=3840; spi=1; batch=32;; office) res=256; spi=64; batch=64;; esac
    echo -n "sort=$sort $fx: "; IGB200_WAVE_SORT=$sort FIXTURE=$fx BATCH=$batch python scripts/gpu_render_perf.py $res $spi 2 | sed 's/.*: //'
  done
done
done
