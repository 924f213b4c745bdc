#!/bin/bash
# compute-sanitizer passes (SURVEY.md section 5 test plan): memcheck, racecheck, synccheck and initcheck over
# __graft_entry__.smoke() (closest-hit + any-hit trace and one path-traced Cornell-box iteration through the C ABI) and a
# small multi-round render.  Logs -> gpurun_out/<tag>_sanitizer_<tool>.log ; summary lines on stdout.
tag=${1:-r02}
mkdir -p gpurun_out
cat > /tmp/sanitize_target.py <<'PY'
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import __graft_entry__ as g
g.smoke()
pkg = importlib.import_module("ignis-masterthesis_b200")
fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
root = os.path.dirname(os.path.abspath(g.__file__))
for name, w, h in (("diamond", 96, 54), ("many-point-lights", 96, 54)):   # deep glass paths; extended shade kernel
    tables, sh, _, _, _ = fx.load_fixture(os.path.join(root, "tests", "golden", name + ".npz"))
    with pkg.Device(0) as dev:
        dev.assign_scene(tables); dev.assign_shading(sh); dev.resize(w, h)
        dev.render(w, h, spi=2, iteration=0, iterations=3)
        dev.render_partition(w, h, 2, 2, 1, 2, 2)           # tile-row partition path
        film, n = dev.framebuffer()
        print(name, "film mean", float(film.mean()), "iterations", n, dev.render_stats()["rounds"], "rounds")
PY
for tool in memcheck racecheck synccheck initcheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python /tmp/sanitize_target.py > gpurun_out/${tag}_sanitizer_${tool}.log 2>&1
  echo "$tool rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/${tag}_sanitizer_${tool}.log | tail -1)"
done
