"""dev tool: who is right on the tiny-direction rays of tests/test_trace_gpu.py::test_hostile_rays_and_degenerate_triangles?"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
ig = importlib.import_module("ignis-masterthesis_b200")
import pyoracle; pyoracle.load()
W = ig.workloads
v, f, n = W.uv_sphere(12, 24)
b = ig.SceneBuilder(); shape = b.add_trimesh(v, f, n); b.add_entity(shape)
m = np.eye(4); m[:3, 3] = [3, 0, 0]; b.add_entity(shape, m)
tables = b.finish()
rng = np.random.default_rng(5)
N = 200000
org = rng.uniform(-4, 6, (N, 3)).astype(np.float32)
for scale in (1.0, 1e-10, 1e-18, 1e-20, 1e-22):
    d = (rng.normal(size=(N, 3)) * scale).astype(np.float32)
    rays = ig.make_rays(org, d, 0.0, np.finfo(np.float32).max)
    osc = pyoracle.OracleScene(tables)
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        g = dev.trace(rays)
    o = osc.trace(rays)
    br = osc.brute(rays)
    bad = np.flatnonzero((g[0] != o[0]) | (g[1] != o[1]))
    gpu_vs_brute = np.count_nonzero((g[0] != br[0]) | (g[1] != br[1]))
    ora_vs_brute = np.count_nonzero((o[0] != br[0]) | (o[1] != br[1]))
    print(f"scale {scale:g}: gpu!=oracle {bad.size}, gpu!=brute {gpu_vs_brute}, oracle!=brute {ora_vs_brute}")
    for i in bad[:4]:
        print("   ray", rays[i], "gpu", [x[i] for x in g[:3]], "oracle", [x[i] for x in o[:3]], "brute", [x[i] for x in br[:3]])
