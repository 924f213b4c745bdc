"""Row a25: SAH cost and build time of the BVH8 builder variants (host only, no GPU): C3 shape (16 128 triangles), C3-flat
(1 032 192 triangles) and the diamond mesh.  usage: python scripts/bvh_quality.py [threads...]"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ig = importlib.import_module("ignis-masterthesis_b200")
lib = ig._lib.load()


def boxes_of(tables, shape=0):
    off = int(tables.lookups["offset"][shape])
    hdr = tables.shape_data[off:off + 16].view(np.uint32)
    nf, nv, nn = int(hdr[0]), int(hdr[1]), int(hdr[2])
    fl = tables.shape_data[off:].view(np.float32)
    verts = fl[12:12 + 4 * nv].reshape(nv, 4)[:, :3]
    idx = tables.shape_data[off:].view(np.uint32)[12 + 4 * nv + 4 * nn:12 + 4 * nv + 4 * nn + 4 * nf].reshape(nf, 4)[:, :3]
    tri = verts[idx]
    return np.ascontiguousarray(np.concatenate([tri.min(axis=1), tri.max(axis=1)], axis=1), np.float32)


def quality(boxes):
    out = np.zeros(4, np.float64)
    assert lib.igb200_debug_bvh8_quality(boxes.ctypes.data, boxes.shape[0], 3, out.ctypes.data) == 0
    return out


sets = {"c3 shape (16128 tris)": boxes_of(ig.workloads.c3_scene()), "c3-flat (1032192 tris)": boxes_of(ig.workloads.c3_scene(flat=True))}
fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
t, _, _, _, _ = fx.load_fixture(os.path.join(ROOT, "tests", "golden", "diamond.npz"))
big = max(range(t.lookups.size), key=lambda s: int(t.shape_data[int(t.lookups["offset"][s]):][:4].view(np.uint32)[0]))
sets["diamond mesh"] = boxes_of(t, big)
threads = sys.argv[1:] or ["1", str(min(16, os.cpu_count() or 1))]
print(f"{'mesh':26s} {'collapse':7s} {'thr':>3s} {'nodes':>8s} {'sum A_n':>9s} {'sum A_l N_l':>11s} {'SAH(1,0.7)':>10s} {'build ms':>9s}")
for name, b in sets.items():
    for collapse in ("greedy", "dp"):
        for th in threads:
            os.environ["IGB200_BVH_COLLAPSE"], os.environ["IGB200_BVH_THREADS"] = collapse, th
            q = min((quality(b) for _ in range(3)), key=lambda o: o[3])
            print(f"{name:26s} {collapse:7s} {th:>3s} {int(q[2]):8d} {q[0]:9.3f} {q[1]:11.3f} {q[0] + 0.7 * q[1]:10.3f} {q[3]:9.1f}")
