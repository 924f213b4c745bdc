"""Pins the oracle (CPU restatement of the reference traversal) -- runs without a GPU.

(1) the reference's own known-answer tests, src/tests/artic/test_intersection.art:1-134;
(2) the BVH4/Tri4 traversal against the BVH-free brute-force intersector on random rays.
"""
import numpy as np
import pytest

FLT_MAX = np.finfo(np.float32).max


# ---- (1) test_intersection.art ---------------------------------------------------------------------
# make_tri(v0=(0,0,0), e1=(-1,0,0), e2=(0,1,0), n=(0,0,-1)): v1 = v0 - e1 = (1,0,0), v2 = v0 + e2 = (0,1,0)
TRI = [[0, 0, 0], [1, 0, 0], [0, 1, 0]]


def test_kat_triangle_front(oracle):
    hit, tuv = oracle.kat_ray_triangle(TRI, [0.2, 0.4, 1], [0, 0, -1])
    assert hit
    np.testing.assert_allclose(tuv, [1.0, 0.2, 0.4], rtol=1e-6)


def test_kat_triangle_back(oracle):
    # no back-face culling: the same triangle is hit from behind (test_intersection.art: o.z=-1, d=+z)
    hit, tuv = oracle.kat_ray_triangle(TRI, [0.2, 0.4, -1], [0, 0, 1])
    assert hit
    np.testing.assert_allclose(tuv, [1.0, 0.2, 0.4], rtol=1e-6)


def test_kat_triangle_miss(oracle):
    hit, _ = oracle.kat_ray_triangle(TRI, [0.2, 4.4, 1], [0, 0, -1])
    assert not hit


def test_kat_box_hit(oracle):
    hit, t = oracle.kat_ray_box([0, 0, 0, 1, 1, 1], [0.2, 0.4, 2], [0, 0, -1])
    assert hit and t == pytest.approx(1.0)


def test_kat_box_miss(oracle):
    hit, _ = oracle.kat_ray_box([0, 0, 0, 1, 1, 1], [0.2, 3.4, 2], [0, 0, -1])
    assert not hit


def test_kat_box_flat(oracle):
    hit, t = oracle.kat_ray_box([0, 0, 0, 1, 1, 0], [0.2, 0.4, 1], [0, 0, -1])
    assert hit and t == pytest.approx(1.0)


def test_triangle_edge_tolerance(oracle):
    # intersection.art:72: u,v,w >= -flt_eps, so a ray exactly on the shared edge hits
    hit, tuv = oracle.kat_ray_triangle(TRI, [0.5, 0.0, 1], [0, 0, -1])
    assert hit and tuv[2] == 0.0
    hit, _ = oracle.kat_ray_triangle(TRI, [0.5, -1e-3, 1], [0, 0, -1])
    assert not hit


def test_triangle_parallel_ray_rejected(oracle):
    # det == 0 -> inv_det = inf -> NaN comparisons reject (intersection.art:78-86)
    hit, _ = oracle.kat_ray_triangle(TRI, [0.2, 0.4, 1], [1, 0, 0])
    assert not hit


def test_triangle_range(oracle):
    assert not oracle.kat_ray_triangle(TRI, [0.2, 0.4, 1], [0, 0, -1], tmin=0.0, tmax=0.5)[0]
    assert not oracle.kat_ray_triangle(TRI, [0.2, 0.4, 1], [0, 0, -1], tmin=1.5, tmax=FLT_MAX)[0]
    assert oracle.kat_ray_triangle(TRI, [0.2, 0.4, 1], [0, 0, -1], tmin=1.0, tmax=1.0)[0]


# ---- (2) BVH4 traversal == brute force ---------------------------------------------------------------
def test_c1_hit_rate_and_ids(ig, oracle):
    tables = ig.workloads.c1_scene()
    rays = ig.workloads.c1_rays(n=100_000)
    osc = oracle.OracleScene(tables)
    ent, prim, t, u, v = osc.trace(rays)
    hit = prim >= 0
    assert 0.11 < hit.mean() < 0.14  # triangle covers 1/8 of the target square
    assert np.all(ent[hit] == 0) and np.all(prim[hit] == 0)
    assert np.all(ent[~hit] == -1) and np.all(t[~hit] == FLT_MAX)
    # analytic check: the hit point must be origin + t*dir on z = 0, inside the triangle
    r = osc.normalized(rays)
    p = r["org"][hit] + r["dir"][hit] * t[hit, None]
    assert np.abs(p[:, 2]).max() < 1e-6
    assert np.all(p[:, 0] >= -1e-6) and np.all(p[:, 1] >= -1e-6) and np.all(p[:, 0] + p[:, 1] <= 1 + 1e-6)
    np.testing.assert_allclose(u[hit], p[:, 0], atol=2e-6)
    np.testing.assert_allclose(v[hit], p[:, 1], atol=2e-6)


@pytest.mark.parametrize("flat", [False, True])
def test_bvh4_matches_brute_force(ig, oracle, flat):
    tables = ig.workloads.c3_scene(grid=2, stacks=12, slices=24, flat=flat)
    rays = ig.workloads.c3_rays(tables, n=4000, seed=0x51)
    osc = oracle.OracleScene(tables)
    ent, prim, t, u, v = osc.trace(rays)
    bent, bprim, bt, bu, bv, tie, _ = osc.brute(rays)
    ok = (ent == bent) & (prim == bprim)
    assert np.all(ok | tie)
    sel = ok & (prim >= 0)
    assert sel.sum() > 400
    assert np.array_equal(t[sel], bt[sel]) and np.array_equal(u[sel], bu[sel]) and np.array_equal(v[sel], bv[sel])


def test_any_hit_agrees_with_closest_hit_occlusion(ig, oracle):
    tables = ig.workloads.c3_scene(grid=2, stacks=12, slices=24)
    rays = ig.workloads.c3_rays(tables, n=4000, seed=0x52)
    osc = oracle.OracleScene(tables)
    _, prim, _, _, _ = osc.trace(rays)
    _, aprim, _, _, _ = osc.trace(rays, any_hit=True)
    assert np.array_equal(prim >= 0, aprim >= 0)


def test_visibility_flags(ig, oracle):
    # entity visible to camera rays only (LoaderEntity.cpp:122-131, ray.art:51)
    b = ig.SceneBuilder()
    s = b.add_trimesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    b.add_entity(s, flags=0x1)
    tables = b.finish()
    rays = ig.make_rays([[0.2, 0.2, 1]], [[0, 0, -1]])
    osc = oracle.OracleScene(tables)
    assert osc.trace(rays, ray_flags=0x1)[1][0] == 0      # camera ray sees it
    assert osc.trace(rays, ray_flags=0x8)[1][0] == -1     # shadow ray does not
    assert osc.trace(rays, ray_flags=0x0)[1][0] == 0      # igtrace rays carry no type bits


def test_sphere_entity(ig, oracle):
    b = ig.SceneBuilder()
    s = b.add_sphere((0, 0, 0), 0.5)
    m = np.eye(4); m[:3, 3] = (0, 0, -3); m[0, 0] = m[1, 1] = m[2, 2] = 2.0  # scaled: world radius 1
    b.add_entity(s, m)
    tables = b.finish()
    rays = ig.make_rays([[0, 0, 0], [0, 0, 0]], [[0, 0, -1], [0, 1, 0]])
    ent, prim, t, u, v = oracle.OracleScene(tables).trace(rays)
    assert prim[0] == 0 and ent[0] == 0 and t[0] == pytest.approx(2.0, rel=1e-6)
    assert prim[1] == -1


def test_empty_ray_list(ig, oracle):
    tables = ig.workloads.c1_scene()
    ent, prim, t, u, v = oracle.OracleScene(tables).trace(np.zeros(0, ig.tables.RAY_DTYPE))
    assert ent.size == 0


def test_normalization_matches_reference_rule(oracle, ig):
    rays = ig.make_rays([[0, 0, 0], [0, 0, 0]], [[0, 0, -2], [0, 0, 0]])
    r = oracle.OracleScene.normalized(rays)
    np.testing.assert_array_equal(r["dir"][0], [0, 0, -1])
    np.testing.assert_array_equal(r["dir"][1], [0, 0, 0])  # zero direction: norm forced to 1 (Device.cpp:589-592)
