"""Parity of the sm_100a traversal kernels against the oracle, through the C ABI.  Needs a B200.

Contract (BASELINE.json north_star): closest-hit entity/primitive IDs bit-exact on non-degenerate rays,
hit distance within 1e-5 relative (here: bit-exact whenever the same triangle wins).  A ray is degenerate
when the brute-force intersector finds a second primitive within 1e-6 relative of the winner (exact ties
resolve by tree order in the reference, src/artic/traversal/mapping_cpu.art:469) -- those are counted,
reported and excluded; they must stay below 1e-4 of the batch.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
FLT_MAX = np.finfo(np.float32).max


@pytest.fixture(scope="module")
def dev(ig):
    d = ig.Device(0, acquire_stats=True)
    yield d
    d.close()


def _compare(ig, oracle, dev, tables, rays, max_degenerate=1e-4):
    dev.assign_scene(tables)
    ent, prim, t, u, v = dev.trace(rays)
    osc = oracle.OracleScene(tables)
    oent, oprim, ot, ou, ov = osc.trace(rays)
    same = (ent == oent) & (prim == oprim)
    bad = np.flatnonzero(~same)
    degenerate = 0
    if bad.size:
        assert bad.size <= max(4, max_degenerate * rays.size), f"{bad.size} id mismatches"
        # Classify every mismatch with the BVH-free intersector.  Accepted only if
        #  (a) the ray is a tie / near-tie (second-best t within 1e-6 relative: degenerate by the contract), or
        #  (b) the GPU result IS the brute-force result (ids and t bit-exact) -- then it is the oracle's BVH that culled a
        #      hit its own triangle test accepts (box-test rounding), not the GPU that is wrong.
        bent, bprim, bt, _, _, tie, second = osc.brute(rays[bad])
        gap = np.abs(second - bt) <= 1e-6 * np.abs(bt)
        gpu_is_brute = (ent[bad] == bent) & (prim[bad] == bprim) & ((t[bad] == bt) | (bprim < 0))
        assert np.all(tie | gap | gpu_is_brute), "GPU differs from the oracle AND from brute force on a non-degenerate ray"
        # a (near-)tie must still be a hit on one of the tied primitives: same t to 1e-6
        near = (tie | gap) & ~gpu_is_brute
        assert np.all(np.abs(t[bad][near] - bt[near]) <= 1e-6 * np.abs(bt[near])), "tie resolved to a different distance"
        degenerate = int(np.count_nonzero(tie | gap))
    _compare.last = dict(mismatches=int(bad.size), degenerate=degenerate, rays=int(rays.size))
    hit = same & (prim >= 0)
    assert np.array_equal(t[hit], ot[hit]), "t not bit-identical for the same triangle"
    assert np.array_equal(u[hit], ou[hit]) and np.array_equal(v[hit], ov[hit])
    miss = same & (prim < 0)
    assert np.all(ent[miss] == -1) and np.array_equal(t[miss], ot[miss])
    return same.mean(), hit.sum()


def test_c1_single_triangle_1m_rays(ig, oracle, dev):
    tables = ig.workloads.c1_scene()
    rays = ig.workloads.c1_rays(1_000_000)
    frac, hits = _compare(ig, oracle, dev, tables, rays)
    assert frac == 1.0 and 110_000 < hits < 140_000


@pytest.mark.parametrize("flat", [False, True])
def test_c3_small_instanced_and_flat(ig, oracle, dev, flat):
    tables = ig.workloads.c3_scene(grid=2, stacks=16, slices=32, flat=flat)
    rays = ig.workloads.c3_rays(tables, n=200_000, seed=0x77)
    frac, hits = _compare(ig, oracle, dev, tables, rays)
    assert frac > 0.9999 and hits > 20_000


def test_c3_full_scene_2m_rays(ig, oracle, dev):
    tables = ig.workloads.c3_scene()
    rays = ig.workloads.c3_rays(tables, n=1 << 21)
    frac, hits = _compare(ig, oracle, dev, tables, rays)
    assert frac > 0.9999
    st = dev.trace_stats()
    assert st["rays"] == rays.size and st["nodes_visited"] > rays.size and st["tris_tested"] > 0


def test_c3_benchmark_batch_itself_16m_rays(ig, oracle, dev, capsys):
    """The batch bench.py times (BASELINE.json configs[2]: 16 777 216 rays, seed 0xB200, generation order) against the
    oracle, closest-hit and any-hit.  Prints the mismatch / degenerate counts DESIGN.md section 4 quotes."""
    tables = ig.workloads.c3_scene()
    rays = ig.workloads.c3_rays(tables)
    assert rays.size == 1 << 24
    frac, hits = _compare(ig, oracle, dev, tables, rays)
    info = dict(_compare.last, hits=int(hits))
    aprim = dev.trace(rays, any_hit=True)[1]
    oprim = oracle.OracleScene(tables).trace(rays, any_hit=True)[1]
    occl = int(np.count_nonzero((aprim >= 0) != (oprim >= 0)))
    info["any_hit_occlusion_mismatches"] = occl
    with capsys.disabled():
        print(f"\nC3 16Mi parity: {info}")
    assert frac > 0.99999 and occl <= 4    # a grazing ray may be occluded by one side of a tie only


@pytest.mark.parametrize("flat", [False, True])
def test_c3_dense_2m_rays(ig, oracle, dev, flat):
    """bench.py's c3-dense variants: rays aimed at the spheres (hit rate ~1, deep traversal) against the oracle, closest-hit
    ids / t / u / v and any-hit occlusion, instanced and flattened scene."""
    tables = ig.workloads.c3_scene(flat=flat)
    rays = ig.workloads.c3_dense_rays(tables, n=1 << 21)
    frac, hits = _compare(ig, oracle, dev, tables, rays)
    assert frac > 0.9999 and hits > 0.99 * rays.size
    st = dev.trace_stats()
    assert st["nodes_visited"] > 8 * rays.size     # several times deeper than C3's 5 node steps per ray
    aprim = dev.trace(rays, any_hit=True)[1]
    oprim = oracle.OracleScene(tables).trace(rays, any_hit=True)[1]
    assert np.count_nonzero((aprim >= 0) != (oprim >= 0)) <= 2


def test_c3_full_size_instanced_equals_flattened(ig, dev):
    """Size-independent property at BASELINE.json's full scene size: the 64-instance scene and the same 1 032 192
    triangles baked to world space (SURVEY 8(d) 'C3-flat') are one geometry, so entity e / primitive p of the
    two-level traversal must be primitive e * 16128 + p of the single 1 M-triangle BVH.  t differs only by the f32
    rounding of the baked vertices (bar: 99.9 % within 1e-4 relative, median < 2e-6)."""
    inst = ig.workloads.c3_scene()
    flat = ig.workloads.c3_scene(flat=True)
    rays = ig.workloads.c3_rays(inst, n=1 << 22)
    dev.assign_scene(inst)
    ent, prim, t, _, _ = dev.trace(rays)
    aprim = dev.trace(rays, any_hit=True)[1]
    dev.assign_scene(flat)
    fent, fprim, ft, _, _ = dev.trace(rays)
    faprim = dev.trace(rays, any_hit=True)[1]
    per_shape = 2 * 64 * 128 - 2 * 128  # triangles of one uv sphere: 16 128
    assert inst.entity_count == 64 and flat.entity_count == 1
    hit, fhit = prim >= 0, fprim >= 0
    assert np.count_nonzero(hit != fhit) <= 4e-5 * rays.size  # grazing rays on the silhouette
    both = hit & fhit
    same = (ent[both].astype(np.int64) * per_shape + prim[both]) == fprim[both]
    assert same.mean() > 0.9999
    rel = np.abs(t[both] - ft[both]) / np.maximum(np.abs(t[both]), 1e-6)
    # grazing hits amplify the vertex rounding (t of a near-tangent ray moves by up to ~1e-2 relative on the same triangle)
    assert np.median(rel) < 2e-6 and np.quantile(rel[same], 0.999) < 1e-4 and rel[same].max() < 0.05
    assert np.count_nonzero((aprim >= 0) != (faprim >= 0)) <= 4e-5 * rays.size
    assert np.array_equal(aprim >= 0, hit)  # any-hit occlusion == closest-hit hit/miss on the same scene
    dev.assign_scene(inst)


def test_against_brute_force(ig, oracle, dev):
    tables = ig.workloads.c3_scene(grid=2, stacks=12, slices=24)
    rays = ig.workloads.c3_rays(tables, n=5000, seed=0x99)
    dev.assign_scene(tables)
    ent, prim, t, u, v = dev.trace(rays)
    bent, bprim, bt, bu, bv, tie, _ = oracle.OracleScene(tables).brute(rays)
    ok = (ent == bent) & (prim == bprim)
    assert np.all(ok | tie)
    sel = ok & (prim >= 0)
    assert np.array_equal(t[sel], bt[sel]) and np.array_equal(u[sel], bu[sel]) and np.array_equal(v[sel], bv[sel])


def test_any_hit_occlusion(ig, oracle, dev):
    tables = ig.workloads.c3_scene(grid=2, stacks=16, slices=32)
    rays = ig.workloads.c3_rays(tables, n=200_000, seed=0x78)
    dev.assign_scene(tables)
    _, prim, _, _, _ = dev.trace(rays)
    aent, aprim, at, _, _ = dev.trace(rays, any_hit=True)
    _, oprim, _, _, _ = oracle.OracleScene(tables).trace(rays, any_hit=True)
    assert np.array_equal(aprim >= 0, prim >= 0)
    assert np.array_equal(aprim >= 0, oprim >= 0)
    assert np.all(at[aprim < 0] == FLT_MAX)


def test_ray_range_and_visibility(ig, oracle, dev):
    b = ig.SceneBuilder()
    s = b.add_trimesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    b.add_entity(s, flags=0x1)
    m = np.eye(4); m[2, 3] = -1.0
    b.add_entity(s, m, flags=0xF)
    tables = b.finish()
    dev.assign_scene(tables)
    osc = oracle.OracleScene(tables)
    rays = ig.make_rays([[0.2, 0.2, 1]] * 4, [[0, 0, -1]] * 4)
    rays["tmin"] = [0, 1.5, 0, 2.5]
    rays["tmax"] = [FLT_MAX, FLT_MAX, 0.5, FLT_MAX]
    for flags in (0x0, 0x1, 0x8):
        got = dev.trace(rays, ray_flags=flags)
        want = osc.trace(rays, ray_flags=flags)
        for g, w in zip(got, want):
            assert np.array_equal(g, w), (flags, g, w)
    ent = dev.trace(rays, ray_flags=0x8)[0]
    assert ent.tolist() == [1, 1, -1, -1]


def test_sphere_and_mixed_scene(ig, oracle, dev):
    b = ig.SceneBuilder()
    v, f, n = ig.workloads.uv_sphere(8, 16)
    mesh = b.add_trimesh(v, f, n)
    sph = b.add_sphere((0, 0, 0), 1.0)
    m1 = np.eye(4); m1[:3, 3] = (-1.5, 0, 0)
    m2 = np.eye(4); m2[:3, 3] = (1.5, 0, 0); m2[0, 0] = m2[1, 1] = m2[2, 2] = 0.5
    b.add_entity(mesh, m1)
    b.add_entity(sph, m2)
    tables = b.finish()
    rays = ig.workloads.c3_rays(tables, n=50_000, seed=0x5)
    dev.assign_scene(tables)
    ent, prim, t, u, v = dev.trace(rays)
    oent, oprim, ot, ou, ov = oracle.OracleScene(tables).trace(rays)
    assert np.array_equal(ent, oent) and np.array_equal(prim, oprim)
    assert (ent == 1).sum() > 100
    tri = ent == 0
    assert np.array_equal(t[tri], ot[tri]) and np.array_equal(u[tri], ou[tri])
    sp = ent == 1
    np.testing.assert_allclose(t[sp], ot[sp], rtol=1e-5)      # sqrt/div chain: not bit-pinned across libm/CUDA
    np.testing.assert_allclose(u[sp], ou[sp], atol=1e-5)
    np.testing.assert_allclose(v[sp], ov[sp], atol=1e-5)


def test_edge_cases_empty_and_ragged(ig, oracle, dev):
    tables = ig.workloads.c1_scene()
    dev.assign_scene(tables)
    ent, prim, t, u, v = dev.trace(np.zeros(0, ig.tables.RAY_DTYPE))
    assert ent.size == 0
    for n in (1, 31, 33, 127, 129):  # ragged vs warp/block size
        rays = ig.workloads.c1_rays(n)
        got = dev.trace(rays)
        want = oracle.OracleScene(tables).trace(rays)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
    # unnormalised directions are normalised on upload like Interface::loadRayList
    rays = ig.make_rays([[0.2, 0.2, 1]], [[0, 0, -5]])
    assert dev.trace(rays)[2][0] == pytest.approx(1.0)
    # empty scene: everything misses
    dev.assign_scene(ig.SceneBuilder().finish())
    ent, prim, t, _, _ = dev.trace(ig.workloads.c1_rays(100))
    assert np.all(ent == -1) and np.all(prim == -1) and np.all(t == FLT_MAX)


def test_chunked_pipeline_schedules_agree(ig, dev, monkeypatch):
    """igb200_trace_rays with many small chunks and a ramp (IGB200_TRACE_CHUNK / _MIN), with downloads queued at upload
    boundaries (default) and right behind the traversal (IGB200_PIPE_ALIGN=0), a column subset, a ragged ray count:
    all identical to the one-chunk call."""
    tables = ig.workloads.c3_scene(grid=2, stacks=16, slices=32)
    rays = ig.workloads.c3_rays(tables, n=300_001, seed=0x51)
    dev.assign_scene(tables)
    want = dev.trace(rays)  # default chunk size: one chunk
    assert int((want[1] >= 0).sum()) > 1000
    monkeypatch.setenv("IGB200_TRACE_CHUNK", "16384")
    monkeypatch.setenv("IGB200_TRACE_CHUNK_MIN", "4096")
    for align in ("1", "0"):
        monkeypatch.setenv("IGB200_PIPE_ALIGN", align)
        with ig.Device(0) as d:
            d.assign_scene(tables)
            for n in (rays.size, 40_000, 4097):
                got = d.trace(rays[:n])
                for g, w in zip(got, want):
                    assert np.array_equal(g, w[:n]), (align, n)
            ent = np.full(rays.size, 7, np.int32)
            t = np.full(rays.size, 7, np.float32)
            d.trace_into(rays, ent, None, t, None, None)  # only two of the five columns
            assert np.array_equal(ent, want[0]) and np.array_equal(t, want[2])


def test_trace_before_scene_is_an_error(ig):
    with ig.Device(0) as d:
        with pytest.raises(ig._lib.BackendError) as e:
            d.trace(ig.workloads.c1_rays(4))
        assert e.value.code == -5


def test_device_resident_call_matches_host_call(ig, dev):
    import torch
    tables = ig.workloads.c3_scene(grid=2, stacks=16, slices=32)
    rays = ig.workloads.c3_rays(tables, n=100_000, seed=0x11)
    dev.assign_scene(tables)
    ent, prim, t, u, v = dev.trace(rays)
    d_rays = torch.from_numpy(rays.view(np.float32).reshape(-1, 8).copy()).cuda()
    n = rays.size
    d_ent = torch.empty(n, dtype=torch.int32, device="cuda")
    d_prim = torch.empty(n, dtype=torch.int32, device="cuda")
    d_t, d_u, d_v = (torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(3))
    s = torch.cuda.current_stream()
    dev.trace_device(d_rays.data_ptr(), n, d_ent.data_ptr(), d_prim.data_ptr(), d_t.data_ptr(), d_u.data_ptr(),
                     d_v.data_ptr(), stream=s.cuda_stream)
    s.synchronize()
    # the device-resident call does not re-normalise directions, so allow last-bit differences in t
    same = (d_ent.cpu().numpy() == ent) & (d_prim.cpu().numpy() == prim)
    assert same.mean() > 0.9999
    np.testing.assert_allclose(d_t.cpu().numpy()[same], t[same], rtol=1e-5)


def test_hostile_rays_and_degenerate_triangles(ig, oracle, dev):
    """Inputs a production ray list will contain sooner or later: axis-aligned and zero directions, empty and inverted
    ranges, NaN / infinite components, huge and denormal values, and a mesh with zero-area triangles.  The kernel must
    terminate, agree with the oracle wherever the reference's arithmetic defines an answer, and report a miss for
    rays whose components are not finite."""
    rng = np.random.default_rng(77)
    W = ig.workloads
    v, f, n = W.uv_sphere(12, 24)
    # zero-area triangles: repeated vertex, collinear vertices
    v = np.concatenate([v, [[0, 0, 2], [0, 0, 2], [1, 0, 2], [2, 0, 2]]]).astype(np.float32)
    nv = v.shape[0]
    f = np.concatenate([f, [[nv - 4, nv - 3, nv - 2], [nv - 4, nv - 2, nv - 1]]]).astype(np.uint32)
    n = np.concatenate([n, np.tile([[0, 0, 1]], (4, 1))]).astype(np.float32)
    b = ig.SceneBuilder()
    shape = b.add_trimesh(v, f, n)
    b.add_entity(shape)
    m = np.eye(4)
    m[:3, 3] = [3, 0, 0]
    b.add_entity(shape, m)
    tables = b.finish()
    dev.assign_scene(tables)
    osc = oracle.OracleScene(tables)

    # ---- defined cases: must equal the oracle exactly
    org = rng.uniform(-4, 6, (4000, 3)).astype(np.float32)
    d = rng.normal(size=(4000, 3)).astype(np.float32)
    d[:1000, 0] = 0                      # axis-aligned in one, two ...
    d[1000:2000, 0:2] = 0
    d[1000:2000, 2] = np.where(rng.random(1000) < 0.5, 1, -1)
    org[1000:1500, 0:2] = 0              # ... straight through the poles / the sphere centre
    d[2000:2100] *= 1e-20                # tiny but non-zero directions (normalised on upload)
    d[2100:2200] *= 1e18
    rays = ig.make_rays(org, d, 0.0, FLT_MAX)
    rays["tmin"][2200:2400] = 5.0        # window that may start behind the first hit
    rays["tmax"][2400:2600] = 0.5        # window that may end before the first hit
    rays["tmin"][2600:2700], rays["tmax"][2600:2700] = 3.0, 1.0    # inverted: nothing can be hit
    got, want = dev.trace(rays), osc.trace(rays)
    same = (got[0] == want[0]) & (got[1] == want[1])
    groups = {"one zero component": slice(0, 1000), "through the poles": slice(1000, 1500), "axis-aligned": slice(1500, 2000),
              "tiny directions": slice(2000, 2100), "huge directions": slice(2100, 2200), "late tmin": slice(2200, 2400),
              "early tmax": slice(2400, 2600), "inverted range": slice(2600, 2700), "plain": slice(2700, 4000)}
    report = {k: int(np.count_nonzero(~same[g])) for k, g in groups.items()}
    # rays through a pole hit the vertex shared by a whole fan of triangles: an exact-t tie, resolved by primitive id here and by
    # leaf order in the reference -- there the ids may differ, hit / miss and t may not
    pole = groups["through the poles"]
    assert np.array_equal(got[1][pole] >= 0, want[1][pole] >= 0), report
    assert np.array_equal(got[2][pole], want[2][pole]), report
    # directions whose NORM is below flt_eps are not normalised on upload (Device.cpp:588-592), and safe_rcp (core/common.art:95-98)
    # then treats every component (< 1e-8) as zero: the reference's BVH traversal -- and the oracle with it -- misses what its own
    # triangle test would hit.  The kernel's box test keeps working down to |d| ~ 2^-80, so for this degenerate group it is
    # held to the BVH-free brute-force intersector (same Moeller-Trumbore arithmetic) instead.
    tiny = groups["tiny directions"]
    brute = osc.brute(rays[tiny])
    assert np.array_equal(got[0][tiny], brute[0]) and np.array_equal(got[1][tiny], brute[1]) and np.array_equal(got[2][tiny], brute[2]), report
    assert np.count_nonzero(brute[1] >= 0) > 0 and np.all(want[1][tiny] == -1)
    others = np.ones(rays.size, bool)
    others[pole] = False
    others[tiny] = False
    assert np.count_nonzero(~same & others) == 0, report
    hit = same & (got[1] >= 0)
    assert hit.any() and np.array_equal(got[2][hit], want[2][hit])
    assert np.all(got[1][2600:2700] == -1)
    agot, awant = dev.trace(rays, any_hit=True), osc.trace(rays)
    defined = np.ones(rays.size, bool)
    defined[tiny] = False
    assert np.array_equal((agot[1] >= 0)[defined], (awant[1] >= 0)[defined])

    # ---- undefined cases: non-finite components and the zero direction terminate as misses (zero: norm -> 1, dir stays 0)
    bad = ig.make_rays(np.zeros((8, 3)), np.tile([[0, 0, 1]], (8, 1)), 0.0, FLT_MAX)
    bad["org"][0, 0] = np.nan
    bad["dir"][1, 2] = np.nan
    bad["org"][2, 1] = np.inf
    bad["dir"][3] = [np.inf, 0, 0]
    bad["dir"][4] = [0, 0, 0]
    bad["tmin"][5] = np.nan
    bad["tmax"][6] = np.nan
    bad["org"][7] = [1e30, -1e30, 1e30]
    for any_hit in (False, True):
        ent, prim, t, _, _ = dev.trace(bad, any_hit=any_hit)
        assert np.all(prim[[0, 1, 2, 3, 4, 7]] == -1) and np.all(ent[[0, 1, 2, 3, 4, 7]] == -1)


def test_device_resident_launches_on_two_streams_are_ordered(ig):
    """igb200_trace_rays_device from two CUDA streams on ONE handle: the kernel's work counters belong to the handle, so the
    library orders the launches on the device (VERDICT r1 'weak' 16: silent wrong answers before).  Results of interleaved
    launches on two streams must equal those of the same launches on one stream."""
    import torch
    tables = ig.workloads.c3_scene(grid=2, stacks=16, slices=32)
    n = 1 << 18
    batches = [ig.workloads.c3_rays(tables, n=n, seed=s) for s in (1, 2)]
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        d_rays = [torch.from_numpy(b.view(np.float32).reshape(-1, 8)).cuda() for b in batches]
        def outs():
            return [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(2)] + [torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(3)]
        ref = [outs(), outs()]
        s0 = torch.cuda.current_stream()
        for k in range(2):
            dev.trace_device(d_rays[k].data_ptr(), n, *[x.data_ptr() for x in ref[k]], stream=s0.cuda_stream)
        torch.cuda.synchronize()
        streams = [torch.cuda.Stream(), torch.cuda.Stream()]
        got = [outs(), outs()]
        for rep in range(8):  # back-to-back launches alternating between the two streams
            for k in range(2):
                dev.trace_device(d_rays[k].data_ptr(), n, *[x.data_ptr() for x in got[k]], stream=streams[k].cuda_stream)
        torch.cuda.synchronize()
        for k in range(2):
            for a, b in zip(ref[k], got[k]):
                assert torch.equal(a, b)
        assert int((ref[0][1] >= 0).sum()) > 0


def test_bvh_disk_cache_round_trip(ig, tmp_path, monkeypatch):
    """IGB200_CACHE_DIR: the BVH8 of a mesh above 500 000 faces is stored on the first igb200_assign_scene and loaded on the next
    (the counterpart of setup_bvh's cache, src/runtime/shape/TriMeshProvider.cpp:329-353).  Loaded and built trees give the same
    hits; a damaged cache file is ignored and rewritten."""
    import glob, time
    tables = ig.workloads.c3_scene(flat=True)  # one mesh of 1 032 192 triangles
    rays = ig.workloads.c3_rays(tables, n=1 << 18, seed=11)
    with ig.Device(0) as dev:  # no cache
        dev.assign_scene(tables)
        ref = dev.trace(rays)
    monkeypatch.setenv("IGB200_CACHE_DIR", str(tmp_path))
    times = []
    for _ in range(2):
        with ig.Device(0) as dev:
            t0 = time.perf_counter()
            dev.assign_scene(tables)
            times.append(time.perf_counter() - t0)
            got = dev.trace(rays)
        for a, b in zip(ref, got):
            assert np.array_equal(a, b)
    files = glob.glob(str(tmp_path / "_bvh8_*.bin"))
    assert len(files) == 1 and os.path.getsize(files[0]) > 1 << 20
    print(f"\nBVH cache: build + store {times[0]:.2f} s, load {times[1]:.2f} s")
    raw = bytearray(open(files[0], "rb").read())
    raw[len(raw) // 2:] = b"\xff" * (len(raw) - len(raw) // 2)  # prim indices out of range
    open(files[0], "wb").write(bytes(raw))
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        got = dev.trace(rays)
    for a, b in zip(ref, got):
        assert np.array_equal(a, b)
