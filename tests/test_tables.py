"""Host-side table builders reproduce the reference's SceneDatabase layouts."""
import numpy as np


def test_trimesh_table_layout(ig):
    b = ig.SceneBuilder()
    s0 = b.add_trimesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    s1 = b.add_trimesh([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0]], [[0, 1, 2], [1, 3, 2]])
    b.add_entity(s0)
    t = b.finish()
    assert (s0, s1) == (0, 1)
    # TriMeshProvider.cpp:577-598: header, bbox (inflated 1e-5), vec4 verts, vec4 normals, 4xu32 faces, uv
    d = t.shape_data
    hdr = d[:16].view(np.uint32)
    assert hdr.tolist() == [1, 3, 3, 3]
    bbox = d[16:48].view(np.float32)
    np.testing.assert_allclose(bbox, [-1e-5, -1e-5, -1e-5, 0, 1 + 1e-5, 1 + 1e-5, 1e-5, 0], atol=1e-7)
    size0 = 4 * (12 + 3 * 4 + 3 * 4 + 1 * 4) + 3 * 8
    # DynTable::addLookup pads by `16 - size % 16` (DynTable.h:18-21)
    assert t.lookups["offset"][1] == size0 + (16 - size0 % 16)
    faces = d[4 * (12 + 24):4 * (12 + 24) + 16].view(np.uint32)
    assert faces.tolist() == [0, 1, 2, 0]


def test_entity_row_and_leaf(ig):
    b = ig.SceneBuilder()
    s = b.add_trimesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    m = np.eye(4); m[:3, 3] = (1, 2, 3); m[0, 0] = 2
    e0 = b.add_entity(s, m, material="a", flags=0x5)
    e1 = b.add_entity(s, None, material="b")
    t = b.finish()
    assert (e0, e1) == (0, 1) and t.entity_count == 2 and t.material_count == 2
    row = t.entities[:36]
    # LoaderEntity.cpp:150-158: toLocal 3x4 col-major, toGlobal, normal 3x3, shape, material, pad
    to_global = row[12:24].reshape(4, 3).T
    np.testing.assert_allclose(to_global, m[:3])
    to_local = row[0:12].reshape(4, 3).T
    np.testing.assert_allclose(to_local @ np.append(m[:3, 3], 1), [0, 0, 0], atol=1e-6)
    assert row[33:36].view(np.uint32).tolist() == [0, 0, 0]
    leaf = t.leaves[0][0]
    assert leaf["flags"] == 0x5 and leaf["entity_id"] == 0 and leaf["shape_id"] == 0
    np.testing.assert_allclose(leaf["bbox_min"], [1 - 2e-5, 2 - 1e-5, 3 - 1e-5], atol=1e-6)
    np.testing.assert_allclose(leaf["bbox_max"], [3 + 2e-5, 3 + 1e-5, 3 + 1e-5], atol=1e-6)
    np.testing.assert_allclose(leaf["local"], row[0:12])
    assert t.entity_per_material == [1, 1]


def test_workload_determinism(ig):
    W = ig.workloads
    # splitmix64 reference values for seed 0 (public test vector of the generator)
    z = W.splitmix64(0, 0, 3)
    assert [int(x) for x in z] == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F]
    r1 = W.c1_rays(1000)
    r2 = W.c1_rays(2000)[:1000]
    assert r1.tobytes() == r2.tobytes()
    sc = W.c3_scene()
    assert sc.leaves[0].size == 64 and sc.lookups.size == 1
    a = W.c3_rays(sc, n=3000, chunk=1000)
    bb = W.c3_rays(sc, n=3000, chunk=4096)
    assert a.tobytes() == bb.tobytes()
    np.testing.assert_allclose(np.linalg.norm(a["dir"], axis=1), 1, atol=1e-6)
