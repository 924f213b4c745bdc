"""The C++ scene loader of the host side (ignis-masterthesis_b200/host/SceneLoader.cpp, SceneMesh.cpp, PngReader.cpp, Json.h)
against the Python loader (sceneio.load_scene), which the golden-image tests pin to the reference: both read the same
scene JSON and must produce the same SceneDatabase blobs and shading tables.

Bar: byte-identical sections, with two stated exceptions -- (1) +0 / -0 where a matrix inverse comes out of a different
elimination order, (2) procedural spheres / disks / cones, where numpy's float32 sin / cos / acos differ from libm's by a
few ulp (the C++ side calls libm like the reference's std::sin on float): those words must agree within 32 ulp.
Scenes of the reference are read from /root/reference (build container only); the repository's own demo scene always runs."""
import importlib
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/scenes"
INT_SECTIONS = {"lookups", "material_count", "entity_per_material", "materials", "selector", "film", "technique"}


def _read_igdb(path):
    d = open(path, "rb").read()
    assert d[:8] == b"IGB200DB"
    _, n = struct.unpack("<II", d[8:16])
    out = {}
    for i in range(n):
        e = d[16 + 40 * i:56 + 40 * i]
        nbytes, off = struct.unpack("<QQ", e[24:40])
        out[e[:24].rstrip(b"\0").decode()] = d[off:off + nbytes]
    return out


def _compare(ig, host_bin, scene, tmp_path, max_ulp):
    tables, sh, w, h, _ = importlib.import_module("ignis-masterthesis_b200.sceneio").load_scene(scene)
    py, cc = str(tmp_path / "py.igdb"), str(tmp_path / "cc.igdb")
    ig.fixtures.save_igdb(py, tables, sh, w, h)
    r = subprocess.run([os.path.join(host_bin, "igdbtool"), "json2igdb", scene, cc], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    a, b = _read_igdb(py), _read_igdb(cc)
    assert set(a) == set(b)
    worst = 0
    for name in a:
        assert len(a[name]) == len(b[name]), name
        if a[name] == b[name]:
            continue
        assert name not in INT_SECTIONS, f"{name} differs"
        x, y = np.frombuffer(a[name], np.uint8), np.frombuffer(b[name], np.uint8)
        words = len(a[name]) // 4
        assert np.array_equal(x[4 * words:], y[4 * words:])
        ix, iy = x[:4 * words].view(np.int32), y[:4 * words].view(np.int32)
        d = np.flatnonzero(ix != iy)
        fx, fy = ix.view(np.float32)[d], iy.view(np.float32)[d]
        assert np.all(np.isfinite(fx)) and np.all(np.isfinite(fy)), f"{name}: a non-float word differs"
        nonzero = ~((fx == 0) & (fy == 0))  # +0 / -0
        # ulp distance through the sign-magnitude -> two's complement map, so that it is meaningful across zero
        key = lambda v: np.where(v < 0, np.int64(-(2 ** 31)) - v.astype(np.int64), v.astype(np.int64))
        ulp = np.abs(key(ix[d][nonzero]) - key(iy[d][nonzero]))
        near_zero = np.abs(fx[nonzero] - fy[nonzero]) <= 1e-6  # cos(pi / 2)-like values: absolute, not ulp
        bad = ulp[~near_zero]
        if bad.size:
            worst = max(worst, int(bad.max()))
    assert worst <= max_ulp, f"{scene}: {worst} ulp"
    return worst


@pytest.mark.parametrize("scene,max_ulp", [
    ("single_triangle.json", 0),              # BASELINE.json configs[0]
    ("primitives.json", 32),                  # configs[1]: every procedural primitive + subdivision
    ("diamond_scene.json", 0),                # configs[3]: PLY meshes, dielectric, area light as a plane
    ("many_point_lights.json", 0),            # configs[4]: PNG texture, bump map, rough conductor, sky bake + CDF, light hierarchy
    ("pillar.json", 0),
    ("evaluation/cbox-d6.json", 0),           # OBJ
    ("evaluation/room.json", 0),
    ("evaluation/multilight-hierarchy.json", 0),
    ("evaluation/multilight-simple.json", 0),
    ("evaluation/env.json", 0),               # textured environment: baked image + compensated 2-D CDF
    ("evaluation/sky-clear.json", 0),         # CIE sky, sun position from date / time / place
    ("evaluation/sky-intermediate.json", 0),
    ("evaluation/sky-perez1.json", 0),        # Perez model from irradiances
    ("evaluation/sphere-light-pure.json", 0),
    ("evaluation/sphere-light-ico.json", 32),
    ("evaluation/two-planes-mirror.json", 0),
    ("evaluation/flipped-prim-glass.json", 32),
    ("evaluation/emissive-plane-scale.json", 0),
    ("radiance/office/office_scene.json", 0),  # externals, fishlens camera, thin dielectric, qrotate
    ("showcase/light/cie_clear.json", 0),      # the underscore spellings of the CIE sky types (LoaderLight.cpp:90-97)
    ("showcase/light/cie_cloudy.json", 0),
    ("sun.json", 0),
    ("sky_fishlens.json", 0),
    ("perez.json", 0),
])
def test_reference_scenes_cpp_loader_equals_python_loader(ig, host_bin, tmp_path, scene, max_ulp):
    if not os.path.isdir(REF):
        pytest.skip("the reference tree exists in the build container only")
    _compare(ig, host_bin, os.path.join(REF, scene), tmp_path, max_ulp)


def _write_pyramid(path):
    verts = [(-0.5, -0.5, 0), (0.5, -0.5, 0), (0.5, 0.5, 0), (-0.5, 0.5, 0), (0, 0, 0.9)]
    faces = [(0, 1, 4), (1, 2, 4), (2, 3, 4), (3, 0, 4), (3, 2, 1, 0)]  # one quad: fan triangulation
    with open(path, "wb") as f:
        f.write(b"ply\nformat binary_little_endian 1.0\nelement vertex 5\nproperty float x\nproperty float y\nproperty float z\n"
                b"element face 5\nproperty list uchar int vertex_indices\nend_header\n")
        for v in verts:
            f.write(struct.pack("<3f", *v))
        for fc in faces:
            f.write(struct.pack("<B", len(fc)) + struct.pack(f"<{len(fc)}i", *fc))


@pytest.fixture()
def demo_scene(tmp_path):
    src = os.path.join(ROOT, "tests", "scenes", "loader_demo.json")
    dst = str(tmp_path / "loader_demo.json")
    open(dst, "w").write(open(src).read())
    _write_pyramid(str(tmp_path / "loader_demo_pyramid.ply"))
    return dst


def test_demo_scene_cpp_loader_equals_python_loader(ig, host_bin, tmp_path, demo_scene):
    _compare(ig, host_bin, demo_scene, tmp_path, 32)


def test_cpp_loader_error_behaviour(host_bin, tmp_path, demo_scene):
    tool = os.path.join(host_bin, "igdbtool")
    out = str(tmp_path / "out.igdb")
    text = open(demo_scene).read()
    cases = {
        "unknown_shape.json": text.replace('"type": "disk"', '"type": "teapot"'),
        "unknown_bsdf.json": text.replace('"type": "mirror"', '"type": "velvet"'),
        "missing_shape.json": text.replace('"shape": "Ball"', '"shape": "NoSuchShape"'),
        "broken.json": text[: len(text) // 2],
        "missing_mesh.json": text.replace("loader_demo_pyramid.ply", "no_such_file.ply"),
    }
    for name, body in cases.items():
        path = str(tmp_path / name)
        open(path, "w").write(body)
        r = subprocess.run([tool, "json2igdb", path, out], capture_output=True, text=True, timeout=60)
        assert r.returncode != 0 and "[ERROR]" in r.stderr, name
    r = subprocess.run([tool, "json2igdb", str(tmp_path / "absent.json"), out], capture_output=True, text=True, timeout=60)
    assert r.returncode != 0


def test_sun_position_cpp_equals_python(ig):
    import ctypes as C
    host = ig._lib.load_host()
    for args in ((2020, 5, 6, 12, 0, 0.0, -6.9965744, 49.235422, -2.0), (2022, 11, 18, 13, 0, 0.0, -6.9965744, 49.235422, -2.0),
                 (2000, 1, 1, 8, 30, 15.0, 13.4, 52.5, -1.0), (2031, 12, 31, 23, 59, 59.0, 151.2, -33.9, -10.0)):
        el, az = C.c_float(), C.c_float()
        y, mo, d, h, mi, s, lon, lat, tz = args
        host.ig_b200_sun_ea(y, mo, d, h, mi, C.c_float(s), C.c_float(lon), C.c_float(lat), C.c_float(tz), C.byref(el), C.byref(az))
        pel, paz = importlib.import_module("ignis-masterthesis_b200.sceneio").compute_sun_ea(*args)
        assert el.value == float(pel) and az.value == float(paz), args
