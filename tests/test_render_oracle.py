"""The oracle's restated render iteration against the reference's own golden values (CPU only).

Pins:
  * src/tests/integrator/test_lights.py:5-44 -- closed-form scene averages of the flat scene;
  * scenes/evaluation/references/*.exr + the error metric / thresholds of scripts/RunEvaluations.py:79-113;
  * src/tests/integrator/test_reproducibility.py:5-20 -- same seed => identical image, spi changes the image.
Two of the reference's expectations cannot be met by its sources as written (recorded in DESIGN.md section 4):
  * test_point_light expects 0.005100456 for "intensity": 1, which is the value for "power": 1
    (intensity is used undivided, src/runtime/light/PointLight.cpp:39-49) -> tested with intensity = 1/(4 pi);
  * test_env_light expects 1.0 and plane-d6 matches Mitsuba only with a one-sided lambertian eval; the reference's
    eval is |cos| (src/artic/bsdf/diffuse.art:3) and yields 2.0 -> the flag IGB200_TECHFLAG_ONE_SIDED_DIFFUSE.
"""
import numpy as np
import pytest

from render_common import eval_error, flat_scene, load_golden, room_scene


def _tea(v0, v1):
    s, M = 0, 0xFFFFFFFF
    for _ in range(4):
        s = (s + 0x9e3779b9) & M
        v0 = (v0 + ((((v1 << 4) & M) + 0xa341316c) ^ ((v1 + s) & M) ^ ((v1 >> 5) + 0xc8013ea4))) & M
        v1 = (v1 + ((((v0 << 4) & M) + 0xad90777d) ^ ((v0 + s) & M) ^ ((v0 >> 5) + 0x7e95761e))) & M
    return v1


def _seed(*vals):
    h = 0x811C9DC5
    for d in vals:
        d &= 0xFFFFFFFF
        for k in range(4):
            h = ((h * 16777619) & 0xFFFFFFFF) ^ ((d >> (8 * k)) & 0xFF)
    return h


def test_rng_known_answers(oracle):
    # src/artic/core/random.art:4-24,34-43,65-87 restated independently in pure Python integers
    for args in ((0, 0, 0, 0, 0, 0), (3, 7, 1, 640, 480, 42), (63, 1023, 0, 3839, 2159, -5)):
        seed, raw, f32 = oracle.kat_rng(*args, n=6)
        assert seed == _seed(*args)
        exp = [_tea(seed, c) for c in range(1, 7)]
        assert list(raw) == exp
        expf = (np.array([(x & 0x7FFFFF) | 0x3F800000 for x in exp], np.uint32).view(np.float32) - np.float32(1))
        assert np.array_equal(f32, expf) and np.all((f32 >= 0) & (f32 < 1))


@pytest.mark.parametrize("light,one_sided,expected,tol", [
    ("none", False, 0.0, 0.0),
    ("point", False, 0.005100456, 1e-3),            # test_point_light (power semantics, see module docstring)
    ("point_intensity", False, 0.005100456 * 4 * np.pi, 1e-3),
    ("env", True, 1.0, 3e-3),                       # test_env_light
    ("env", False, 2.0, 3e-3),                      # the reference's sources as written (back-side NEE leak)
])
def test_flat_scene_closed_forms(ig, oracle, light, one_sided, expected, tol):
    tables, sh = flat_scene(ig, light, 200, one_sided)
    film = oracle.OracleScene(tables).render(sh, 200, 200, spi=4, iterations=4, seed=0)
    value = film.mean() / 4
    assert value == pytest.approx(expected, rel=tol, abs=1e-9)


def test_spot_light_against_quadrature(ig, oracle):
    # light/spot.art:8-40: intensity * cos_l * cos_s / (pi d^2) inside the cone (no blend range, nothing cut off)
    tables, sh = flat_scene(ig, "spot", 200)
    value = oracle.OracleScene(tables).render(sh, 200, 200, spi=4, iterations=2).mean() / 2
    x = (np.arange(2000) + 0.5) / 2000
    X, Y = np.meshgrid(x, x)
    d2 = X * X + Y * Y + 4
    assert value == pytest.approx(float((4 / (d2 * d2)).mean() / np.pi), rel=1e-3)


@pytest.mark.parametrize("name,one_sided,spp,threshold", [
    ("point", False, 64, 1e-3),            # default threshold of RunEvaluations.py:158
    ("emissive-plane", False, 256, 1e-3),
    ("cbox-d1", False, 32, 5e-3),          # predef_eps, RunEvaluations.py:91-93
    ("plane-d6", True, 64, 1e-3),
    ("cbox-d6", True, 512, 8e-3),          # 5e-3 is met at 1024 spp (GPU test); CPU budget here is 512 spp
    ("plane-d1", False, 64, 1e-3),
    ("flipped-prim-glass", False, 64, 1e-3),      # mirrored entity transform (negative determinant) around a dielectric
    ("emissive-plane-scale", False, 256, 1e-3),   # area light on a scaled entity ({"scale": 0.5} transform form)
    ("room", False, 512, 1e-3),                   # predef_eps "room", RunEvaluations.py:97
    ("sphere-light-ico", False, 1024, 2e-3),      # icosphere mesh as area light; predef_eps "sphere-light-ico"
    ("sphere-light-pure", False, 256, 3e-3),      # make_sphere_area_emitter; predef_eps "sphere-light-pure"
    ("two-planes-mirror", False, 256, 1e-3),      # sphere emitter + mirror against the Radiance reference
    ("sky-uniform", False, 16, 1e-3),             # make_cie_sky_light through the fishlens camera vs gensky -u
    ("sky-cloudy", False, 16, 1e-3),              # ... -c
    ("sky-clear", False, 16, 1e-3),               # make_cie_sunny_light, sun position from date / time / place vs gensky +s
    ("sky-intermediate", False, 16, 1e-3),        # ... +i
    ("sky-perez1", False, 16, 1e-3),                # make_perez_light, coefficients from irradiances (PerezModel.cpp) vs gendaylit
])
def test_golden_images(ig, oracle, name, one_sided, spp, threshold):
    tables, sh, w, h, ref = load_golden(ig, name)
    if one_sided:
        sh.technique.flags = 1
    spi = 8
    film = oracle.OracleScene(tables).render(sh, w, h, spi=spi, iterations=spp // spi, seed=1) / (spp // spi)
    assert np.all(np.isfinite(film))
    assert eval_error(film, ref) < threshold


def test_reproducibility(ig, oracle):
    # src/tests/integrator/test_reproducibility.py:5-20
    tables, sh = room_scene(ig, 48)
    osc = oracle.OracleScene(tables)
    a = osc.render(sh, 48, 48, spi=2, iterations=2, seed=3)
    b = osc.render(sh, 48, 48, spi=2, iterations=2, seed=3, threads=1)
    assert np.array_equal(a, b)
    c = osc.render(sh, 48, 48, spi=4, iterations=1, seed=3)
    assert not np.array_equal(a, c)
    d = osc.render(sh, 48, 48, spi=2, iterations=2, seed=4)
    assert not np.array_equal(a, d)


def test_igtrace_list_mode_debug_ids(ig, oracle):
    # igtrace contract (src/frontend/trace/main.cpp): --technique debug, __debug_mode 15 / 17 / 13
    tables = ig.workloads.c3_scene(grid=2, stacks=8, slices=16)
    rays = ig.workloads.c3_rays(tables, n=4096)
    osc = oracle.OracleScene(tables)
    ent, prim, t, _, _ = osc.trace(rays)
    sh = ig.Shading()
    sh.add_material(sh.add_diffuse())
    for mode, col in ((15, prim), (17, ent), (13, t)):
        sh.set_debug(mode)
        film = osc.render(sh, 0, 0, spi=1, rays=rays)[0, :, 0]
        hit = prim >= 0
        assert np.array_equal(film[hit], col[hit].astype(np.float32)) and np.all(film[~hit] == 0)
