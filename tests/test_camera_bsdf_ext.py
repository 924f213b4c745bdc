"""Fishlens camera (camera/fishlens.art) and thin dielectric (bsdf/dielectric.art:27-47) in the oracle: the two pieces
scenes/radiance/office/office_scene.json -- the scene of the reference's only published timings -- needs beyond the
five BASELINE configs; plus the loader features that file uses (JSON comments, repeated transform keys, qrotate)."""
import json

import numpy as np
import pytest


def _fish_dirs(nx, ny, view, xasp=1.0, yasp=1.0):
    fx, fy = nx * xasp, ny * yasp
    r = np.sqrt(fx * fx + fy * fy)
    theta = r * np.pi / 2
    with np.errstate(invalid="ignore", divide="ignore"):
        sp, cp = np.where(r < 1e-7, 0, fy / r), np.where(r < 1e-7, 0, fx / r)
    d = np.stack([np.sin(theta) * cp, np.sin(theta) * sp, np.cos(theta)], axis=-1) @ view.T
    return d / np.linalg.norm(d, axis=-1, keepdims=True)


def _unit_sphere_around_eye(ig, eye):
    b = ig.SceneBuilder()
    b.add_entity(b.add_sphere(eye, 1.0), material="m")
    return b.finish()


def test_fishlens_ray_directions(ig, oracle):
    """A unit sphere centred on the eye: debug mode 10 (world hit point) minus the eye IS the camera ray direction."""
    eye = np.array([0.5, -1.0, 2.0])
    m = np.eye(4)
    m[:3, 3] = eye
    m[:3, :3] = np.array([[0, 0, 1], [1, 0, 0], [0, 1, 0]], float).T  # right-handed frame: dir = col 2, up = col 1
    res, spp = 32, 256
    sh = ig.Shading()
    sh.set_fishlens_from_transform(m, res, res, 0.0, 100.0)
    sh.add_material(sh.add_diffuse())
    sh.set_debug(10)
    assert sh.camera.type == ig.shading.CAMERA_FISHLENS and tuple(sh.camera.scale) == (1.0, 1.0)
    film = oracle.OracleScene(_unit_sphere_around_eye(ig, eye)).render(sh, res, res, spi=spp, iterations=1, seed=3)  # one iteration = mean over spi
    got = film - eye.astype(np.float32)
    up, cdir = m[:3, 1], m[:3, 2]
    right = np.cross(cdir, up)
    view = np.stack([right / np.linalg.norm(right), up, cdir], axis=1)
    c = (np.arange(res) + 0.5) / res
    nx, ny = np.meshgrid(2 * c - 1, 1 - 2 * c)
    want = _fish_dirs(nx, ny, view)
    # the pixel sampler jitters inside the pixel: the mean of 256 unit vectors sits within ~half a pixel of the centre direction
    assert np.abs(got - want).max() < 0.06 and np.abs(got - want).mean() < 5e-3
    centre = got[res // 2 - 1:res // 2 + 1, res // 2 - 1:res // 2 + 1].mean(axis=(0, 1))
    assert np.allclose(centre / np.linalg.norm(centre), cdir, atol=2e-2)
    assert np.allclose(got[res // 2, -1] / np.linalg.norm(got[res // 2, -1]), view[:, 0], atol=0.08)  # nx = 1: 90 degrees off axis


def test_fishlens_aspect_modes(ig):
    sh = ig.Shading()
    sh.set_fishlens_from_transform(np.eye(4), 200, 100)                 # circular: the circle fits the short side
    assert tuple(sh.camera.scale) == (2.0, 1.0)
    sh.set_fishlens_from_transform(np.eye(4), 100, 200)
    assert tuple(sh.camera.scale) == (1.0, 0.5)
    sh.set_fishlens_from_transform(np.eye(4), 200, 100, mode="cropped")
    assert tuple(sh.camera.scale) == (1.0, 0.5)
    sh.set_fishlens_from_transform(np.eye(4), 200, 100, mode="full")    # the circle passes through the corners
    f = np.float32(np.sqrt(5.0) * 100) / np.float32(100)
    assert np.allclose(tuple(sh.camera.scale), (f / 2, f))


def test_thin_dielectric_closed_form(ig, oracle):
    """A thin pane in front of the camera under a unit environment is a white furnace (ks = kt = 1): film == 1.
    With kt = 0.5 and a black reflection the film is 0.5 * (1 - F) with F = f + (1 - f) f / (f + 1) at the pixel's angle."""
    b = ig.SceneBuilder()
    v = np.array([[-50, -50, 0], [50, -50, 0], [-50, 50, 0], [50, 50, 0]], np.float32)
    b.add_entity(b.add_trimesh(v, np.array([[0, 1, 3], [0, 3, 2]], np.uint32), np.tile(np.array([0, 0, -1], np.float32), (4, 1))), material="pane")
    tables = b.finish()
    cam = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, -1], [0, 0, 0, 1]], float)
    for kt, ks in ((1.0, 1.0), (0.5, 0.0)):
        sh = ig.Shading()
        sh.set_perspective_from_transform(cam, 40, 16, 16, 0.01, 100)
        sh.set_path(max_depth=8)
        sh.add_material(sh.add_dielectric(1.52, 1.0, (ks,) * 3, (kt,) * 3, thin=True))
        sh.add_env_light((1, 1, 1))
        assert sh.bsdfs[0].type == ig.shading.BSDF_THIN_DIELECTRIC
        film = oracle.OracleScene(tables).render(sh, 16, 16, spi=256, iterations=1, seed=9)
        if ks == 1.0:
            assert np.allclose(film, 1.0, atol=1e-5)
        else:
            k = 1 / 1.52                                   # near the axis: cos ~ 1
            cos_t = np.sqrt(1 - (1 - 1.0) * k * k)
            rs, rp = (k - cos_t) / (k + cos_t), (1 - k * cos_t) / (1 + k * cos_t)
            f = 0.5 * (rs * rs + rp * rp)
            F = f + (1 - f) * f / (f + 1)
            assert film[7:9, 7:9].mean() == pytest.approx(0.5 * (1 - F), rel=2e-2)


def test_loader_reads_the_office_scene_dialect(ig, tmp_path):
    sio = ig.sceneio if hasattr(ig, "sceneio") else __import__("importlib").import_module("ignis-masterthesis_b200.sceneio")
    text = '''{ // comment
      "camera": {"type": "fishlens", "near_clip": 0.1, /* block */ "far_clip": 1000,
                 "transform": {"translate": [1, 2, 3], "qrotate": [0.7071067690849304, 0, 0.7071067690849304, 0],
                               "qrotate": [0.7071067690849304, -0.7071067690849304, 0, 0], "scale": [1, 1, -1]}},
      "film": {"size": [64, 32]},
      "bsdfs": [{"type": "dielectric", "name": "w", "int_ior": 1.52, "thin": true,},],
      "shapes": [{"type": "rectangle", "name": "r"}],
      "entities": [{"name": "e", "shape": "r", "bsdf": "w"}]
    }'''
    scene = sio.load_json(text, is_text=True)
    ops = scene["camera"]["transform"].pairs
    assert [k for k, _ in ops] == ["translate", "qrotate", "qrotate", "scale"]     # both quaternions survive
    tables, sh, w, h, _ = sio.load_scene(scene, str(tmp_path))
    assert (w, h) == (64, 32) and sh.camera.type == ig.shading.CAMERA_FISHLENS and sh.bsdfs[0].type == ig.shading.BSDF_THIN_DIELECTRIC
    # T * Ry(90) * Rx(-90) * S(1,1,-1): eye = translation, dir = -(R col 2), up = R col 1
    ry = np.array([[0, 0, 1], [0, 1, 0], [-1, 0, 0]], float)
    rx = np.array([[1, 0, 0], [0, 0, 1], [0, -1, 0]], float)
    r = ry @ rx
    assert np.allclose(sh.camera.eye, [1, 2, 3]) and np.allclose(sh.camera.dir, -r[:, 2], atol=1e-6) and np.allclose(sh.camera.up, r[:, 1], atol=1e-6)
    assert json.loads(sio._strip_json_comments('{"a": "// not a comment", "b": [1, 2,], }')) == {"a": "// not a comment", "b": [1, 2]}


@pytest.mark.parametrize("kind,variant,has_ground", [(0, True, True), (1, True, True), (1, False, False), (0, False, False), (2, False, True), (2, False, False)])
def test_sky_function_sampling_is_consistent_with_its_emission(ig, oracle, kind, variant, has_ground):
    """make_environment_light_function (light/env.art:24-108): next event estimation (sample_direct + pdf_direct, sphere or
    cosine-hemisphere sampling) and pure BSDF sampling (emission on miss) must estimate the same image."""
    from render_common import flat_scene
    tables, sh = flat_scene(ig, "none", 48, one_sided=True, depth=2)  # |cos| + sphere NEE lights the back side too (DESIGN 4.1 note 2)
    sun = np.array([0.3, 0.5, -0.8])
    tilt = np.array([[1, 0, 0], [0, 0, -1], [0, 1, 0]], float)                       # tilt the sky: its up axis is world -z
    if kind == 2:  # Perez: always sphere sampling, black below the horizon without ground
        sh.add_perez_sky(sun / np.linalg.norm(sun), (1.0, 0.9, 0.8), (0.3, 0.4, 0.2), (-1.0, -0.32, 6.0, -3.0, 0.3), has_ground, tilt)
        assert bool(sh.infinite[0].flags & ig.shading.SKYFLAG_HAS_GROUND) == has_ground and sh.infinite[0].tex == ig.shading.SKY_PEREZ
    else:
        sh.add_sky_function(kind, variant, has_ground, scale=(1, 1, 1), zenith=(1.0, 0.9, 0.8), ground=(0.3, 0.4, 0.2), zenith_brightness=1.7,
                            ground_brightness=0.2, c2=0.9, sun_dir=sun / np.linalg.norm(sun), transform=tilt)
    assert sh.infinite[0].type == ig.shading.LIGHT_ENV_FUNCTION
    osc = oracle.OracleScene(tables)
    means = []
    for nee in (True, False):
        sh.technique.enable_nee = int(nee)
        film = osc.render(sh, 48, 48, spi=64, iterations=4, seed=4) / 4
        assert np.all(np.isfinite(film)) and film.min() >= 0
        means.append(film.reshape(-1, 3).mean(axis=0))
    assert means[0].max() > 0.05
    assert np.allclose(means[0], means[1], rtol=2e-2)


def test_sun_light_closed_form_and_directional_limit(ig, oracle):
    """make_sun_light (light/sun.art:4-28) over the flat scene of the reference's light tests: a white plane lit head-on.
    Each sample carries color / (sun_area * pdf) = 2 (1 - c) / tan^2(A) and the plane's cosine is the sampled z, so the mean
    radiance is (1 / pi) * 2 (1 - c) / tan^2 A * E[z] with E[z] = (1 + c) / 2 for a uniform cone: (1 - c^2) / (pi tan^2 A)
    = cos^2(A) / pi.  As the cone closes this tends to the directional light's 1 / pi."""
    from render_common import flat_scene
    tables, sh = flat_scene(ig, "none", 24, one_sided=False, depth=2)
    osc = oracle.OracleScene(tables)
    for angle in (60.0, 11.4, 0.5):
        sh.infinite.clear()
        sh.add_sun_light((0, 0, 1), (1, 1, 1), angle)
        assert sh.infinite[0].type == ig.shading.LIGHT_SUN and sh.infinite[0].v[3] == pytest.approx(np.cos(np.radians(angle / 2)), rel=1e-6)
        film = osc.render(sh, 24, 24, spi=256, iterations=1, seed=2)
        want = np.cos(np.radians(angle / 2)) ** 2 / np.pi
        assert film.mean() == pytest.approx(want, rel=5e-3)
    sh.infinite.clear()
    sh.add_directional_light((0, 0, 1), (1, 1, 1))
    assert osc.render(sh, 24, 24, spi=16, iterations=1, seed=2).mean() == pytest.approx(1 / np.pi, rel=1e-4)
    sh.infinite.clear()
    sh.add_sun_light((0, 0, 1), (1, 1, 1), radius=0.1)   # sun_cos_angle_from_radius
    assert sh.infinite[0].v[3] == pytest.approx(1 / np.sqrt(1.01), rel=1e-6)
