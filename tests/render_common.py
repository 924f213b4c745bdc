"""Shared helpers of the render tests: scenes, fixtures and the reference's image error metric."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def eval_error(img, ref):
    """error_image() of the reference, scripts/RunEvaluations.py:79-86, literally: per ELEMENT (pixel and channel) the
    relative squared error, absolute where ref == 0; clipped at the 99th percentile over all elements; mean."""
    img, ref = np.asarray(img, np.float64), np.asarray(ref, np.float64)
    mask = ref != 0
    err = np.zeros_like(ref)
    err[mask] = np.square((img[mask] - ref[mask]) / ref[mask])
    err[~mask] = np.square(img[~mask])
    return float(np.average(np.clip(err, 0, np.percentile(err, 99))))


def flat_scene(ig, light, res, one_sided=False, depth=2):
    """create_flat_scene() of src/tests/integrator/common/__init__.py:34-62: a 2x2 rectangle (flip_normals) at z = 0
    seen by a 90-degree camera at z = -1, white diffuse, path depth 2."""
    b = ig.SceneBuilder()
    v = np.array([[-1, -1, 0], [1, -1, 0], [-1, 1, 0], [1, 1, 0]], np.float32)
    n = -np.tile(np.array([0, 0, 1], np.float32), (4, 1))
    t = np.array([[0, 0], [1, 0], [0, 1], [1, 1]], np.float32)
    f = np.array([[0, 1, 3], [0, 3, 2]], np.uint32)[:, [0, 2, 1]]
    b.add_entity(b.add_trimesh(v, f, n, t), material="ground")
    tables = b.finish()
    sh = ig.Shading()
    sh.set_perspective_from_transform(np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, -1], [0, 0, 0, 1]], float), 90, res, res, 0.01, 100)
    sh.set_path(max_depth=depth, one_sided_diffuse=one_sided)
    sh.add_material(sh.add_diffuse((1, 1, 1)))
    if light == "point":  # "power": 1 -> intensity 1/(4 pi) (src/runtime/light/PointLight.cpp:16-49)
        sh.add_point_light((0, 0, -2), np.float32(1 / (4 * np.pi)))
    elif light == "point_intensity":
        sh.add_point_light((0, 0, -2), 1.0)
    elif light == "spot":
        sh.add_spot_light((0, 0, -2), (0, 0, 1), 45, 45, 1.0)
    elif light == "env":
        sh.add_env_light((1, 1, 1))
    elif light == "directional":
        sh.add_directional_light((0, 0, 1), 1.0)
    return tables, sh


def room_scene(ig, res, depth=8):
    """Closed room: diffuse walls, downward plane area light, glass + mirror sphere meshes, analytic sphere,
    conductor box top -- exercises every BSDF / light-hit path of the supported subset."""
    W = ig.workloads
    b = ig.SceneBuilder()

    def quad(p0, ex, ey):
        p0, ex, ey = [np.asarray(a, np.float32) for a in (p0, ex, ey)]
        v = np.stack([p0, p0 + ex, p0 + ey, p0 + ex + ey])
        n = np.cross(ex, ey)
        return v, np.array([[0, 1, 3], [0, 3, 2]], np.uint32), np.tile(n / np.linalg.norm(n), (4, 1))

    walls = [((-2, -2, -2), (4, 0, 0), (0, 0, 4), "white"), ((-2, 2, -2), (0, 0, 4), (4, 0, 0), "white"),
             ((-2, -2, -2), (0, 0, 4), (0, 4, 0), "red"), ((2, -2, -2), (0, 4, 0), (0, 0, 4), "green"),
             ((-2, -2, 2), (4, 0, 0), (0, 4, 0), "white"), ((-1.2, -1.2, -1.0), (0.8, 0, 0), (0, 0, 0.8), "gold")]
    for p, ex, ey, m in walls:
        v, f, n = quad(p, ex, ey)
        b.add_entity(b.add_trimesh(v, f, n), material=m)
    lv, lf, ln = quad((-0.5, 1.99, -0.5), (1, 0, 0), (0, 0, 1))
    light_ent = b.add_entity(b.add_trimesh(lv, lf, ln), material="light")
    sv, sf, sn = W.uv_sphere(16, 32, 0.6)
    shape = b.add_trimesh(sv, sf, sn)
    b.add_entity(shape, np.array([[1, 0, 0, -0.9], [0, 1, 0, -1.4], [0, 0, 1, 0.3], [0, 0, 0, 1]], float), material="glass")
    b.add_entity(shape, np.array([[1, 0, 0, 0.9], [0, 1, 0, -1.4], [0, 0, 1, -0.4], [0, 0, 0, 1]], float), material="mirror")
    b.add_entity(b.add_sphere((0, 0, 0), 0.4), np.array([[1, 0, 0, 0.0], [0, 1, 0, -1.6], [0, 0, 1, 1.2], [0, 0, 0, 1]], float),
                 material="white")
    tables = b.finish()
    sh = ig.Shading()
    sh.set_perspective((0, 0, -6.5), (0, 0, 1), (0, 1, 0), 40, res, res, 0.1, 100)
    sh.set_path(max_depth=depth)
    bs = {"white": sh.add_diffuse((0.8, 0.8, 0.8)), "red": sh.add_diffuse((0.8, 0.1, 0.1)), "green": sh.add_diffuse((0.1, 0.8, 0.1)),
          "light": sh.add_diffuse((0, 0, 0)), "glass": sh.add_dielectric(1.5), "mirror": sh.add_mirror((0.9, 0.9, 0.9)),
          "gold": sh.add_conductor((0.143, 0.375, 1.442), (3.983, 2.386, 1.603))}
    lid = sh.add_area_light(light_ent, (20, 20, 20), plane=(lv[0], lv[1] - lv[0], lv[2] - lv[0]))
    sh.add_point_light((1.5, 1.0, -1.5), (0.5, 0.5, 1.0))
    for name in b._materials:  # material ids are handed out in first-use order
        sh.add_material(bs[name], lid if name == "light" else -1)
    return tables, sh


def features_scene(ig, res, depth=6):
    """Open stage exercising the ABI v3 shading paths: checkerboard and image textures (nearest / bilinear /
    bicubic, repeat / mirror / clamp), a rough conductor (GGX VNDF), a bump-mapped rough conductor, a smooth
    conductor, a textured environment with a 2-D CDF and the light hierarchy over point + plane lights."""
    import importlib
    io = importlib.import_module("ignis-masterthesis_b200.sceneio")
    S = ig.shading
    W = ig.workloads
    b = ig.SceneBuilder()
    uv4 = np.array([[0, 0], [1, 0], [0, 1], [1, 1]], np.float32)

    def quad(p0, ex, ey):
        p0, ex, ey = [np.asarray(a, np.float32) for a in (p0, ex, ey)]
        v = np.stack([p0, p0 + ex, p0 + ey, p0 + ex + ey])
        n = np.cross(ex, ey)
        return v, np.array([[0, 1, 3], [0, 3, 2]], np.uint32), np.tile(n / np.linalg.norm(n), (4, 1))

    for p, ex, ey, m in [((-3, -2, -3), (6, 0, 0), (0, 0, 6), "floor"), ((-3, -2, 3), (6, 0, 0), (0, 4, 0), "back"),
                         ((-3, -2, -3), (0, 0, 6), (0, 4, 0), "side")]:
        v, f, n = quad(p, ex, ey)
        b.add_entity(b.add_trimesh(v, f, n, uv4), material=m)
    lv, lf, ln = quad((-0.5, 2.5, -0.5), (1, 0, 0), (0, 0, 1))
    light_ent = b.add_entity(b.add_trimesh(lv, lf, ln, uv4), material="light")
    sv, sf, sn = W.uv_sphere(24, 48, 0.7)
    suv = np.stack([np.arctan2(sn[:, 1], sn[:, 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(sn[:, 2], -1, 1)) / np.pi], axis=1).astype(np.float32)
    shape = b.add_trimesh(sv, sf, sn, suv)
    T = lambda x, y, z: np.array([[1, 0, 0, x], [0, 1, 0, y], [0, 0, 1, z], [0, 0, 0, 1]], float)
    b.add_entity(shape, T(-1.6, -1.3, 0.5), material="rough")
    b.add_entity(shape, T(0.0, -1.3, 0.0), material="bumped")
    b.add_entity(shape, T(1.6, -1.3, 0.5), material="smooth")
    tables = b.finish()

    sh = ig.Shading()
    sh.set_perspective((0, 0.3, -7.0), (0, -0.05, 1), (0, 1, 0), 45, res, res, 0.1, 100)
    sh.set_path(max_depth=depth)
    rng = np.random.default_rng(7)
    img8 = rng.integers(0, 256, (32, 64, 4), dtype=np.uint32)
    packed = (img8[:, :, 0] | (img8[:, :, 1] << 8) | (img8[:, :, 2] << 16) | (img8[:, :, 3] << 24)).astype(np.uint32)
    yy, xx = np.mgrid[0:48, 0:48]
    bump8 = (127 + 120 * np.sin(xx * 0.7) * np.cos(yy * 0.5)).astype(np.uint32)
    bump = (bump8 | (bump8 << 8) | (bump8 << 16) | (np.uint32(255) << 24)).astype(np.uint32)
    t_check = sh.add_checkerboard((6, 6), (0.2, 0.2, 0.25), (0.9, 0.9, 0.8))
    t_img = sh.add_image_texture(packed, S.FILTER_BICUBIC, (S.WRAP_REPEAT, S.WRAP_MIRROR), np.array([[2, 0, 0.25], [0, 2, 0.1], [0, 0, 1]]))
    t_side = sh.add_image_texture(packed, S.FILTER_NEAREST, (S.WRAP_CLAMP, S.WRAP_REPEAT), np.array([[1.5, 0, -0.2], [0, 1.5, 0], [0, 0, 1]]), opaque=True)
    t_bump = sh.add_image_texture(bump, S.FILTER_BILINEAR, (S.WRAP_REPEAT, S.WRAP_REPEAT), np.array([[4, 0, 0], [0, 4, 0], [0, 0, 1]]))
    env = np.zeros((16, 32, 4), np.float32)
    env[:, :, :3] = 0.15
    env[10:13, 5:9, :3] = (6.0, 5.0, 3.0)
    env[:, :, 3] = 1
    t_env = sh.add_image_texture(env, S.FILTER_BILINEAR, (S.WRAP_REPEAT, S.WRAP_REPEAT))
    gold = ((0.143, 0.375, 1.442), (3.983, 2.386, 1.603))
    inner = sh.add_conductor(*gold, roughness=0.16)
    bs = {"floor": sh.add_diffuse(texture=t_check), "back": sh.add_diffuse(texture=t_img), "side": sh.add_diffuse(texture=t_side),
          "light": sh.add_diffuse((0, 0, 0)), "rough": sh.add_conductor(*gold, alpha=(0.3, 0.1)),
          "bumped": sh.add_bumpmap(inner, t_bump, 0.2), "smooth": sh.add_conductor(*gold)}
    lid = sh.add_area_light(light_ent, (15, 15, 15), plane=(lv[0], lv[1] - lv[0], lv[2] - lv[0]))
    hier = [((lv[0] + lv[3]) / 2, ln[0], 15.0 * np.pi)]
    for pos, col in [((2.5, 1.0, -2.0), (3, 1, 1)), ((-2.5, 0.5, -1.0), (1, 3, 1)), ((0.0, 1.5, 2.0), (1, 1, 4)), ((2.0, -1.0, 2.5), (2, 2, 2))]:
        sh.add_point_light(pos, col)
        hier.append((pos, None, float(np.mean(col)) * 4 * np.pi))
    marginal, cond = io.compute_cdf2d(env, True, False)
    sh.add_env_textured(t_env, (1.5, 1.5, 1.5), sh.add_cdf2d(marginal, cond))
    sh.set_hierarchy(*io.build_light_hierarchy(hier))
    for name in b._materials:
        sh.add_material(bs[name], lid if name == "light" else -1)
    return tables, sh


def load_golden(ig, name):
    import importlib
    fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
    return fx.load_fixture(os.path.join(GOLDEN, name + ".npz"))
