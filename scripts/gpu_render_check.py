"""GPU-side check of the wavefront path tracer against the oracle on a few synthetic scenes (dev tool)."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
pkg = importlib.import_module("ignis-masterthesis_b200")
import pyoracle

def flat_scene(light, res):
    b = pkg.SceneBuilder()
    v = np.array([[-1,-1,0],[1,-1,0],[-1,1,0],[1,1,0]], np.float32)
    n = -np.tile(np.array([0,0,1],np.float32),(4,1))
    t = np.array([[0,0],[1,0],[0,1],[1,1]], np.float32)
    f = np.array([[0,1,3],[0,3,2]], np.uint32)[:, [0,2,1]]
    b.add_entity(b.add_trimesh(v, f, n, t), material="ground")
    tables = b.finish()
    sh = pkg.Shading()
    sh.set_perspective_from_transform(np.array([[1,0,0,0],[0,1,0,0],[0,0,1,-1],[0,0,0,1]],float), 90, res, res, 0.01, 100)
    sh.set_path(max_depth=2)
    sh.add_material(sh.add_diffuse((1,1,1)))
    if light == "point": sh.add_point_light((0,0,-2), np.float32(1/(4*np.pi)))
    if light == "env": sh.add_env_light((1,1,1))
    return tables, sh

def box_scene(res):
    """closed diffuse room with an area light (plane emitter), a diffuse box, a glass sphere mesh and a mirror."""
    W = pkg.workloads
    b = pkg.SceneBuilder()
    def quad(p0, ex, ey):
        p0, ex, ey = [np.asarray(a, np.float32) for a in (p0, ex, ey)]
        v = np.stack([p0, p0+ex, p0+ey, p0+ex+ey]); n = np.cross(ex, ey); n = n/np.linalg.norm(n)
        return v, np.array([[0,1,3],[0,3,2]], np.uint32), np.tile(n,(4,1))
    walls = [((-2,-2,-2),(4,0,0),(0,0,4)), ((-2,2,-2),(0,0,4),(4,0,0)), ((-2,-2,-2),(0,0,4),(0,4,0)),
             ((2,-2,-2),(0,4,0),(0,0,4)), ((-2,-2,2),(4,0,0),(0,4,0))]
    ents = []
    for k, (p, ex, ey) in enumerate(walls):
        v, f, n = quad(p, ex, ey)
        ents.append(b.add_entity(b.add_trimesh(v, f, n), material="red" if k == 2 else ("green" if k == 3 else "white")))
    v, f, n = quad((-0.5,1.99,-0.5),(1,0,0),(0,0,1))  # normal (0,-1,0): faces down
    light_ent = b.add_entity(b.add_trimesh(v, f, n), material="light")
    sv, sf, sn = W.uv_sphere(16, 32, 0.6)
    b.add_entity(b.add_trimesh(sv, sf, sn), np.array([[1,0,0,-0.9],[0,1,0,-1.4],[0,0,1,0.3],[0,0,0,1]],float), material="glass")
    b.add_entity(b.add_trimesh(sv, sf, sn), np.array([[1,0,0,0.9],[0,1,0,-1.4],[0,0,1,-0.4],[0,0,0,1]],float), material="mirror")
    b.add_entity(b.add_sphere((0,0,0),0.4), np.array([[1,0,0,0.0],[0,1,0,-1.6],[0,0,1,1.2],[0,0,0,1]],float), material="white")
    tables = b.finish()
    sh = pkg.Shading()
    sh.set_perspective((0,0,-6.5),(0,0,1),(0,1,0), 40, res, res, 0.1, 100)
    sh.set_path(max_depth=8)
    bs = {"white": sh.add_diffuse((0.8,0.8,0.8)), "red": sh.add_diffuse((0.8,0.1,0.1)), "green": sh.add_diffuse((0.1,0.8,0.1)),
          "light": sh.add_diffuse((0,0,0)), "glass": sh.add_dielectric(1.5), "mirror": sh.add_mirror((0.9,0.9,0.9))}
    o, x, y = v[0], v[1]-v[0], v[2]-v[0]
    lid = sh.add_area_light(light_ent, (20,20,20), plane=(o, x, y))
    for name in b._materials:  # material ids in first-use order
        sh.add_material(bs[name], lid if name == "light" else -1)
    return tables, sh

def relmse(a, b):
    m = b != 0
    e = np.square(a-b); e[m] /= np.square(b[m]); e = e.mean(axis=2)
    return float(np.clip(e, 0, np.percentile(e, 99)).mean())

def compare(name, tables, sh, res, spi, iters):
    osc = pyoracle.OracleScene(tables)
    t0 = time.time(); ref = osc.render(sh, res, res, spi=spi, iterations=iters, seed=7); t_cpu = time.time()-t0
    with pkg.Device(0) as dev:
        dev.assign_scene(tables); dev.assign_shading(sh)
        t0 = time.time()
        for it in range(iters): dev.render(res, res, spi=spi, iteration=it, seed=7)
        film, n = dev.framebuffer(); t_gpu = time.time()-t0
        st = dev.render_stats()
    d = np.abs(film-ref); rel = d.sum()/max(ref.sum(),1e-30)
    bad = (d > 1e-3*np.maximum(np.abs(ref),1e-3)).mean()
    print(f"{name}: mean gpu {film.mean()/iters:.6f} cpu {ref.mean()/iters:.6f} L1rel {rel:.3e} relmse {relmse(film/iters, ref/iters):.3e} "
          f"px>1e-3 {bad:.4f} stats gpu {st['camera_rays']},{st['bounce_rays']},{st['shadow_rays']} (last iter) cpu {osc.ray_stats/iters} "
          f"rounds {st['rounds']} gpu {t_gpu:.2f}s ({st['device_ms']:.2f} ms/iter) cpu {t_cpu:.2f}s", flush=True)

if __name__ == "__main__":
    res = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    for light in ("none", "point", "env"):
        tables, sh = flat_scene(light, res)
        compare("flat-"+light, tables, sh, res, 4, 2)
    tables, sh = box_scene(res)
    compare("box", tables, sh, res, 4, 4)
    sh.technique.flags = 1
    compare("box-onesided", tables, sh, res, 4, 4)
    sh.set_debug(15)
    compare("box-debug15", tables, sh, res, 1, 1)
