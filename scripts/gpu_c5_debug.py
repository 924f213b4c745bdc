"""dev tool: is the GPU's ray-count expectation on the C5 scene the oracle's?  Independent seeds on both sides."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pyoracle
pkg = importlib.import_module("ignis-masterthesis_b200")
fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
w, h, spi, iters = 192, 108, 16, 4
tables, sh, _, _, _ = fx.load_fixture(os.path.join(ROOT, "tests", "golden", "many-point-lights.npz"))
g, o, gs, os_ = [], [], [], []
osc = pyoracle.OracleScene(tables)
with pkg.Device(0) as dev:
    dev.assign_scene(tables); dev.assign_shading(sh); dev.resize(w, h)
    for k in range(12):
        dev.clear_framebuffer()
        st = np.zeros(3, np.int64)
        for it in range(iters):
            dev.render(w, h, spi=spi, iteration=it, seed=100 + k)
            s = dev.render_stats(); st += (s["camera_rays"], s["bounce_rays"], s["shadow_rays"])
        film, _ = dev.framebuffer()
        g.append(st[1]); gs.append(float(film.sum()))
        ref = osc.render(sh, w, h, spi=spi, iterations=iters, seed=200 + k)
        o.append(int(osc.ray_stats[1])); os_.append(float(ref.sum()))
g, o, gs, os_ = map(np.asarray, (g, o, gs, os_))
print("bounce gpu mean %.0f +- %.0f   oracle mean %.0f +- %.0f" % (g.mean(), g.std() / np.sqrt(len(g)), o.mean(), o.std() / np.sqrt(len(o))))
print("sum    gpu mean %.1f +- %.1f   oracle mean %.1f +- %.1f" % (gs.mean(), gs.std() / np.sqrt(len(g)), os_.mean(), os_.std() / np.sqrt(len(o))))
