"""dev tool: count non-finite pixels of the C5 scene on the GPU at growing sample counts, and per max depth."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("ignis-masterthesis_b200")
fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
tables, sh, _, _, _ = fx.load_fixture(os.path.join(ROOT, "tests", "golden", "many-point-lights.npz"))
w, h = 3840, 2160
for md, its in ((64, 64), (64, 1024)):
    sh.technique.max_depth = md
    with pkg.Device(0) as dev:
        dev.assign_scene(tables); dev.assign_shading(sh); dev.resize(w, h)
        dev.render(w, h, spi=1, iteration=0, seed=0, iterations=its)
        film, n = dev.framebuffer()
        bad = ~np.isfinite(film).all(axis=2)
        ys, xs = np.nonzero(bad)
        print("depth", md, "its", its, "bad px", int(bad.sum()), "nan", int(np.isnan(film).sum()), "inf", int(np.isinf(film).sum()),
              list(zip(xs[:5], ys[:5])), flush=True)
