"""Throughput of the wavefront path tracer on the synthetic box scene (dev tool)."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
pkg = importlib.import_module("ignis-masterthesis_b200")
import gpu_render_check as G
res = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
spi = int(sys.argv[2]) if len(sys.argv) > 2 else 8
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 10
depth = int(sys.argv[4]) if len(sys.argv) > 4 else 8
fixture = os.environ.get("FIXTURE")
if fixture:
    fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
    tables, sh, w0, h0, _ = fx.load_fixture(os.path.join(ROOT, "tests", "golden", fixture + ".npz"))
    resy = res * h0 // w0
else:
    tables, sh = G.box_scene(res)
    sh.set_path(max_depth=depth)
    resy = res
with pkg.Device(0) as dev:
    dev.assign_scene(tables); dev.assign_shading(sh)
    batch = int(os.environ.get("BATCH", "1"))
    for it in range(int(os.environ.get('WARM','2'))): dev.render(res, resy, spi=spi, iteration=1_000_000 + it * batch, iterations=batch)  # same shape as the timed calls
    ms, rays = [], 0
    t0 = time.time()
    pipe = os.environ.get("PIPE", "0") == "1"  # igb200_render_pipelined, one iteration per call; wall includes the final flush
    for it in range(iters):
        dev.render(res, resy, spi=spi, iteration=2+it*batch, iterations=batch, pipelined=pipe)
        st = dev.render_stats(); ms.append(st["device_ms"]/batch); rays = (st["camera_rays"]+st["bounce_rays"]+st["shadow_rays"])/batch
    if pipe: dev.flush()
    wall = time.time()-t0
    ms = np.array(ms)
    print(f"res {res} spi {spi} depth {depth} cap {os.environ.get('IGB200_WAVE_CAPACITY','1M')} sort {os.environ.get('IGB200_WAVE_SORT','1')}: "
          f"{res*resy*spi/np.median(ms)/1e3:.1f} Msamples/s (median {np.median(ms):.2f} ms/iter, wall {wall/iters*1e3:.2f} ms/iter), "
          f"{rays/np.median(ms)/1e3:.1f} Mrays/s, rounds {st['rounds']} launches {st['kernel_launches']}"
          + (f", pipelined: {res*resy*spi*iters/wall/1e6:.1f} Msamples/s by wall clock incl. flush" if pipe else ""), flush=True)
