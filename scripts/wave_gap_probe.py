"""Are there gaps between the kernels of a wavefront round?  The same batched render once through the captured round graph (device
time of the whole call) and once with per-kernel events (--stats path: sum of the kernel times).  usage: FIXTURE=diamond python scripts/wave_gap_probe.py"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("ignis-masterthesis_b200")
fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
B = int(os.environ.get("BATCH", "0"))
for name, res, spi, batch in (("diamond", 1920, 4, B or 32), ("many-point-lights", 3840, 1, B or 32), ("office", 256, 64, B or 64)):
    tables, sh, w0, h0, _ = fx.load_fixture(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    resy = res * h0 // w0
    out = {}
    for stats in (False, True):
        with pkg.Device(0, acquire_stats=stats) as dev:
            dev.assign_scene(tables); dev.assign_shading(sh)
            dev.render(res, resy, spi=spi, iteration=1000, iterations=batch)
            if stats:
                dev.section_stats(reset=True)
            dev.render(res, resy, spi=spi, iteration=0, iterations=batch)
            st = dev.render_stats()
            if stats:
                sec = dev.section_stats()
                out["kernels"] = {k: round(v["ms"], 2) for k, v in sec.items() if isinstance(v, dict) and "ms" in v and v["ms"] > 0}
                out["sum_ms"] = sum(out["kernels"].values())
            else:
                out["graph_ms"] = st["device_ms"]; out["rounds"] = st["rounds"]; out["launches"] = st["kernel_launches"]
    print(name, f"graph {out['graph_ms']:.2f} ms, sum of kernels {out['sum_ms']:.2f} ms, rounds {out['rounds']}, launches {out['launches']}: "
          f"{(out['graph_ms'] - out['sum_ms']) / out['launches'] * 1e3:.1f} us per launch unaccounted", out["kernels"], flush=True)
