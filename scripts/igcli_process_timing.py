"""Whole-process timing of bin/igcli (the reference's published benchmark is `time igcli --spp 16384` on office_scene.json): writes the
committed fixtures as .igdb scene files and runs igcli per step (default: pipelined steps) and with --batch.  One GPU."""
import importlib, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("ignis-masterthesis_b200")
fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
igcli = os.path.join(ROOT, "ignis-masterthesis_b200", "bin", "igcli")
tmp = tempfile.mkdtemp()
for name, spp, spi in (("office", 16384, 64), ("diamond", 256, 4), ("many-point-lights", 64, 1)):
    tables, sh, w, h, _ = fx.load_fixture(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    scene = os.path.join(tmp, name + ".igdb")
    fx.save_igdb(scene, tables, sh, w, h)
    for extra in ([], ["--batch"]):
        t0 = time.time()
        r = subprocess.run([igcli, scene, "-o", os.path.join(tmp, "out.exr"), "--spp", str(spp), "--spi", str(spi)] + extra, capture_output=True, text=True)
        wall = time.time() - t0
        tail = [l.strip() for l in r.stdout.splitlines() if "Render>" in l or l.startswith("#")]
        print(f"{name} {w}x{h} spp {spp} spi {spi} {' '.join(extra) or 'per step'}: rc {r.returncode}, process {wall:.2f} s; " + "; ".join(tail), flush=True)
