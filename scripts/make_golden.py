"""Generates tests/golden/*.npz from the reference tree (run in the build container only; /root/reference is
read-only and absent on the GPU box).  Each fixture = the table blobs `sceneio.load_scene` builds from the
reference's scene file + the reference's own golden image where it ships one."""
import importlib, os, sys
os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
import cv2
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("ignis-masterthesis_b200")
sceneio = importlib.import_module("ignis-masterthesis_b200.sceneio")
fixtures = importlib.import_module("ignis-masterthesis_b200.fixtures")
REF = "/root/reference/scenes"
OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)

def exr(name):
    img = cv2.imread(f"{REF}/evaluation/references/{name}", cv2.IMREAD_UNCHANGED)
    return np.ascontiguousarray(img[:, :, :3][:, :, ::-1]).astype(np.float32)

JOBS = [  # (fixture, scene file, reference image, width, height)
    ("cbox-d6", "evaluation/cbox-d6.json", "ref-cbox-d6-4096.exr", None, None),
    ("cbox-d1", "evaluation/cbox-d1.json", "ref-cbox-d1-4096.exr", None, None),
    ("point", "evaluation/point.json", "ref-point-4096.exr", None, None),
    ("emissive-plane", "evaluation/emissive-plane.json", "ref-emissive-plane-4096.exr", None, None),
    ("plane-d6", "evaluation/plane-d6.json", "ref-plane-d6-4096.exr", None, None),
    ("diamond", "diamond_scene.json", None, 1920, 1080),
    ("primitives", "primitives.json", None, 1024, 1024),  # BASELINE.json configs[1]
    ("multilight-uniform", "evaluation/multilight-uniform.json", "ref-multilight-4096.exr", None, None),
    ("multilight-hierarchy", "evaluation/multilight-hierarchy.json", "ref-multilight-4096.exr", None, None),
    # more of scripts/RunEvaluations.py's scene list that needs nothing beyond the supported subset
    ("plane-d1", "evaluation/plane-d1.json", "ref-plane-d1-4096.exr", None, None),
    ("room", "evaluation/room.json", "ref-room-4096.exr", None, None),
    ("flipped-prim-glass", "evaluation/flipped-prim-glass.json", "ref-flipped-prim-glass-4096.exr", None, None),
    ("emissive-plane-scale", "evaluation/emissive-plane-scale.json", "ref-emissive-plane-scale-4096.exr", None, None),
    ("sphere-light-ico", "evaluation/sphere-light-ico.json", "ref-sphere-light-4096.exr", None, None),  # mesh area light
    ("sphere-light-pure", "evaluation/sphere-light-pure.json", "ref-sphere-light-4096.exr", None, None),  # analytic sphere emitter
    ("two-planes-mirror", "evaluation/two-planes-mirror.json", "ref-two-planes-mirror-rad.exr", None, None),  # sphere light + mirror, Radiance reference
    # the scene of the reference's only published timings (scenes/radiance/office/benchmark.txt): fishlens camera, thin dielectric
    ("office", "radiance/office/office_scene.json", None, None, None),
    # analytic CIE skies through the fishlens camera, against the Radiance (gensky) references
    ("sky-uniform", "evaluation/sky-uniform.json", "ref-sky-uniform-rad.exr", None, None),
    ("sky-cloudy", "evaluation/sky-cloudy.json", "ref-sky-cloudy-rad.exr", None, None),
    ("sky-clear", "evaluation/sky-clear.json", "ref-sky-clear-rad.exr", None, None),
    ("sky-intermediate", "evaluation/sky-intermediate.json", "ref-sky-intermediate-rad.exr", None, None),
    ("sky-perez1", "evaluation/sky-perez1.json", "ref-sky-perez1-rad.exr", None, None),  # Perez all-weather sky vs gendaylit
    ("multilight-simple", "evaluation/multilight-simple.json", "ref-multilight-4096.exr", None, None),  # CDF ("simple") selector
    ("many-point-lights", "many_point_lights.json", None, 3840, 2160),  # BASELINE.json configs[4]
    # textured environment + 2-D CDF.  The shipped Mitsuba image (ref-env-4096.exr) is kept for the record but is NOT a
    # pin: it is 2.16x below the analytic irradiance of the one lit texel (note 5).
    ("env", "evaluation/env.json", "ref-env-4096.exr", None, None),
]
only = set(sys.argv[1:])
for name, scene, ref, w, h in JOBS:
    if only and name not in only:
        continue
    tables, sh, w, h, info = sceneio.load_scene(f"{REF}/{scene}", width=w, height=h)
    fixtures.save_fixture(os.path.join(OUT, name + ".npz"), tables, sh, w, h, exr(ref) if ref else None,
                          meta=f"{scene}; reference image {ref}")
    print(name, w, h, "entities", tables.entity_count, "materials", tables.material_count, os.path.getsize(os.path.join(OUT, name + ".npz")))
