"""The C++ frontends (ignis-masterthesis_b200/bin/igtrace, igcli) end to end on the GPU, against the same calls made
through the Python binding of the C ABI: the igtrace ray-list-in / contribution-out contract of
src/frontend/trace/main.cpp and igcli's iteration loop + image output (src/frontend/cli/main.cpp)."""
import os
import subprocess

import numpy as np
import pytest

from render_common import load_golden

pytestmark = pytest.mark.gpu


def _run(*cmd):
    r = subprocess.run(list(cmd), capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    return r


def _write_rays(path, rays):
    # the igtrace text format: `ox oy oz dx dy dz tmin tmax` per line; %.9g round-trips f32 exactly
    cols = np.concatenate([rays["org"], rays["dir"], rays["tmin"][:, None], rays["tmax"][:, None]], axis=1)
    np.savetxt(path, cols, fmt="%.9g")


def _debug_shading(ig):
    sh = ig.Shading()
    sh.add_material(sh.add_diffuse())
    return sh


def test_igtrace_c1_first_10000_rays_hit_contract(ig, host_bin, tmp_path):
    """BASELINE.json configs[0] through the CLI: single triangle, the first 10 000 rays of the 1 M-ray list as text."""
    tables = ig.workloads.c1_scene()
    rays = ig.workloads.c1_rays(n=10_000)
    scene, rfile, out = str(tmp_path / "c1.igdb"), str(tmp_path / "rays.txt"), str(tmp_path / "out.txt")
    ig.fixtures.save_igdb(scene, tables, _debug_shading(ig), rays.size, 1)
    _write_rays(rfile, rays)
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        ent, prim, t, u, v = dev.trace(rays)
    hit = prim >= 0
    assert 0.08 < hit.mean() < 0.17  # expected hit rate 12.5 %
    # debug mode 13 = hit distance (debugtracer.art:76-78): `t t t` per hit, `0 0 0` per miss
    _run(os.path.join(host_bin, "igtrace"), scene, "-i", rfile, "-o", out, "--technique", "debug", "--debug-mode", "13")
    got = np.loadtxt(out, dtype=np.float64).reshape(-1, 3)
    assert got.shape == (rays.size, 3)
    assert np.all(got[~hit] == 0) and np.all(got[hit, 0] > 0)  # hit / miss decisions identical
    np.testing.assert_allclose(got[hit, 0], t[hit], rtol=1e-6)   # std::scientific prints 7 significant digits
    first = open(out).readline()
    assert first.count("\t") == 2 and "e" in first               # `r\tg\tb` in scientific notation (trace/main.cpp:63-67)


def test_igtrace_ids_on_instanced_scene_and_spp_average(ig, host_bin, tmp_path):
    tables = ig.workloads.c3_scene(grid=2, stacks=16, slices=32)
    rays = ig.workloads.c3_rays(tables, n=20_000)
    scene, rfile, out = str(tmp_path / "c3.igdb"), str(tmp_path / "rays.txt"), str(tmp_path / "out.txt")
    ig.fixtures.save_igdb(scene, tables, _debug_shading(ig), rays.size, 1)
    _write_rays(rfile, rays)
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        ent, prim, t, _, _ = dev.trace(rays)
    hit = prim >= 0
    for mode, col in ((15, prim), (17, ent)):  # DEBUG_PRIM_ID_RAW / DEBUG_ENTITY_ID_RAW
        # --spp 3 = three iterations over the same rays; the output is the film divided by the iteration count
        _run(os.path.join(host_bin, "igtrace"), scene, "-i", rfile, "-o", out, "--technique", "debug", "--debug-mode", str(mode), "--spp", "3")
        got = np.loadtxt(out, dtype=np.float64).reshape(-1, 3)
        assert np.array_equal(got[hit, 0], col[hit].astype(np.float64)) and np.all(got[~hit] == 0)
    # stdin / stdout form, short lines ignored, list ends at the first empty line (trace/main.cpp:23-38)
    text = open(rfile).read().splitlines()
    feed = "\n".join(text[:5] + ["1 2 3"] + text[5:10] + ["", text[10]]) + "\n"
    r = subprocess.run([os.path.join(host_bin, "igtrace"), scene, "--technique", "debug", "--debug-mode", "15"], input=feed,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    got = np.loadtxt(r.stdout.splitlines(), dtype=np.float64).reshape(-1, 3)
    assert got.shape[0] == 10
    assert np.array_equal(got[:, 0], np.where(hit[:10], prim[:10], 0).astype(np.float64))


def test_igtrace_path_contributions_match_the_binding(ig, host_bin, tmp_path):
    tables, sh, w, h, _ = load_golden(ig, "cbox-d6")
    rng = np.random.default_rng(5)
    n = 4096
    rays = ig.make_rays(rng.uniform([100, 100, 100], [450, 450, 450], (n, 3)), rng.normal(size=(n, 3)), 0.0, 0.0)
    rays["tmax"] = np.finfo(np.float32).max
    scene, rfile, out = str(tmp_path / "cbox.igdb"), str(tmp_path / "rays.txt"), str(tmp_path / "out.txt")
    ig.fixtures.save_igdb(scene, tables, sh, w, h)
    _write_rays(rfile, rays)
    _run(os.path.join(host_bin, "igtrace"), scene, "-i", rfile, "-o", out, "--spp", "4", "--seed", "11")
    got = np.loadtxt(out, dtype=np.float64).reshape(-1, 3)
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        dev.assign_shading(sh)
        for it in range(4):
            dev.render(spi=1, rays=rays, iteration=it, seed=11)
        film, iters = dev.framebuffer()
    assert iters == 4 and got.shape == (n, 3) and got.sum() > 0
    np.testing.assert_allclose(got, film[0].astype(np.float64) / 4, rtol=2e-6, atol=1e-9)


@pytest.mark.parametrize("name,size,batch", [("cbox-d6", (128, 128), False), ("cbox-d6", (128, 128), True),
                                             ("many-point-lights", (480, 270), True)])  # textures, sky, hierarchy in the blob
def test_igcli_image_matches_the_binding(ig, host_bin, tmp_path, name, size, batch):
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    import cv2
    tables, sh, w, h, _ = load_golden(ig, name)
    scene, out = str(tmp_path / "scene.igdb"), str(tmp_path / "out.exr")
    ig.fixtures.save_igdb(scene, tables, sh, w, h)
    sw, sh_ = size
    cmd = [os.path.join(host_bin, "igcli"), scene, "-o", out, "--spp", "8", "--spi", "4", "--seed", "7", "--width", str(sw), "--height", str(sh_)]
    r = _run(*(cmd + (["--batch"] if batch else [])))
    assert "Iterations: 2" in r.stdout and "SPP: 8" in r.stdout and "(min/med/max Msamples/s)" in r.stdout  # cli/main.cpp:170-196
    img = cv2.imread(out, cv2.IMREAD_UNCHANGED)[:, :, ::-1]
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        dev.assign_shading(sh)
        dev.resize(sw, sh_)
        for it in range(2):
            dev.render(sw, sh_, spi=4, iteration=it, seed=7)
        film, iters = dev.framebuffer()
    assert iters == 2 and img.shape == (sh_, sw, 3) and img.mean() > 0.01
    # the same samples; only the order of the f32 atomic additions into the film differs
    np.testing.assert_allclose(img, film / 2, rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("batch", [False, True])
def test_igcli_gpus_2_image_equals_one_gpu_image(ig, host_bin, tmp_path, batch):
    """`igcli --gpus 2` (IG::Device over igb200_group_*): per-iteration steps split by tile rows (1 iteration < 2 GPUs), the
    --batch form by iteration index; either way the EXR equals the one-GPU EXR up to f32 summation order."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    import cv2
    tables, sh, w, h, _ = load_golden(ig, "cbox-d6")
    scene = str(tmp_path / "scene.igdb")
    ig.fixtures.save_igdb(scene, tables, sh, w, h)
    imgs = []
    for gpus in (1, 2):
        out = str(tmp_path / f"out{gpus}.exr")
        cmd = [os.path.join(host_bin, "igcli"), scene, "-o", out, "--spp", "16", "--spi", "4", "--seed", "7", "--width", "160", "--height", "120",
               "--gpus", str(gpus)] + (["--batch"] if batch else [])
        r = _run(*cmd)
        assert "Iterations: 4" in r.stdout and ("GPUs: 2" in r.stdout) == (gpus == 2)
        imgs.append(cv2.imread(out, cv2.IMREAD_UNCHANGED))
    np.testing.assert_allclose(imgs[1], imgs[0], rtol=1e-4, atol=1e-6)


def test_igcli_stats_reports_the_reference_categories(ig, host_bin, tmp_path):
    """`--stats`: Device::getStatistics with the reference's ShaderType / SectionType / Quantity names (Statistics.h:9-63),
    filled from CUDA events between the kernels of the wavefront rounds (igb200_get_section_stats)."""
    tables, sh, w, h, _ = load_golden(ig, "cbox-d6")
    scene, out = str(tmp_path / "scene.igdb"), str(tmp_path / "out.exr")
    ig.fixtures.save_igdb(scene, tables, sh, w, h)
    r = _run(os.path.join(host_bin, "igcli"), scene, "-o", out, "--spp", "8", "--spi", "4", "--width", "96", "--height", "96", "--stats")
    for name in ("PrimaryTraversal", "SecondaryTraversal", "RayGeneration", "Hit", "GPUSortPrimary", "FramebufferHostUpdate"):
        assert name in r.stdout, r.stdout
    assert f"CameraRays> {96 * 96 * 8}" in r.stdout
    # the stats mode renders the same image as the graph mode
    out2 = str(tmp_path / "out2.exr")
    _run(os.path.join(host_bin, "igcli"), scene, "-o", out2, "--spp", "8", "--spi", "4", "--width", "96", "--height", "96")
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    import cv2
    np.testing.assert_allclose(cv2.imread(out, cv2.IMREAD_UNCHANGED), cv2.imread(out2, cv2.IMREAD_UNCHANGED), rtol=1e-4, atol=1e-6)


def test_igcli_renders_a_scene_json_without_python(ig, host_bin, tmp_path):
    """`igcli scene.json`: the C++ loader (host/SceneLoader.cpp) builds the tables, no .igdb and no Python in the path.  The image
    equals the one rendered from the tables the C++ loader wrote out (json2igdb -> binding); the demo scene is this repo's own."""
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    import cv2
    from test_scene_loader_cpp import _write_pyramid
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    scene = str(tmp_path / "loader_demo.json")
    open(scene, "w").write(open(os.path.join(root, "tests", "scenes", "loader_demo.json")).read())
    _write_pyramid(str(tmp_path / "loader_demo_pyramid.ply"))
    out, igdb = str(tmp_path / "out.exr"), str(tmp_path / "demo.igdb")
    r = _run(os.path.join(host_bin, "igcli"), scene, "-o", out, "--spp", "16", "--spi", "4", "--seed", "3")
    assert "Iterations: 4" in r.stdout
    img = cv2.imread(out, cv2.IMREAD_UNCHANGED)[:, :, ::-1]
    assert img.shape == (120, 160, 3) and np.all(np.isfinite(img)) and img.mean() > 0.02
    _run(os.path.join(host_bin, "igdbtool"), "json2igdb", scene, igdb)
    out2 = str(tmp_path / "out2.exr")
    _run(os.path.join(host_bin, "igcli"), igdb, "-o", out2, "--spp", "16", "--spi", "4", "--seed", "3")
    np.testing.assert_allclose(img, cv2.imread(out2, cv2.IMREAD_UNCHANGED)[:, :, ::-1], rtol=1e-4, atol=1e-6)
    # every light type of the scene contributes: the lamp (area), the point and spot lights and the sky
    assert img[:, :, 2].mean() > 0.01 and img.max() > 1.0
