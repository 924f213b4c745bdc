"""SURVEY N4 on the GPU, through the C ABI: MIS AOVs, infobuffer AOVs, named framebuffer access with per-image iteration
counts (Device.cpp:1300-1307,1389-1403,1460-1472) and the tonemap pipeline -- each against the oracle."""
import numpy as np
import pytest

from render_common import flat_scene, load_golden, room_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev(ig):
    d = ig.Device(0)
    yield d
    d.close()


def _close(a, b, far_limit=5e-3):
    scale = np.maximum(np.abs(b), 1e-3 * max(float(np.abs(b).mean()), 1e-12))
    rel = np.abs(a - b) / scale
    assert float((rel > 1e-3).mean()) <= far_limit and float(np.median(rel)) <= 1e-5, (float((rel > 1e-3).mean()), float(np.median(rel)))


def test_mis_aovs_match_oracle_and_partition_the_film(ig, oracle, dev):
    tables, sh, _, _, _ = load_golden(ig, "cbox-d6")
    w = h = 128
    dev.assign_scene(tables)
    dev.assign_shading(sh)
    dev.resize(w, h)
    with pytest.raises(RuntimeError, match="unknown AOV"):
        dev.framebuffer(name="Direct Weights")               # not enabled by this technique (Device.cpp:1303-1305)
    for it in range(2):
        dev.render(w, h, spi=4, iteration=it, seed=7)
    plain, _ = dev.framebuffer()
    sh.technique.flags |= ig.shading.TECHFLAG_AOV_MIS
    dev.assign_shading(sh)
    dev.clear_framebuffer()
    for it in range(2):
        dev.render(w, h, spi=4, iteration=it, seed=7)
    film, n = dev.framebuffer()
    di, n_di = dev.framebuffer(name="Direct Weights")
    nee, n_nee = dev.framebuffer(name="NEE Weights")
    assert (n, n_di, n_nee) == (2, 2, 2)
    assert np.allclose(film, plain, rtol=1e-5, atol=1e-6)    # same samples; only the atomic add order may differ
    assert np.allclose(di + nee, film, rtol=1e-4, atol=1e-5)
    aov = [np.zeros((h, w, 3), np.float32) for _ in range(2)]
    ref = oracle.OracleScene(tables).render(sh, w, h, spi=4, iterations=2, seed=7, aovs=aov)
    _close(film, ref); _close(di, aov[0]); _close(nee, aov[1])
    with pytest.raises(RuntimeError, match="unknown AOV"):
        dev.framebuffer(name="Normals")
    # Device::clearFramebuffer(name): one image, its own counter
    dev.clear_framebuffer("NEE Weights")
    nee, n_nee = dev.framebuffer(name="NEE Weights")
    assert n_nee == 0 and not nee.any() and dev.framebuffer()[1] == 2 and dev.framebuffer(name="Direct Weights")[1] == 2


@pytest.mark.parametrize("handle_specular", [False, True])
def test_infobuffer_matches_oracle_and_locks_the_film(ig, oracle, dev, handle_specular):
    tables, sh = room_scene(ig, 96)
    w = h = 96
    dev.assign_scene(tables)
    sh.set_path(max_depth=4)
    dev.assign_shading(sh)
    dev.resize(w, h)
    dev.render(w, h, spi=2, iteration=0, seed=1)
    before, n_before = dev.framebuffer()
    sh.set_infobuffer(max_depth=8, handle_specular=handle_specular)
    dev.assign_shading(sh)                                   # the extra technique variant of the denoiser flow
    dev.render(w, h, spi=1, iteration=0, seed=1)             # OverrideSPI = 1 (InfoBufferTechnique.cpp:37)
    after, n_after = dev.framebuffer()
    assert n_after == n_before == 1 and np.array_equal(before, after)   # LockFramebuffer
    st = dev.render_stats()
    aov = [np.zeros((h, w, 3), np.float32) for _ in range(3)]
    osc = oracle.OracleScene(tables)
    osc.render(sh, w, h, spi=1, iterations=1, seed=1, aovs=aov)
    assert (st["camera_rays"], st["shadow_rays"]) == (w * h, 0) and abs(st["bounce_rays"] - int(osc.ray_stats[1])) <= 2
    assert (st["bounce_rays"] > 0) == handle_specular
    for k, name in enumerate(("Normals", "Albedo", "Depth")):
        img, n = dev.framebuffer(name=name)
        assert n == 1
        _close(img, aov[k])


def test_infobuffer_closed_form_full_size(ig, dev):
    res = 1000
    tables, sh = flat_scene(ig, "env", res)
    sh.set_infobuffer(max_depth=8)
    dev.assign_scene(tables)
    dev.assign_shading(sh)
    dev.resize(res, res)
    dev.render(res, res, spi=1, iteration=0)
    nrm, _ = dev.framebuffer(name="Normals")
    alb, _ = dev.framebuffer(name="Albedo")
    dep, _ = dev.framebuffer(name="Depth")
    assert np.allclose(nrm, np.broadcast_to(np.float32([0, 0, -1]), nrm.shape), atol=1e-6)
    ok = np.isclose(alb, 1 / np.pi, rtol=1e-5)               # a cosine sample with cos = 0 is rejected and drops 1 / 16
    assert ok.mean() > 0.999 and alb.max() <= (1 + 1e-5) / np.pi and alb.min() >= (14 / 16) / np.pi
    # distance to the plane z = 0 from (0, 0, -1) through the pixel (jittered inside it): 1 / cos
    y, x = np.mgrid[0:res, 0:res]
    lo = np.sqrt(1 + (np.minimum(np.abs(2 * x / res - 1), np.abs(2 * (x + 1) / res - 1))) ** 2
                 + (np.minimum(np.abs(2 * y / res - 1), np.abs(2 * (y + 1) / res - 1))) ** 2)
    hi = np.sqrt(1 + (np.maximum(np.abs(2 * x / res - 1), np.abs(2 * (x + 1) / res - 1))) ** 2
                 + (np.maximum(np.abs(2 * y / res - 1), np.abs(2 * (y + 1) / res - 1))) ** 2)
    assert np.all(dep[..., 0] >= lo - 1e-5) and np.all(dep[..., 0] <= hi + 1e-5)


@pytest.mark.parametrize("method", [0, 1, 2, 3, 4])
@pytest.mark.parametrize("gamma", [False, True])
def test_tonemap_matches_oracle(ig, oracle, dev, method, gamma):
    tables, sh, _, _, _ = load_golden(ig, "cbox-d6")
    w = h = 256
    dev.assign_scene(tables)
    dev.assign_shading(sh)
    dev.resize(w, h)
    for it in range(3):
        dev.render(w, h, spi=4, iteration=it, seed=2)
    film, n = dev.framebuffer()
    got = dev.tonemap(method=method, use_gamma=gamma, scale=1.5, exposure_factor=1.25, exposure_offset=0.01)
    want = oracle.tonemap(film, n, method=method, use_gamma=gamma, scale=1.5, exposure_factor=1.25, exposure_offset=0.01)
    assert got.shape == (h, w) and np.all(got >> 24 == 255)
    gb = got.view(np.uint8).reshape(h, w, 4).astype(np.int32)
    wb = want.view(np.uint8).reshape(h, w, 4).astype(np.int32)
    # powf / division rounding may move a value across an integer boundary of c * 255: at most one step, rarely
    assert np.abs(gb - wb).max() <= 1 and float((gb != wb).mean()) < 1e-3
    assert len(np.unique(got)) > 1000


def test_tonemap_marks_bad_pixels_and_reads_aovs(ig, oracle, dev):
    import torch
    tables, sh = flat_scene(ig, "env", 64)
    sh.set_infobuffer(max_depth=2)
    dev.assign_scene(tables)
    dev.assign_shading(sh)
    dev.resize(64, 64)
    dev.render(64, 64, spi=1, iteration=0)
    n = dev.tonemap(aov="Normals", method=0, use_gamma=False)
    assert np.all(n == 0xFF000000)                           # (0, 0, -1): X + Y + Z <= eps reads as black (core/color.art:98)
    a = dev.tonemap(aov="Albedo", method=0, use_gamma=False)
    assert np.all(a == (0xFF000000 | (81 << 16) | (81 << 8) | 81))   # 1 / pi * 255 = 81.17
    bad = torch.zeros(64 * 64 * 3, dtype=torch.float32, device="cuda")
    bad[0] = float("nan"); bad[3] = float("inf"); bad[6:9] = 0.5; bad[12] = -1.0; bad[14] = 2.0
    sh.set_path(max_depth=2)
    dev.assign_shading(sh)
    dev.render(64, 64, spi=1, iteration=0)                   # the film exists and counts one iteration
    dev.set_framebuffer(bad.data_ptr(), bad.numel(), 1)
    t = dev.tonemap(method=0, use_gamma=False).reshape(-1)
    assert t[0] == 0xFF00FFFF and t[1] == 0xFFFF0096 and t[2] == 0xFF7F7F7F and t[3] == 0xFF000000
    assert t[4] == 0xFFFFFF00                                # (-1, 0, 2): negative chromaticity reads yellow (tonemap.art:27-28)
