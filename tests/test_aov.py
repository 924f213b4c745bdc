"""SURVEY N4 on the oracle: the path technique's MIS AOVs, the infobuffer technique (denoiser AOVs) and the tonemap
pipeline, pinned by construction properties and closed forms (the reference ships no golden for them)."""
import numpy as np
import pytest

from render_common import flat_scene, load_golden, room_scene


def test_mis_aovs_partition_the_film(ig, oracle):
    # pathtracer.art:128,158,206: every film splat also goes to exactly one of the two AOVs
    tables, sh, _, _, _ = load_golden(ig, "cbox-d6")
    osc = oracle.OracleScene(tables)
    plain = osc.render(sh, 64, 64, spi=8, iterations=2, seed=3)
    sh.technique.flags |= ig.shading.TECHFLAG_AOV_MIS
    aov = [np.zeros((64, 64, 3), np.float32) for _ in range(2)]
    film = osc.render(sh, 64, 64, spi=8, iterations=2, seed=3, aovs=aov)
    assert np.array_equal(film, plain)                       # the AOVs do not perturb the estimator
    assert aov[0].sum() > 0 and aov[1].sum() > 0
    assert np.allclose(aov[0] + aov[1], film, rtol=1e-5, atol=1e-6)
    lamp = film.max(axis=2) > 0.5 * film.max()               # the directly visible emitter lives in "Direct Weights"
    assert lamp.any() and np.all(aov[0][lamp].sum(axis=1) > 2 * aov[1][lamp].sum(axis=1))


def test_infobuffer_closed_form_on_the_flat_scene(ig, oracle):
    # technique/infobuffer.art:36-63 on a white lambertian plane at distance 1: shading normal, kd / pi (the BSDF sample
    # weight kd times pdf / cos), and the hit distance; the film stays locked
    res = 32
    tables, sh = flat_scene(ig, "env", res)
    sh.set_infobuffer(max_depth=8)
    aov = [np.zeros((res, res, 3), np.float32) for _ in range(3)]
    film = oracle.OracleScene(tables).render(sh, res, res, spi=1, iterations=1, seed=5, aovs=aov)
    assert not film.any()
    assert np.allclose(aov[0], np.broadcast_to(np.float32([0, 0, -1]), aov[0].shape), atol=1e-6)
    assert np.allclose(aov[1], 1 / np.pi, rtol=1e-5)
    d = aov[2][..., 0]
    assert np.array_equal(aov[2][..., 1], d) and d.min() >= 1 and d.max() <= np.sqrt(3) + 1e-5
    assert d[res // 2, res // 2] < 1.01 and d[0, 0] > 1.6


def test_infobuffer_follows_specular_paths_only_when_asked(ig, oracle):
    tables, sh = room_scene(ig, 48)
    osc = oracle.OracleScene(tables)
    out = {}
    for hs in (False, True):
        sh.set_infobuffer(max_depth=8, handle_specular=hs)
        aov = [np.zeros((48, 48, 3), np.float32) for _ in range(3)]
        osc.render(sh, 48, 48, spi=1, iterations=1, seed=5, aovs=aov)
        out[hs] = aov, osc.ray_stats.copy()
    (n0, a0, d0), s0 = out[False]
    (n1, a1, d1), s1 = out[True]
    assert s0[1] == 0 and s1[1] > 0                          # bounce rays only through specular surfaces
    changed = np.any(n0 != n1, axis=2)
    assert 0 < changed.mean() < 0.5                          # the glass / mirror pixels, nothing else
    ended = changed & (d1[..., 0] > 0)                       # chains that reach a non-specular surface within max_depth
    assert ended.any() and np.all(d1[ended][:, 0] > d0[ended][:, 0])   # the distance accumulates along the chain
    assert a0.max() <= 1 and a1.max() <= 1                   # color_saturate(albedo, 1); a delta BSDF gives weight / cos
    ln = np.linalg.norm(n1, axis=2)
    assert np.all((np.abs(ln - 1) < 1e-4) | (ln == 0))


def test_tonemap_known_answers(oracle):
    px = np.zeros((1, 8, 3), np.float32)
    px[0, 0] = 0.5; px[0, 1] = np.nan; px[0, 2] = np.inf; px[0, 3] = (-1, 0, 2); px[0, 7] = (-1, -1, -1); px[0, 4] = 4.0; px[0, 5] = (1, 0, 0)
    out = oracle.tonemap(px, 1, method=0, use_gamma=False)[0]
    assert out[0] == 0xFF7F7F7F                              # 0.5 * 255 = 127.5 truncates (make_byte_color)
    assert out[1] == 0xFF00FFFF and out[2] == 0xFFFF0096 and out[3] == 0xFFFFFF00   # NaN cyan, Inf pink, negative yellow
    assert out[4] == 0xFFFFFFFF and out[6] == 0xFF000000 and out[7] == 0xFF000000   # X + Y + Z <= eps reads as black (color.art:98)
    r = out[5]
    assert (r >> 16) & 0xFF >= 254 and (r >> 8) & 0xFF <= 1 and r & 0xFF <= 1       # xyY round trip of pure red
    # iteration count and scale divide / multiply the sums
    assert oracle.tonemap(px * 4, 4, method=0, use_gamma=False)[0, 0] == 0xFF7F7F7F
    assert oracle.tonemap(px, 1, method=0, use_gamma=False, scale=2.0)[0, 0] == 0xFFFFFFFF
    # sRGB transfer curve: 0.5 -> 1.055 * 0.5^(1/2.4) - 0.055 = 0.7354 -> 187
    assert oracle.tonemap(px, 1, method=0, use_gamma=True)[0, 0] == 0xFFBBBBBB
    # reinhard: L / (1 + L) on the luminance, chromaticity kept: grey 4.0 -> 0.8 -> 204
    assert oracle.tonemap(px, 1, method=1, use_gamma=False)[0, 4] & 0xFF in (203, 204)
    curves = [oracle.tonemap(px, 1, method=m, use_gamma=False)[0, 4] & 0xFF for m in range(5)]
    assert curves[0] == 255 and len(set(curves)) >= 4
