"""Multi-GPU product path (ABI v6) on hardware.

One-device tests (always run): the partitions of igb200_render_partition, rendered one after the other into ONE film, are the
single-device image -- this pins the iteration split and the interleaved 16-pixel tile-row split themselves.
Two-device tests (skipped when the box has one GPU; run with `gpurun --gpus 2`): DeviceGroup (one process, ncclCommInitAll,
one host thread + stream per device, ncclReduce onto member 0) equals the single-device film to f32 summation order
(SURVEY.md S7 gate: max relative difference <= 1e-5), for both partitionings, plus the ray-range split of igtrace batches."""
import numpy as np
import pytest

from render_common import room_scene

pytestmark = pytest.mark.gpu


def _gpu_count():
    import torch
    return torch.cuda.device_count()


def _single(ig, tables, sh, w, h, spi, iterations, seed):
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        dev.assign_shading(sh)
        dev.resize(w, h)
        dev.render(w, h, spi=spi, iteration=0, seed=seed, iterations=iterations)
        film, n = dev.framebuffer()
        st = dev.render_stats()
    return film, n, st


def _close(a, b):
    scale = np.maximum(np.abs(b), 1e-3 * float(np.abs(b).mean()))
    return float(np.max(np.abs(a - b) / scale))


@pytest.mark.parametrize("mode,parts,h", [(1, 3, 72), (2, 3, 72), (2, 4, 70), (2, 8, 40)])   # 70: partial last tile row; 8 > tile rows of 40
def test_partitions_on_one_device_sum_to_the_whole(ig, mode, parts, h):
    w, spi, iterations, seed = 96, 2, 6, 17
    tables, sh = room_scene(ig, w, depth=4)
    ref, n_ref, st_ref = _single(ig, tables, sh, w, h, spi, iterations, seed)
    cam = 0
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        dev.assign_shading(sh)
        dev.resize(w, h)
        for p in range(parts):
            dev.render_partition(w, h, spi, iterations, p, parts, mode, seed=seed)
            cam += dev.render_stats()["camera_rays"]
        film, n = dev.framebuffer()
    assert cam == st_ref["camera_rays"] == w * h * spi * iterations    # every sample exactly once
    assert n == (iterations if mode == 1 else iterations * parts)       # per-call counts; the reducer sets the group's
    assert _close(film, ref) <= 1e-5


def test_tile_rows_are_disjoint_and_interleaved(ig):
    w, h = 64, 80   # 5 tile rows
    tables, sh = room_scene(ig, w, depth=2)
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        dev.assign_shading(sh)
        owned = []
        for p in range(2):
            dev.resize(w, h)
            dev.render_partition(w, h, 1, 1, p, 2, 2, seed=3)
            film, _ = dev.framebuffer()
            owned.append(film.sum(axis=(1, 2)) > 0)
    rows = np.arange(h) // 16
    assert np.array_equal(owned[0], rows % 2 == 0) and np.array_equal(owned[1], rows % 2 == 1)


@pytest.mark.skipif(_gpu_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("mode,iterations", [(1, 6), (2, 6), (0, 1), (0, 5)])
def test_group_film_equals_single_device_film(ig, mode, iterations):
    n_dev = min(_gpu_count(), 8)
    w, h, spi, seed = 160, 120, 2, 23
    tables, sh = room_scene(ig, w, depth=6)
    ref, n_ref, st_ref = _single(ig, tables, sh, w, h, spi, iterations, seed)
    with ig.DeviceGroup(n_dev) as grp:
        grp.assign_scene(tables)
        grp.assign_shading(sh)
        grp.resize(w, h)
        grp.render(w, h, spi, iterations, mode, seed=seed)
        film, n = grp.framebuffer()
        st = grp.render_stats()
        assert n == iterations == n_ref
        assert _close(film, ref) <= 1e-5
        assert (st["camera_rays"], st["bounce_rays"], st["shadow_rays"]) == tuple(st_ref[k] for k in ("camera_rays", "bounce_rays", "shadow_rays"))
        assert st["film_reduce_ms"] > 0
        # a second request accumulates on the root like a second Device::render would
        grp.render(w, h, spi, iterations, mode, iteration=iterations, seed=seed)
        film2, n2 = grp.framebuffer()
        assert n2 == 2 * iterations and film2.sum() > 1.5 * film.sum()
        for m in grp.members[1:]:
            assert m.framebuffer()[1] == 0 and not m.framebuffer()[0].any()   # non-root films start from zero again


@pytest.mark.skipif(_gpu_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_group_trace_equals_single_device_trace(ig):
    tables = ig.workloads.c3_scene(grid=2, stacks=16, slices=32)
    rays = ig.workloads.c3_rays(tables, n=300_001)
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        ref = dev.trace(rays)
    with ig.DeviceGroup(min(_gpu_count(), 8)) as grp:
        grp.assign_scene(tables)
        got = grp.trace(rays)
    for a, b in zip(got, ref):
        assert np.array_equal(a, b)
