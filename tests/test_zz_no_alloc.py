"""igb200_trace_rays_noalloc (no counterpart in the reference): the host-buffer
trace with all scratch in a static device array, against the existing igb200_trace_rays through the C ABI.

Same kernel, same inputs => the tolerance is the one tests/test_trace_gpu.py uses for the same triangle: bit-exact ids, t, u, v.
Two calls in a row (the second smaller, on other rays and any-hit as well) show that the shared array's left-over contents
do not leak into a later call.  The file name makes pytest collect it last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _existing(dev, rays, any_hit=False):
    return dev.trace(rays, any_hit=any_hit)


def _noalloc(dev, rays, any_hit=False, poison=True):
    n = rays.size
    ent, prim = np.empty(n, np.int32), np.empty(n, np.int32)
    t, u, v = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
    if poison:
        ent.fill(-77); prim.fill(-77); t.fill(np.nan); u.fill(np.nan); v.fill(np.nan)
    dev.trace_noalloc_into(rays, ent, prim, t, u, v, any_hit=any_hit)
    return ent, prim, t, u, v


def _same(a, b):
    for x, y in zip(a, b):
        assert x.dtype == y.dtype and np.array_equal(x, y)   # bit-exact: ids, t, u, v


def test_noalloc_equals_existing_entry_on_two_calls_in_a_row(ig):
    tables = ig.workloads.c3_scene(grid=2, stacks=16, slices=32)
    rays_a = ig.workloads.c3_rays(tables, n=ig.Device.NOALLOC_MAX_RAYS, seed=0xA110C)     # exactly the maximum
    rays_b = ig.workloads.c3_dense_rays(tables, n=20_001, seed=0xB0B, grid=2)              # smaller, nearly all hit
    with ig.Device(0) as dev:
        dev.assign_scene(tables)
        ref_a, ref_b = _existing(dev, rays_a), _existing(dev, rays_b)
        ref_b_any = _existing(dev, rays_b, any_hit=True)
        assert (ref_a[1] >= 0).any() and (ref_a[1] < 0).any() and (ref_b[1] >= 0).mean() > 0.9
        _same(_noalloc(dev, rays_a), ref_a)          # call 1
        _same(_noalloc(dev, rays_b), ref_b)          # call 2, right after: the array still holds call 1's rays and hits
        got_any = _noalloc(dev, rays_b, any_hit=True)
        assert np.array_equal(got_any[1] >= 0, ref_b_any[1] >= 0)
        _same(_noalloc(dev, rays_a), ref_a)          # and back to the large batch
        # a second handle on the same GPU shares the same array (serialised here): same answers
        with ig.Device(0) as other:
            other.assign_scene(tables)
            _same(_noalloc(other, rays_b), ref_b)
        # subset of the columns (NULL = not wanted)
        prim = np.empty(rays_b.size, np.int32)
        tt = np.empty(rays_b.size, np.float32)
        dev.trace_noalloc_into(rays_b, None, prim, tt, None, None)
        assert np.array_equal(prim, ref_b[1]) and np.array_equal(tt, ref_b[2])


def test_noalloc_limits_and_errors(ig):
    B = ig._lib.BackendError
    tables = ig.workloads.c1_scene()
    with ig.Device(0) as dev:
        with pytest.raises(B) as e:                      # no scene yet
            _noalloc(dev, ig.workloads.c1_rays(8), poison=False)
        assert e.value.code == -5
        dev.assign_scene(tables)
        too_many = ig.workloads.c1_rays(ig.Device.NOALLOC_MAX_RAYS + 1)
        with pytest.raises(B) as e:
            _noalloc(dev, too_many, poison=False)
        assert e.value.code == -6 and "65536" in str(e.value) and "exceed" in str(e.value)
        _noalloc(dev, too_many[:0])                      # n == 0: nothing to do, no error
        _same(_noalloc(dev, too_many[:1000]), _existing(dev, too_many[:1000]))   # still works after the refusals
