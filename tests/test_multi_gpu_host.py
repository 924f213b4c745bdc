"""N > 1 host logic on CPU (gloo, world_size 2): the product's partitioning of a request (igb200_partition_share) summed onto
rank 0 gives the single-process image (the seed hashes iteration index and pixel,
src/artic/core/random.art:34-43, so the union of the ranks' samples IS the single-GPU sample set)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from render_common import room_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("mode,port", [(1, 29631), (2, 29632), (0, 29633)])
def test_partitioned_request_reduces_to_single_process_image(ig, oracle, tmp_path, mode, port):
    out = str(tmp_path / "film.npz")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                           "--master-addr", "127.0.0.1", "--master-port", str(port),
                           os.path.join(ROOT, "tests", "dist_film_worker.py"), out, str(mode)], env=env, timeout=600)
    z = np.load(out)
    iterations = int(z["requested"])
    assert int(z["iterations"]) == iterations
    tables, sh = room_scene(ig, 40, depth=4)
    ref = oracle.OracleScene(tables).render(sh, 40, 40, spi=2, iterations=iterations, seed=4)
    np.testing.assert_allclose(z["film"], ref, rtol=1e-5, atol=1e-6)  # f32 summation order only


@pytest.mark.parametrize("total,parts", [(0, 3), (1, 4), (7, 3), (1024, 8), (5, 5)])
def test_partition_share_covers_the_request_exactly_once(ig, total, parts):
    # ITERATIONS: contiguous, disjoint, balanced blocks
    blocks = [ig.partition_share(total, p, parts, 1) for p in range(parts)]
    covered = np.zeros(total, int)
    for first, count, tr, tc in blocks:
        assert (tr, tc) == (0, 1)
        covered[first:first + count] += 1
    assert np.all(covered == 1)
    counts = [b[1] for b in blocks]
    assert max(counts) - min(counts) <= 1
    # TILE_ROWS: everyone renders everything of its own rows
    assert [ig.partition_share(total, p, parts, 2) for p in range(parts)] == [(0, total, p, parts) for p in range(parts)]
    # AUTO: iterations when there are enough of them
    auto = [ig.partition_share(total, p, parts, 0) for p in range(parts)]
    assert auto == (blocks if total >= parts else [(0, total, p, parts) for p in range(parts)])
    with pytest.raises(ig._lib.BackendError):
        ig.partition_share(total, parts, parts, 1)
