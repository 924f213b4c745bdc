"""N > 1 host logic on CPU (gloo, world_size 2): splitting the iterations over ranks and summing the films on rank 0
gives the single-process image (the seed hashes the iteration index, src/artic/core/random.art:34-43,
so the union of the ranks' samples IS the single-GPU sample set)."""
import os
import subprocess
import sys

import numpy as np

from render_common import room_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_iteration_split_reduces_to_single_process_image(ig, oracle, tmp_path):
    out = str(tmp_path / "film.npz")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                           "--master-addr", "127.0.0.1", "--master-port", "29631",
                           os.path.join(ROOT, "tests", "dist_film_worker.py"), out], env=env, timeout=600)
    z = np.load(out)
    assert int(z["iterations"]) == 6
    tables, sh = room_scene(ig, 40, depth=4)
    ref = oracle.OracleScene(tables).render(sh, 40, 40, spi=2, iterations=6, seed=4)
    np.testing.assert_allclose(z["film"], ref, rtol=1e-5, atol=1e-6)  # f32 summation order only
