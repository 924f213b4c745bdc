"""Worker of tests/test_multi_gpu_host.py: the N-rank film combine of bench.py (iteration split + sum onto rank 0),
run on CPU with the gloo backend and the oracle standing in for the device."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    out = sys.argv[1]
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    ig = importlib.import_module("ignis-masterthesis_b200")
    import pyoracle
    from render_common import room_scene
    tables, sh = room_scene(ig, 40, depth=4)
    osc = pyoracle.OracleScene(tables)
    iterations = 6
    film = np.zeros((40, 40, 3), np.float32)
    per = iterations // world  # the split of bench.py::bench_path_tracer: rank r renders [r * per, (r + 1) * per)
    osc.render(sh, 40, 40, spi=2, iterations=per, seed=4, first_iteration=rank * per, film=film, threads=2)
    t = torch.from_numpy(film)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    counts = torch.tensor([per])
    dist.reduce(counts, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.savez(out, film=t.numpy(), iterations=int(counts[0]))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
