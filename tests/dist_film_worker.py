"""Worker of tests/test_multi_gpu_host.py: the product's request partitioning (igb200_partition_share through the package:
iteration blocks or interleaved 16-pixel tile rows) + a sum onto rank 0, run on CPU with the gloo backend.  The ORACLE
stands in for the device here (test infrastructure): what is under test is the host arithmetic of the split and the
iteration-count bookkeeping, not a renderer."""
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
    out, mode = sys.argv[1], int(sys.argv[2])
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    ig = importlib.import_module("ignis-masterthesis_b200")
    import pyoracle
    from render_common import room_scene
    w = h = 40
    tables, sh = room_scene(ig, w, depth=4)
    osc = pyoracle.OracleScene(tables)
    iterations = 5 if mode != 2 else 2
    first, count, tile_rank, tile_count = ig.partition_share(iterations, rank, world, mode)
    film = np.zeros((h, w, 3), np.float32)
    if count:
        osc.render(sh, w, h, spi=2, iterations=count, seed=4, first_iteration=first, film=film, threads=2)
    if tile_count > 1:  # samples are per pixel (the seed hashes x, y): rendering only the owned rows == masking the others
        rows = np.arange(h) // 16
        film[rows % tile_count != tile_rank] = 0
    t = torch.from_numpy(film)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    counts = torch.tensor([count if tile_count == 1 else (count if rank == 0 else 0)])
    dist.reduce(counts, dst=0, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.savez(out, film=t.numpy(), iterations=int(counts[0]), requested=iterations)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
