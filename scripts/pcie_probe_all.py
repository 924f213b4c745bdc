"""Per-GPU PCIe bandwidth with 1 and with N GPUs copying at once (does the host path of the 8-GPU VM saturate?).
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P scripts/pcie_probe_all.py
Every rank moves the bytes of one 16 Mi-ray igb200_trace_rays step (512 MiB up, 320 MiB down, both directions at once, pinned
host memory) -- first one rank at a time while the others idle, then all ranks together.  Rank 0 prints one table."""
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")
up, down = 512 << 20, 320 << 20
h_in, h_out = torch.empty(up, dtype=torch.uint8).pin_memory(), torch.empty(down, dtype=torch.uint8).pin_memory()
d_in, d_out = torch.empty(up, dtype=torch.uint8, device="cuda"), torch.empty(down, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def both():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def timed(reps):
    both()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        both()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps


alone = 0.0
for r in range(world):  # one rank at a time
    dist.barrier()
    if r == rank:
        alone = timed(20)
dist.barrier()
together = timed(60)  # all ranks at once
dist.barrier()
res = torch.tensor([alone, together], dtype=torch.float64)
out = [torch.zeros(2, dtype=torch.float64) for _ in range(world)]
dist.all_gather(out, res)
if rank == 0:
    print(f"# {world} GPUs; per step 512 MiB H2D + 320 MiB D2H at once; GB/s = (up + down) bytes / time; Mrays/s = 16 Mi rays / time")
    print("rank  alone ms  alone GB/s  alone Mrays/s | together ms  GB/s  Mrays/s")
    tot = 0.0
    for r, o in enumerate(out):
        a, t = float(o[0]), float(o[1])
        tot += (up + down) / t / 1e9
        print(f"{r:4d}  {a * 1e3:8.2f}  {(up + down) / a / 1e9:10.1f}  {16.777216 / a:13.0f} | {t * 1e3:11.2f}  {(up + down) / t / 1e9:4.1f}  {16.777216 / t:7.0f}")
    print(f"aggregate with all {world} GPUs copying: {tot:.1f} GB/s; sum of the 'alone' rates: {sum((up + down) / float(o[0]) / 1e9 for o in out):.1f} GB/s")
dist.destroy_process_group()
