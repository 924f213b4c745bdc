"""Chunk schedule of the host-buffer pipeline (igb200_trace_rays): sweep IGB200_TRACE_CHUNK / IGB200_TRACE_CHUNK_MIN on the
benchmark batch (C3, 16 Mi rays, pinned host memory in and out, all five hit columns).  One GPU.
usage: python scripts/e2e_chunk_sweep.py ["big:min[:sms] big:min[:sms] ..."]     (ray counts, IGB200_PIPE_SMS; default list below)"""
import importlib
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("ignis-masterthesis_b200")
W = pkg.workloads
configs = (sys.argv[1] if len(sys.argv) > 1 else "1048576:262144 1048576:1048576 2097152:262144 4194304:262144 524288:131072 1048576:65536 2097152:524288 16777216:16777216").split()
n = 16_777_216
tables = W.c3_scene()
rays = W.c3_rays(tables, n=n, seed=0xB200)
h_rays = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).pin_memory()
h_rays_np = h_rays.numpy().view(pkg.tables.RAY_DTYPE).reshape(-1)
outs = [torch.empty(n, dtype=torch.int32).pin_memory().numpy() for _ in range(2)] + [torch.empty(n, dtype=torch.float32).pin_memory().numpy() for _ in range(3)]
for rep in range(int(os.environ.get('REPS', '2'))):
    for cfg in configs:
        big, small, sms = (cfg.split(":") + ["0"])[:3]
        os.environ["IGB200_TRACE_CHUNK"], os.environ["IGB200_TRACE_CHUNK_MIN"] = big, small
        os.environ.pop("IGB200_PIPE_SMS", None)
        if sms != "0":
            os.environ["IGB200_PIPE_SMS"] = sms
        dev = pkg.Device(0)
        dev.assign_scene(tables)
        for _ in range(2):
            dev.trace_into(h_rays_np, *outs)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            dev.trace_into(h_rays_np, *outs)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 5
        print(f"chunk {int(big):>9d} min {int(small):>9d} sms {int(sms):>3d}: {dt * 1e3:7.3f} ms per 16 Mi rays = {n / dt / 1e6:7.1f} Mrays/s", flush=True)
        dev.close()
