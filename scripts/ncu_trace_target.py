"""ncu target: device-resident closest-hit launches over the full C3 batch (the kernel bench.py's roofline describes).
usage: python scripts/ncu_trace_target.py [launches] [rays]"""
import importlib, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("ignis-masterthesis_b200")
launches = int(sys.argv[1]) if len(sys.argv) > 1 else 3
n = int(sys.argv[2]) if len(sys.argv) > 2 else 16 * 1024 * 1024
tables = pkg.workloads.c3_scene()
rays = pkg.workloads.c3_rays(tables, n=n)
with pkg.Device(0) as dev:
    dev.assign_scene(tables)
    d_rays = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).cuda()
    d_i = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(2)]
    d_f = [torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(3)]
    s = torch.cuda.current_stream()
    for _ in range(launches):
        dev.trace_device(d_rays.data_ptr(), n, d_i[0].data_ptr(), d_i[1].data_ptr(), d_f[0].data_ptr(), d_f[1].data_ptr(),
                         d_f[2].data_ptr(), stream=s.cuda_stream)
    torch.cuda.synchronize()
    print("hits", int((d_i[1] >= 0).sum()))
