"""Device-resident closest-hit throughput of small launches (C3 scene): rays per launch 64 Ki .. 16 Mi.  One GPU."""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("ignis-masterthesis_b200")
W = pkg.workloads
tables = W.c3_scene()
rays = W.c3_rays(tables, n=1 << 24, seed=0xB200)
dev = pkg.Device(0); dev.assign_scene(tables)
d_rays = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).cuda()
n_max = rays.size
d_i = [torch.empty(n_max, dtype=torch.int32, device="cuda") for _ in range(2)]
d_f = [torch.empty(n_max, dtype=torch.float32, device="cuda") for _ in range(3)]
s = torch.cuda.current_stream()
for n in (1 << 16, 1 << 17, 1 << 18, 1 << 20, 1 << 21, 1 << 22, 1 << 24):
    def one():
        dev.trace_device(d_rays.data_ptr(), n, d_i[0].data_ptr(), d_i[1].data_ptr(), d_f[0].data_ptr(), d_f[1].data_ptr(), d_f[2].data_ptr(), stream=s.cuda_stream)
    for _ in range(5): one()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(5, min(200, (1 << 26) // n))
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): one()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{n:>9d} rays per launch: {ms * 1e3:8.1f} us, {n / ms / 1e3:7.0f} Mrays/s", flush=True)
dev.close()
