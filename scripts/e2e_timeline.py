"""Per-chunk timeline of one igb200_trace_rays call on the benchmark batch (IGB200_PIPE_TIMELINE=1 -> stderr).  One GPU."""
import importlib, os, sys
import numpy as np, torch
os.environ["IGB200_PIPE_TIMELINE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("ignis-masterthesis_b200")
W = pkg.workloads
n = 16_777_216
tables = W.c3_scene()
rays = W.c3_rays(tables, n=n, seed=0xB200)
h_rays = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).pin_memory()
h_rays_np = h_rays.numpy().view(pkg.tables.RAY_DTYPE).reshape(-1)
outs = [torch.empty(n, dtype=torch.int32).pin_memory().numpy() for _ in range(2)] + [torch.empty(n, dtype=torch.float32).pin_memory().numpy() for _ in range(3)]
dev = pkg.Device(0)
dev.assign_scene(tables)
for _ in range(3):
    dev.trace_into(h_rays_np, *outs)
print("# ids + t only", file=sys.stderr, flush=True)
dev.trace_into(h_rays_np, outs[0], outs[1], outs[2], None, None)
dev.close()
