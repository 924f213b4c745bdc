"""bench.py contract on CPU: the reference arm (--impl reference = the oracle on the host cores) prints ONE JSON line with the keys
the driver reads, on the GPU arm's metric / unit / config; and the GPU arm fails loudly (no CPU fallback) when there is no GPU."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=600, cwd=ROOT)


def test_reference_arm_line():
    p = _run("--impl", "reference", "--rays", "200000", "--steps", "2", "--warmup", "1", "--pt-iterations", "0", "--c5-spp", "0",
             "--office-spp", "0")
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1                                   # exactly one JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Mrays/s closest-hit (incoherent)" and d["unit"] == "Mrays/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1
    assert d["config"]["workload"] == "c3-instanced" and d["config"]["rays_per_step"] == 200000
    assert d["value"] > 0 and abs(d["value"] - 200000 / d["ms_per_step"] / 1e3) < 1e-6 * d["value"] + 1e-9
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "200000" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0 and d["vs_baseline"] is None


def test_gpu_arm_needs_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = _run("--rays", "1000", "--steps", "1", "--warmup", "0", "--pt-iterations", "0", "--c5-spp", "0", "--office-spp", "0")
    assert p.returncode != 0                                 # the product path has no CPU fallback
    assert not [l for l in p.stdout.splitlines() if l.strip().startswith("{")]


def test_reference_arm_under_torchrun_prints_once():
    """Launched like the GPU arm for N > 1 (torchrun, one process per rank): rank 0 alone runs and prints, the others exit 0."""
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29541", os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--rays", "100000",
                        "--steps", "1", "--warmup", "1", "--pt-iterations", "0", "--c5-spp", "0", "--office-spp", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip().startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["config"]["rays_per_step"] == 100000
