#!/usr/bin/env python
"""bench.py -- closest-hit ray throughput on the BASELINE.json config-3 workload.

Workload (config.workload "c3-instanced"): 64 instances of a 16 128-triangle sphere mesh = 1 032 192 triangles,
16 777 216 incoherent rays per step per GPU (uniform directions, generation order), closest-hit.
A step = one pass of the traversal hot path over one ray batch.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--rays R] [--impl ours|reference]

Prints ONE JSON line (rank 0).  `value` = whole-job Mrays/s with rays resident in HBM; `e2e` = the same
metric through the host-buffer C-ABI call igb200_trace_rays (H2D of the rays + D2H of the five hit columns
inside the timed region); `roofline` relates the kernel to measured HBM bandwidth via the algorithmic bytes
per ray of SURVEY.md 8(d); `cpu_baseline` = the oracle (restated reference CPU traversal) on the host cores.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s closest-hit (incoherent)"
UNIT = "Mrays/s"
FULL_RAYS = 16_777_216


def load_pkg():
    return importlib.import_module("ignis-masterthesis_b200")


def load_oracle():
    # bench.py's cpu_baseline / --impl reference legs are the only product-side users of oracle/
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle
    pyoracle.load()
    return pyoracle


def measured_traffic(n):
    """DRAM bytes (read + write) per launch of the dominant kernel from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            t = json.load(f)
        return (t["dram_bytes_read"] + t["dram_bytes_write"]) * (n / t["rays_per_launch"])
    except (OSError, KeyError, ValueError):
        return None


def ncu_counters():
    """The committed ncu capture of the dominant kernel (profiles/roofline_traffic.json): the denominators that actually bind
    it.  The kernel's working set lives in L1 / L2, so algorithmic bytes over HBM peak (the contract's `frac`) says little;
    issue-slot utilisation x SIMT efficiency and the L1 / L2 throughputs say where the time goes."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            t = json.load(f)
        keys = ("kernel_ms", "issue_slot_pct", "threads_per_inst", "alu_pipe_pct", "fma_pipe_pct", "l1_bytes", "l2_bytes",
                "l1_hit_pct", "l2_hit_pct", "local_memory_l1_bytes", "dram_bytes_read", "dram_bytes_write", "warp_insts", "source")
        out = {k: t[k] for k in keys if k in t}
        if "kernel_ms" in t:
            sec = t["kernel_ms"] * 1e-3
            out["dram_gbs"] = (t["dram_bytes_read"] + t["dram_bytes_write"]) / sec / 1e9
            if "l2_bytes" in t:
                out["l2_gbs"] = t["l2_bytes"] / sec / 1e9
            if "l1_bytes" in t:
                out["l1_gbs"] = t["l1_bytes"] / sec / 1e9
        if "issue_slot_pct" in t and "threads_per_inst" in t:
            out["issue_x_simt"] = t["issue_slot_pct"] / 100.0 * t["threads_per_inst"] / 32.0
        return out
    except (OSError, KeyError, ValueError):
        return None


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clocks / throttle reasons during the timed region (B200_PROFILING.md recipe) through NVML
    (same counters as `nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.*`), every 5 ms."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self._halt = index, [], threading.Event()
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        while not self._halt.is_set():
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((float(mhz), int(reasons)))
            except Exception:
                pass
            self._halt.wait(0.005)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        sm = [s[0] for s in self.samples]
        bits = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}
        reasons = sorted({n for _, r in self.samples for n, b in bits.items() if r & b})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(self.samples)}


def oracle_throughput(pkg, tables, rays, budget_rays):
    """Times the oracle on a bounded sample of the workload; returns (Mrays/s, cores, sample text)."""
    oracle = load_oracle()
    osc = oracle.OracleScene(tables)
    sample = oracle.OracleScene.normalized(rays[:budget_rays])
    cores = os.cpu_count() or 1
    osc.trace(sample[: min(65536, sample.size)], threads=cores, normalize=False)  # warm-up
    t0 = time.perf_counter()
    osc.trace(sample, threads=cores, normalize=False)
    dt = time.perf_counter() - t0
    return sample.size / dt / 1e6, cores, f"first {sample.size} rays of the c3-instanced batch, closest-hit"


PT_WORKLOAD = "c4-diamond"  # BASELINE.json configs[3]: scenes/diamond_scene.json, 1920x1080 (fixture tests/golden/diamond.npz)


class _DevArray:
    """Exposes a raw device pointer to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}


PT_FIXTURE = {"c4-diamond": "diamond", "c5-many-point-lights": "many-point-lights", "office-fisheye": "office"}
# The reference's only published timing: `igcli --spp 16384` on scenes/radiance/office/office_scene.json
# (256 x 256 fisheye, depth 8) took 13.516 s on a GeForce RTX 3080 -- WHOLE igcli wall clock, scene load + JIT + save included.
OFFICE_PUBLISHED = {"value": 256 * 256 * 16384 / 13.516 / 1e6, "unit": "Msamples/s", "hardware": "GeForce RTX 3080 12 GB",
                    "source": "scenes/radiance/office/benchmark.txt:17", "includes": "whole igcli wall clock: scene load + shader JIT + render + save"}


def load_pt_workload(pkg, workload=None):
    fx = importlib.import_module("ignis-masterthesis_b200.fixtures")
    tables, sh, w, h, _ = fx.load_fixture(os.path.join(ROOT, "tests", "golden", PT_FIXTURE[workload or PT_WORKLOAD] + ".npz"))
    spi = max(1, min(64, int(np.ceil(8 / ((w / 1000.0) * (h / 1000.0))))))  # recommendSPI (GPU), Runtime.cpp:61-69
    return tables, sh, w, h, spi


_COMM = {}


def product_comm(pkg, dev, rank, world):
    """The product's per-process communicator (igb200_comm_*: NCCL next to the device handle).  torch.distributed is only
    the launcher-side plumbing that hands rank 0's 128-byte id to the other ranks."""
    import torch
    import torch.distributed as dist
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid = torch.frombuffer(bytearray(pkg.Comm.unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(uid, src=0)
    return pkg.Comm(dev, bytes(uid.cpu().numpy().tobytes()), world, rank)


def bench_path_tracer(pkg, local_rank, rank, world, total_iterations, warmup, workload=None, mode=None):
    """Path-traced samples/s (igcli's definition: spi * w * h * iterations / seconds, src/frontend/cli/main.cpp:135) of ONE
    request of `total_iterations` iterations over all ranks, through the product's multi-GPU entries: igb200_comm_render
    (the request partitioned by iteration index or by tile rows inside the library, scene replicated) and
    igb200_comm_reduce_film (one ncclReduce of the film onto rank 0 over NVLink, communicator warmed beforehand).
    `value` stops after the reduce (film resident on rank 0's device); `e2e_value` adds the D2H read of the presented
    film through igb200_framebuffer into the device's pinned mirror (no further host copy)."""
    import torch
    import torch.distributed as dist
    workload = workload or PT_WORKLOAD
    tables, sh, w, h, spi = load_pt_workload(pkg, workload)
    dev = pkg.Device(local_rank)
    t_build = time.perf_counter()
    dev.assign_scene(tables)
    build_s = time.perf_counter() - t_build
    dev.assign_shading(sh)
    dev.resize(w, h)
    comm = product_comm(pkg, dev, rank, world) if world > 1 else None
    mode = pkg.device.PARTITION_AUTO if mode is None else mode
    # warm-up in the shape of the timed call: enough iterations per call for the library to size its wavefront streams as the
    # timed request will (they grow on demand up to 32 Mi rays; growing re-allocates and re-captures the round graph)
    share = max(1, total_iterations // world)
    warm_iters = max(1, min(share, -(-(1 << 25) // (w * h * spi))))
    for it in range(warmup):
        dev.render(w, h, spi=spi, iteration=1_000_000 + it * warm_iters, iterations=warm_iters)
    if comm is not None:  # warm the communicator on a film-sized buffer outside the timed region (first-use cost)
        comm.reduce_film(warmup, root=0, mode=pkg.device.PARTITION_ITERATIONS, warm=True)
    dev.clear_framebuffer()
    dev.framebuffer(copy=False)  # pinned mirror allocated and touched before timing
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    if comm is not None:
        comm.render(w, h, spi, total_iterations, mode=mode)
    else:
        dev.render(w, h, spi=spi, iteration=0, iterations=total_iterations)
    st = dev.render_stats()
    ms_iter = st["device_ms"]
    rays = st["camera_rays"] + st["bounce_rays"] + st["shadow_rays"]
    launches, rounds = st["kernel_launches"], st["rounds"]
    reduce_ms = comm.reduce_film(total_iterations, root=0, mode=mode, warm=False) if comm is not None else 0.0
    total_s = time.perf_counter() - t0
    host_film, n_total = dev.framebuffer(copy=False)  # D2H of the presented frame into the pinned mirror: the e2e read
    e2e_s = time.perf_counter() - t0
    film_mean = float(host_film.mean() / max(1, n_total)) if rank == 0 else 0.0
    stats = torch.tensor([total_s, e2e_s, float(rays), float(ms_iter), float(launches), float(reduce_ms)], device="cuda", dtype=torch.float64)
    reduce_ms_min = reduce_ms
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        total_s, e2e_s, rays, launches, reduce_ms, ms_iter = float(mx[0]), float(mx[1]), float(sm[2]), float(sm[4]), float(mx[5]), float(mx[3])
        mn = stats.clone()  # the rank that arrives last does not wait for anybody: its time is the collective itself
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        reduce_ms_min = float(mn[5])
    samples = float(spi) * w * h * total_iterations
    first, count, tr, tc = pkg.partition_share(total_iterations, rank, world, mode)
    out = {
        "workload": workload, "width": w, "height": h, "spi": spi, "iterations": int(total_iterations),
        "max_depth": int(sh.technique.max_depth), "metric": "path-traced samples/s", "unit": "Msamples/s",
        "value": samples / total_s / 1e6, "e2e_value": samples / e2e_s / 1e6,
        "device_ms_max_rank": ms_iter, "mrays_per_s": rays / total_s / 1e6,
        "rays_per_sample": rays / samples, "film_reduce_ms": reduce_ms, "film_reduce_ms_min_rank": reduce_ms_min,
        "film_reduce_wire_ms_at_450GBs": (w * h * 3 * 4) / 450e9 * 1e3 if world > 1 else 0.0,
        "d2h_bytes": w * h * 3 * 4, "partition": "tile rows" if tc > 1 else "iteration blocks",
        "api": "igb200_comm_render + igb200_comm_reduce_film" if world > 1 else "igb200_render_iterations",
        "gpu_launches": int(launches), "rounds_rank0": int(rounds), "iterations_on_root_film": int(n_total), "film_mean": film_mean,
        "bvh_build_s": build_s,
    }
    if world == 1:
        # the reference's frontends step ONE iteration per Device::render call (Runtime.cpp:294-330): the same samples/s metric
        # stepping through igb200_render_pipelined (wall clock, the final igb200_flush and the film read inside), next to plain
        # igb200_render calls, each of which drains its deep paths before it returns
        steps = int(max(1, min(total_iterations, 40)))
        dev.clear_framebuffer()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for it in range(steps):
            dev.render(w, h, spi=spi, iteration=it, pipelined=True)
        dev.flush()
        dev.framebuffer(copy=False)
        piped_s = time.perf_counter() - t0
        plain_steps = int(min(steps, 8))
        t0 = time.perf_counter()
        for it in range(plain_steps):
            dev.render(w, h, spi=spi, iteration=steps + it)
        plain_s = time.perf_counter() - t0
        out["per_step"] = {"api": "igb200_render_pipelined + igb200_flush + igb200_framebuffer", "steps": steps,
                           "value": float(spi) * w * h * steps / piped_s / 1e6,
                           "plain_igb200_render": float(spi) * w * h * plain_steps / plain_s / 1e6, "unit": "Msamples/s"}
    if comm is not None:
        comm.close()
    dev.close()
    return out


def oracle_samples_per_s(pkg, budget_s=12.0, workload=None):
    """The oracle's render iteration on the host cores on a bounded sample of the same workload (a crop-free
    low-resolution render of the same scene: same rays per sample)."""
    oracle = load_oracle()
    workload = workload or PT_WORKLOAD
    tables, sh, w, h, spi = load_pt_workload(pkg, workload)
    osc = oracle.OracleScene(tables)
    cw, ch = w // 4, h // 4
    cores = os.cpu_count() or 1
    osc.render(sh, cw, ch, spi=1, iterations=1)
    t0 = time.perf_counter()
    n = 0
    while time.perf_counter() - t0 < budget_s:
        osc.render(sh, cw, ch, spi=spi, iterations=1, first_iteration=n)
        n += 1
    dt = time.perf_counter() - t0
    return {"value": cw * ch * spi * n / dt / 1e6, "unit": "Msamples/s", "cores": cores, "kind": "port",
            "sample": f"{n} iterations of {workload} at {cw}x{ch} (1/16 of the pixels), spi {spi}"}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU traversal (oracle restatement: the reference itself needs
    AnyDSL/Artic and cannot be built here) on all host cores, bounded sample per step."""
    if rank != 0:
        return
    pkg = load_pkg()
    tables = pkg.workloads.c3_scene()
    n = args.rays  # the SAME batch as the GPU arm (16 777 216 rays, seed 0xB200): ~0.3 s per step on 16 cores
    rays = pkg.workloads.c3_rays(tables, n=n)
    oracle = load_oracle()
    osc = oracle.OracleScene(tables)
    sample = oracle.OracleScene.normalized(rays)
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        osc.trace(sample, threads=cores, normalize=False)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        osc.trace(sample, threads=cores, normalize=False)
    dt = time.perf_counter() - t0
    val = n * args.steps / dt / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "c3-instanced", "triangles": 1032192, "entities": 64, "rays_per_step": n,
                   "ray_order": "generation order (incoherent)"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"all {n} rays of the c3-instanced batch per step"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if args.pt_iterations > 0:
        pt = oracle_samples_per_s(pkg)
        line["path_tracer"] = {"workload": PT_WORKLOAD, "metric": "path-traced samples/s", "unit": "Msamples/s",
                               "value": pt["value"], "e2e_value": pt["value"], "cpu_baseline": pt}
    if args.c5_spp > 0:
        c5 = oracle_samples_per_s(pkg, 8.0, "c5-many-point-lights")
        line["path_tracer_c5"] = {"workload": "c5-many-point-lights", "metric": "path-traced samples/s", "unit": "Msamples/s",
                                  "value": c5["value"], "e2e_value": c5["value"], "cpu_baseline": c5}
    emit(json.dumps(line))


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pkg = load_pkg()
    W = pkg.workloads
    tables = W.c3_scene()
    n = args.rays
    # weak scaling: every rank traces its own batch (distinct seed), scene replicated, no data-path collective
    rays = W.c3_rays(tables, n=n, seed=0xB200 + rank)

    dev = pkg.Device(local_rank)
    t_build = time.perf_counter()
    dev.assign_scene(tables)
    info = dev.bvh_info()
    info["build_s"] = time.perf_counter() - t_build  # host BVH8 build + upload (igb200_assign_scene), one host thread

    d_rays = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).cuda()
    d_ent = torch.empty(n, dtype=torch.int32, device="cuda")
    d_prim = torch.empty(n, dtype=torch.int32, device="cuda")
    d_t, d_u, d_v = (torch.empty(n, dtype=torch.float32, device="cuda") for _ in range(3))
    stream = torch.cuda.current_stream()

    def step():
        dev.trace_device(d_rays.data_ptr(), n, d_ent.data_ptr(), d_prim.data_ptr(), d_t.data_ptr(), d_u.data_ptr(),
                         d_v.data_ptr(), stream=stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin.record(stream)
    for a, b in evs:
        a.record(stream)
        step()
        b.record(stream)
    t_end.record(stream)
    barrier()
    clocks = sampler.stop() if sampler else None
    total_ms = t_begin.elapsed_time(t_end)
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in evs]))
    if world > 1:
        tm = torch.tensor([total_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        total_ms = float(tm.item())
    value = world * n * args.steps / (total_ms * 1e-3) / 1e6
    hits = int((d_prim >= 0).sum().item())

    # ---- BASELINE.json configs[2] also names any-hit on the same batch, and SURVEY 8(d) a flattened variant of the
    # same triangles (one 1 M-triangle BVH, ~75 MB of nodes + triangles: L2- instead of L1-resident) ----
    def timed_variant(device, any_hit, steps, src=None):
        src = d_rays if src is None else src
        def one():
            device.trace_device(src.data_ptr(), n, d_ent.data_ptr(), d_prim.data_ptr(), d_t.data_ptr(), d_u.data_ptr(),
                                d_v.data_ptr(), any_hit=any_hit, stream=stream.cuda_stream)
        for _ in range(3):
            one()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(steps):
            one()
        b.record(stream)
        barrier()
        ms = a.elapsed_time(b) / steps
        if world > 1:
            tm = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            ms = float(tm.item())
        return {"value": world * n / (ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms, "steps": steps,
                "hit_rate": float((d_prim >= 0).sum().item()) / n}

    variants = None
    if args.variant_steps > 0:
        variants = {"c3-instanced any-hit": timed_variant(dev, True, args.variant_steps)}
        fdev = pkg.Device(local_rank)
        ftables = W.c3_scene(flat=True)
        t_build = time.perf_counter()
        fdev.assign_scene(ftables)
        flat_build_s = time.perf_counter() - t_build
        variants["c3-flat closest-hit"] = timed_variant(fdev, False, args.variant_steps)
        variants["c3-flat any-hit"] = timed_variant(fdev, True, args.variant_steps)
        variants["c3-flat bvh"] = dict(fdev.bvh_info(), build_s=flat_build_s)
        # c3-dense: the same two scenes under rays that (nearly) all hit -- aimed at a random sphere each, so they cross
        # the grid and are occluded on the way (workloads.c3_dense_rays): deep traversal instead of C3's 18 % hit rate.
        dense = W.c3_dense_rays(tables, n=n, seed=0xDE5E + rank)
        d_dense = torch.from_numpy(dense.view(np.float32).reshape(-1, 8)).cuda()
        peak_v, _ = measured_peak()
        for name, device, scene_tables in (("c3-dense", dev, tables), ("c3-dense-flat", fdev, ftables)):
            for any_hit in (False, True):
                r = timed_variant(device, any_hit, args.variant_steps, d_dense)
                if rank == 0:  # visit counts of the same kernel (STATS instantiation) on a 2 Mi-ray sample, untimed
                    sdev = pkg.Device(local_rank, acquire_stats=True)
                    sdev.assign_scene(scene_tables)
                    m = min(n, 1 << 21)
                    sdev.trace(dense[:m], any_hit=any_hit)
                    st = sdev.trace_stats()
                    sdev.close()
                    pr = {"nodes": st["nodes_visited"] / m, "tris": st["tris_tested"] / m, "entities": st["entities_entered"] / m}
                    bpr = 32 + (4 if any_hit else 20) + 80 * pr["nodes"] + 48 * pr["tris"] + 96 * pr["entities"]
                    ach = bpr * n / (r["ms_per_step"] * 1e-3) / 1e9
                    r.update(per_ray=pr, bytes_per_ray=bpr, roofline={"bound": "hbm", "achieved": ach, "peak": peak_v, "unit": "GB/s",
                                                                      "frac": ach / peak_v})
                variants[name + (" any-hit" if any_hit else " closest-hit")] = r
        del d_dense, dense
        fdev.close()
        del ftables

    # ---- end to end through the host-buffer C-ABI call (pinned host memory in, pinned host memory out) ----
    h_rays = torch.from_numpy(rays.view(np.float32).reshape(-1, 8)).pin_memory()
    h_i = [torch.empty(n, dtype=torch.int32).pin_memory() for _ in range(2)]
    h_f = [torch.empty(n, dtype=torch.float32).pin_memory() for _ in range(3)]
    outs = [x.numpy() for x in h_i + h_f]
    h_rays_np = h_rays.numpy().view(pkg.tables.RAY_DTYPE).reshape(-1)
    e2e_steps = max(1, min(args.steps, 5))
    for _ in range(2):
        dev.trace_into(h_rays_np, *outs)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        dev.trace_into(h_rays_np, *outs)
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        tm = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        e2e_s = float(tm.item())
    e2e_val = world * n * e2e_steps / e2e_s / 1e6
    assert int((outs[1] >= 0).sum()) > 0
    # the same call asking for the columns igtrace's debug modes 15 / 17 / 13 need (ids + t): 12 B instead of 20 B per ray back
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        dev.trace_into(h_rays_np, outs[0], outs[1], outs[2], None, None)
    barrier()
    e2e_ids_s = time.perf_counter() - t0
    if world > 1:
        tm = torch.tensor([e2e_ids_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        e2e_ids_s = float(tm.item())
    e2e_ids_val = world * n * e2e_steps / e2e_ids_s / 1e6
    dev.close()
    del d_rays, d_ent, d_prim, d_t, d_u, d_v

    # ---- second headline metric: path-traced samples/s (all ranks take part) ----
    # (weak scaling: the request grows with the number of ranks, pt_iterations per GPU)
    pt = bench_path_tracer(pkg, local_rank, rank, world, args.pt_iterations * world, 2) if args.pt_iterations > 0 else None
    if pt is not None:
        pt["scaling"] = "weak"

    # ---- BASELINE.json configs[4]: many_point_lights at 3840x2160, a FIXED number of samples per pixel split across
    # the ranks inside the library (strong scaling: igb200_partition_share, iteration blocks; tile rows when spp < ranks * spi)
    c5 = None
    if args.c5_spp > 0:
        _, _, _, _, c5_spi = load_pt_workload(pkg, "c5-many-point-lights")
        total_it = max(1, args.c5_spp // c5_spi)
        c5 = bench_path_tracer(pkg, local_rank, rank, world, total_it, 2, "c5-many-point-lights")
        c5["scaling"] = "strong"
        c5["spp"] = int(c5["iterations"] * c5["spi"])

    # ---- the reference's own published benchmark scene, at its published sample count (N = 1 only: 1 G samples last ~1 s)
    office = None
    if args.office_spp > 0 and world == 1:
        _, _, _, _, o_spi = load_pt_workload(pkg, "office-fisheye")
        office = bench_path_tracer(pkg, local_rank, rank, world, max(1, args.office_spp // o_spi), 2, "office-fisheye")
        office["spp"] = int(office["iterations"] * office["spi"])
        office["published"] = OFFICE_PUBLISHED
        office["vs_published"] = office["e2e_value"] / OFFICE_PUBLISHED["value"]

    if rank == 0:
        # ---- algorithmic bytes per ray: visit counts from a counting pass over a 2 Mi-ray sample (untimed) ----
        sdev = pkg.Device(local_rank, acquire_stats=True)
        sdev.assign_scene(tables)
        m = min(n, 1 << 21)
        sdev.trace(rays[:m])
        st = sdev.trace_stats()
        sdev.close()
        n_node, n_tri, n_ent = st["nodes_visited"] / m, st["tris_tested"] / m, st["entities_entered"] / m
        bytes_per_ray = 32 + 20 + 80 * n_node + 48 * n_tri + 96 * n_ent
        peak, peak_kind = measured_peak()
        achieved = bytes_per_ray * n / (kernel_ms * 1e-3) / 1e9
        # ---- CPU baseline: oracle on a bounded sample ----
        # (rank 0 at N=1 only: at N>1 the host cores are shared with the other ranks' launch threads)
        cpu = None
        if world == 1:
            cpu_val, cores, sample = oracle_throughput(pkg, tables, rays, min(n, args.cpu_rays))
            cpu = {"value": cpu_val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "c3-instanced", "triangles": 1032192, "entities": 64, "rays_per_step": n,
                       "mesh": "procedural unit UV sphere, 64 x 128: the surface of scenes/meshes/SphereTri.ply (16 128 triangles, the same "
                               "8066 vertex positions to 1e-3, tests/test_primitives.py; generated because the reference tree is absent on the GPU box)",
                       "ray_order": "generation order (incoherent)", "hit_rate": hits / n,
                       "l2": "inputs+outputs per step (%.0f MB) exceed the 126 MB L2" % (n * 52 / 1e6),
                       "bvh": info},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": n * 32, "d2h_bytes_per_step": n * 20,
                    "steps": e2e_steps,
                    "ids_and_t_only": {"value": e2e_ids_val, "unit": UNIT, "d2h_bytes_per_step": n * 12}},
            "gpu_launches": args.steps,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": measured_traffic(n), "peak_kind": peak_kind, "kernel": "trace_kernel<closest>",
                         "kernel_ms": kernel_ms, "bytes_per_ray": bytes_per_ray,
                         "per_ray": {"nodes": n_node, "tris": n_tri, "entities": n_ent},
                         "dram_frac": (measured_traffic(n) / (kernel_ms * 1e-3) / 1e9 / peak) if measured_traffic(n) else None,
                         "ncu": ncu_counters()},
            "cpu_baseline": cpu,
        }
        if variants is not None:
            line["trace_variants"] = variants
            line["gpu_launches"] += 7 * (args.variant_steps + 3) + 4
        if pt is not None:
            pt["cpu_baseline"] = oracle_samples_per_s(pkg) if world == 1 else None
            line["path_tracer"] = pt
            line["gpu_launches"] += pt["gpu_launches"]
        if office is not None:
            line["path_tracer_office"] = office
            line["gpu_launches"] += office["gpu_launches"]
        if c5 is not None:
            c5["cpu_baseline"] = oracle_samples_per_s(pkg, 8.0, "c5-many-point-lights") if world == 1 else None
            line["path_tracer_c5"] = c5
            line["gpu_launches"] += c5["gpu_launches"]
        emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def reserve_stdout():
    """stdout carries exactly ONE JSON line: libraries that chat on fd 1 (NCCL prints its version banner there when the
    first communicator is created) are sent to stderr for the whole run; emit() writes to the real stdout."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(line + "\n")
    out.flush()


def main():
    reserve_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--rays", type=int, default=FULL_RAYS)
    ap.add_argument("--cpu-rays", type=int, default=4_000_000)
    ap.add_argument("--pt-iterations", type=int, default=64,
                    help="path-traced iterations per GPU; 64 x spi 4 = the 256 spp of BASELINE.json configs[3] (0 disables the leg)")
    ap.add_argument("--c5-spp", type=int, default=1024,
                    help="samples per pixel of the many_point_lights leg (BASELINE.json configs[4]: 1024), split across ranks; 0 disables")
    ap.add_argument("--office-spp", type=int, default=16384,
                    help="samples per pixel of the office_scene.json leg (the reference's published run: 16384); 0 disables")
    ap.add_argument("--variant-steps", type=int, default=10,
                    help="timed steps of the any-hit and flattened-mesh variants of the C3 batch (0 disables)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
