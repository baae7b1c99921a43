#!/usr/bin/env python
"""bench.py — the render-loop hot path (renderthread -> color -> hit / scatter / pdf,
Raytracing_n.cpp:55-106, :815-879) on N B200s, next to the reference's own CPU renderer.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config spheres|cornell] [--impl reference]

A *step* is one render of the whole configured image (every pixel, every Sobol index) followed by
the reference's resolve (mean -> sqrt -> 8-bit, Raytracing_n.cpp:848-862).  `value` is paths/s
(path = one color() call from the sample loop, :841) with the scene already resident in HBM; `e2e`
is the same metric through the C-ABI call that takes HOST buffers (srt_scene_create from the host
scene description + srt_render_image into host memory), copies inside the timed region.  With N > 1
(torchrun) the Sobol index range [0, ns) is split across ranks and the float accumulation buffers
are summed with ONE NCCL reduce.

`--impl reference` times the UNMODIFIED reference renderer (oracle/_ref/libsrt_ref.so: the
reference's renderthread pool on all host threads) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SEED = 0x5EED
L2_FLUSH_BYTES = 256 << 20  # > the 126 MB L2

# nominal FP32 peak used only for the informational fp32 fraction: 148 SMs x 128 lanes x 2 flop
FP32_LANES = 148 * 128 * 2


def config_of(name):
    from simple_raytracing_render_b200 import scenes
    nx, ny, ns = scenes.CONFIG_SIZES[name]
    label = {"cornell": "C1 Cornell box 512x512 @ 16 spp",
             "spheres": "C2 random spheres (bvh_node) 1200x800 @ 64 spp",
             "teapot": "C3 Utah teapot 640k triangles (bvh_node), Beckmann, sky sphere 1024x1024 @ 256 spp",
             "soldier": "C4-like textured Beckmann mesh + env map + MERL-material floor + glass 1920x1080 @ 1024 spp",
             "tessellated": f"C5 tessellated terrain (2 x {int(os.environ.get('SRT_TESS_NSIDE', '512'))}^2 triangles) "
                            "+ constant_medium 3840x2160 @ 256 spp"}[name]
    return nx, ny, ns, label


def cpu_sample(nx, ny, ns, target_paths=1_100_000):
    """Bounded sample of a config for the CPU legs: the same scene and camera at 1/f resolution per
    axis and min(ns, 16) samples, f the smallest integer that brings it under ~1.1 M paths."""
    ss = min(ns, 16)
    f = 1
    while (nx // f) * (ny // f) * ss > target_paths:
        f += 1
    return nx // f, ny // f, ss


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="srt_clocks_", suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


class stdout_to_stderr:
    """The reference's own code prints to the C stdout (brdf.h:162); the bench line must be the only thing
    on stdout, so file descriptor 1 points at stderr while the reference runs."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        try:
            import ctypes
            ctypes.CDLL(None).fflush(None)
        except OSError:
            pass
        os.dup2(self.saved, 1)
        os.close(self.saved)
        return False


def algorithmic_work(counters, n_pixels, prim_bytes):
    """bytes and flops of one launch from the kernel's own counters."""
    rays, nodes, prims = counters["rays"], counters["node_visits"], counters["prim_tests"]
    state = 144 * rays + 24 * n_pixels            # ray + hit + path state, accumulation: moves through L2 / HBM
    geometry = 128 * nodes + prim_bytes * prims   # node and primitive fetches: L1 / shared-memory hits for C1-C4
    fl = 12 * rays + 18 * 4 * nodes + 27 * prims + 130 * max(rays - counters["paths"], 0)
    return state + geometry, fl, state


KERNEL_OF_MODE = {0: "k_render", 1: "k_wf_trace + k_wf_shade", 2: "k_wf_persist", 3: "k_wf_sorted"}


def measured_traffic(config, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, from the committed
    ncu --set full capture of this configuration (profiles/traffic.json, written by tools/ncu_traffic.py)."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as f:
            rec = json.load(f).get(config)
    except (OSError, ValueError):
        return None
    if not rec or rec.get("kernel") != kernel:
        return None
    return rec.get("dram_bytes_per_launch")


def run_reference(args):
    """The reference's own renderer on the host cores: bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import ref
    from simple_raytracing_render_b200 import scenes
    nx, ny, ns, label = config_of(args.config)
    if not ref.available():
        emit({"impl": "reference", "unavailable": "oracle/_ref/libsrt_ref.so was not built"})
        return 0
    cores = os.cpu_count() or 1
    sx, sy, ss = cpu_sample(nx, ny, ns)
    paths = sx * sy * ss
    with stdout_to_stderr():
        rs = ref.RefBuilder().realize(*scenes.by_name(args.config))
        for _ in range(max(args.warmup, 1)):
            rs.render_timed(sx, sy, ss, cores)
        times = [rs.render_timed(sx, sy, ss, cores) for _ in range(args.steps)]
        _, nrays = rs.render_float(max(sx // 4, 8), max(sy // 4, 8), ss, mod_ny=sx == sy)
    rpp = nrays / float(max(sx // 4, 8) * max(sy // 4, 8) * ss)
    sec = float(np.mean(times))
    v = paths / sec
    sample = f"{sx}x{sy} @ {ss} spp of the same scene/camera ({paths} paths per step), {cores} std::threads"
    emit({
        "impl": "reference", "metric": "paths/s", "value": v, "unit": "paths/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": label, "sample": sample},
        "mrays_per_s": v * rpp / 1e6, "rays_per_path": rpp,
        "cpu_baseline": {"value": v, "unit": "paths/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": "paths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })
    return 0


REAL_STDOUT = None


def emit(line):
    """The ONE JSON line, on the process's original stdout (see main)."""
    data = (json.dumps(line) + "\n").encode()
    if REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(REAL_STDOUT, data)


def main():
    # Libraries print to the C stdout behind Python's back (NCCL its version banner at communicator
    # creation, the reference's brdf.h a progress line): file descriptor 1 points at stderr for the whole run
    # and the JSON line goes to a private copy of the original stdout.
    global REAL_STDOUT
    sys.stdout.flush()
    REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="spheres", choices=["spheres", "cornell", "teapot", "soldier", "tessellated"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from simple_raytracing_render_b200 import ffi, multigpu, render, scenes, scene as S

    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this backend has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=dev)

    nx, ny, ns, label = config_of(args.config)
    n_pixels = nx * ny
    flat = S.flatten(*scenes.by_name(args.config))
    stream = torch.cuda.Stream(device=dev)
    ctx = render.Context(local_rank, stream.cuda_stream)
    sc = render.Scene(ctx, flat)
    info = sc.info()

    # sample-range shard of this rank: Sobol indices [s0, s1) of every pixel
    s0, s1 = multigpu.sample_shard(ns, rank, world_size)
    params = render.render_params(nx, ny, ns, s0, s1, seed=SEED)
    accum = torch.zeros(n_pixels * 3, dtype=torch.float32, device=dev)
    rgb8 = torch.zeros(n_pixels * 3, dtype=torch.uint8, device=dev)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)

    def step():
        accum.zero_()
        sc.render(params, accum)
        multigpu.reduce_accum(accum, 0)
        if rank == 0:
            ctx.resolve(accum, n_pixels, ns, rgb8)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world_size > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
        sync_all()
        launches0 = ctx.counters()["kernel_launches"]
        clocks = ClockSampler(local_rank)
        if rank == 0:
            clocks.start()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for k in range(args.steps):
            flush.fill_(k & 0xFF)  # L2 flush between timed iterations (untimed)
            if world_size > 1:
                dist.barrier()
            ev[k][0].record(stream)
            accum.zero_()
            kev[k][0].record(stream)
            sc.render(params, accum)
            kev[k][1].record(stream)
            multigpu.reduce_accum(accum, 0)
            if rank == 0:
                ctx.resolve(accum, n_pixels, ns, rgb8)
            ev[k][1].record(stream)
        sync_all()
        clk = clocks.stop() if rank == 0 else None
        launches = ctx.counters()["kernel_launches"] - launches0
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    kernel_ms = sum(a.elapsed_time(b) for a, b in kev) / args.steps
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    paths_per_step = n_pixels * ns
    value = paths_per_step / (ms_per_step * 1e-3)

    # ---- untimed instrumented replay of the same launch: the kernel's own work counters ----------
    with torch.cuda.stream(stream):
        scratch = torch.zeros_like(accum)
        cnt = sc.render_counted(params, scratch)
    rpp = cnt["rays"] / max(cnt["paths"], 1)
    prim_bytes = {"spheres": 16, "cornell": 20, "teapot": 48, "soldier": 48, "tessellated": 48}[args.config]
    abytes, aflops, state_bytes = algorithmic_work(cnt, n_pixels, prim_bytes)
    kernel_name = KERNEL_OF_MODE.get(int(sc.info()["render_variant"]) // 100, "k_render")

    # ---- e2e: HOST buffers through the C ABI, copies inside the timed region ----------------------
    full = render.render_params(nx, ny, ns, s0, s1, seed=SEED)
    host_rgb = torch.empty(n_pixels * 3, dtype=torch.uint8).pin_memory()
    host_acc = torch.empty(n_pixels * 3, dtype=torch.float32).pin_memory() if world_size > 1 else None

    e2e_parts = {"scene_create_ms": 0.0, "render_image_ms": 0.0}

    def e2e_step():
        ta = time.perf_counter()
        sc2 = render.Scene(ctx, flat)  # host scene description -> device (h2d)
        tb = time.perf_counter()
        e2e_parts["scene_create_ms"] += (tb - ta) * 1e3
        if world_size > 1:
            sc2.render_image(full, rgb8=host_rgb.numpy(), accum=host_acc.numpy(), want_accum=True)
            a = host_acc.to(dev, non_blocking=True)
            dist.reduce(a, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                ctx.resolve(a, n_pixels, ns, host_rgb.numpy())
        else:
            sc2.render_image(full, rgb8=host_rgb.numpy())
        e2e_parts["render_image_ms"] += (time.perf_counter() - tb) * 1e3
        sc2.close()

    with torch.cuda.stream(stream):
        for _ in range(2):
            e2e_step()
        sync_all()
        e2e_steps = max(3, min(args.steps, 10))
        e2e_parts = {k: 0.0 for k in e2e_parts}
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        sync_all()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world_size > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())

    if rank == 0:
        hbm_peak, which, sm_max = peaks()
        fp32_meas = ctx.measure_fp32_peak()
        k_s = kernel_ms * 1e-3
        achieved = abytes / k_s / 1e9
        line = {
            "metric": "paths/s", "value": value, "unit": "paths/s", "n_gpus": world_size, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": label, "nx": nx, "ny": ny, "spp": ns, "max_depth": 50,
                       "parallelism": f"sample-range x{world_size}", "l2": "flushed between timed steps (256 MiB write)",
                       "scene_device_bytes": int(info["device_bytes"]), "bvh_nodes": int(info["bvh_nodes"])},
            "mrays_per_s": value * rpp / 1e6, "rays_per_path": rpp,
            "kernel_ms": kernel_ms, "kernel_share_of_step": kernel_ms / ms_per_step,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                         "frac": achieved / hbm_peak, "traffic": measured_traffic(args.config, kernel_name),
                         "peak_source": which, "kernel": kernel_name, "algorithmic_bytes_per_launch": int(abytes),
                         "state_bytes_per_launch": int(state_bytes), "state_gbps": state_bytes / k_s / 1e9,
                         "note": "algorithmic bytes (SURVEY 8d) = path state + node / primitive fetches; the "
                                 "fetches hit L1 / shared memory (scene far below L2), so `achieved` is not DRAM "
                                 "traffic and may exceed `peak`; `traffic` is the measured DRAM bytes. The binding "
                                 "limits are instruction issue and latency: see fp32 and DESIGN.md §8"},
            "fp32": {"achieved_tflops": aflops / k_s / 1e12,
                     "nominal_peak_tflops": FP32_LANES * sm_max * 1e6 / 1e12,
                     "measured_peak_tflops": fp32_meas,
                     "frac": (aflops / k_s / 1e12) / fp32_meas if fp32_meas > 0 else None,
                     "frac_of": "measured FMA-chain peak (srt_measure_fp32_peak)",
                     "algorithmic_flops_per_launch": int(aflops)},
            "counters": {k: cnt[k] for k in ("paths", "rays", "culled_rays", "probe_rays", "node_visits", "prim_tests", "reject_iters")},
            "e2e": {"value": paths_per_step / e2e_s, "unit": "paths/s",
                    "h2d_bytes_per_step": int(info["device_bytes"]),
                    "d2h_bytes_per_step": int(n_pixels * 3 + (n_pixels * 12 if world_size > 1 else 0)),
                    "ms_per_step": e2e_s * 1e3,
                    "parts_ms": {k: v / e2e_steps for k, v in e2e_parts.items()}},
            "gpu_launches": int(launches),
            "clocks": clk,
        }
        if world_size == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args.config)
        emit(line)
    sc.close()
    ctx.close()
    if world_size > 1:
        dist.destroy_process_group()
    return 0


def cpu_baseline(config):
    """oracle/_ref (the unmodified reference renderer) on the host cores, bounded sample."""
    from oracle import ref
    from simple_raytracing_render_b200 import scenes
    if not ref.available():
        return {"value": None, "unit": "paths/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    nx, ny, ns, _ = config_of(config)
    sx, sy, ss = cpu_sample(nx, ny, ns)
    cores = os.cpu_count() or 1
    with stdout_to_stderr():
        rs = ref.RefBuilder().realize(*scenes.by_name(config))
        rs.render_timed(sx, sy, ss, cores)
        secs = [rs.render_timed(sx, sy, ss, cores) for _ in range(3)]
        one = rs.render_timed(sx, sy, ss, 1)
    paths = sx * sy * ss
    return {"value": paths / float(np.mean(secs)), "unit": "paths/s", "cores": cores, "kind": "reference",
            "sample": f"{sx}x{sy} @ {ss} spp of the same scene/camera ({paths} paths), reference renderthread pool",
            "single_thread_value": paths / one}


if __name__ == "__main__":
    sys.exit(main())
