#!/usr/bin/env python
"""bench.py — the render-loop hot path (renderthread -> color -> hit / scatter / pdf,
Raytracing_n.cpp:55-106, :815-879) on N B200s, next to the reference's own CPU renderer.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config spheres|cornell|...] [--impl reference]
                    [--no-per-config] [--no-cpu-baseline]

A *step* is one render of the whole configured image (every pixel, every Sobol index) followed by
the reference's resolve (mean -> sqrt -> 8-bit, Raytracing_n.cpp:848-862).  `value` is paths/s
(path = one color() call from the sample loop, :841) with the scene already resident in HBM; `e2e`
is the same metric through the C-ABI calls that take HOST buffers (srt_scene_create from the host
scene description, render, 8-bit image into pinned host memory), copies inside the timed region.
With N > 1 (torchrun) the Sobol index range [0, ns) is split across ranks and the int64 fixed-point
accumulation buffers are summed exactly, resolved per pixel tile and the 8-bit tiles distributed by ONE kernel
over NVLink peer memory (NCCL reduce-scatter / all-gather as the fallback).

The ONE JSON line describes the headline workload (BASELINE.json:configs[1], C2 random spheres) and
carries `per_config`: the same measurement (value, e2e, cpu_baseline, kernel, roofline, counters) for
every configuration of BASELINE.json — C1 Cornell (the configuration of the north-star target),
C3 teapot, C4 soldier-like, C5 tessellated at its test size and at the full 10 M triangles — with
a few steps each.  Under torchrun per_config holds the configurations BASELINE names for 8 GPUs
(C4 sample-sharded, C5 at 10 M triangles).

`--impl reference` times the UNMODIFIED reference renderer (oracle/_ref/libsrt_ref.so: the
reference's renderthread pool on all host threads) on a bounded sample of the same workload, plus
ONE full-size frame of the headline configuration (`full_frame`).
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
T_START = time.time()

# nominal FP32 peak (informational): 148 SMs x 128 lanes x 2 flop
FP32_LANES = 148 * 128 * 2

# name -> (scene name for scenes.by_name, kwargs, label, bytes per primitive test)
CONFIGS = {
    "cornell": ("cornell", {}, "C1 Cornell box 512x512 @ 16 spp", 20),
    "spheres": ("spheres", {}, "C2 random spheres (bvh_node) 1200x800 @ 64 spp", 16),
    "teapot": ("teapot", {}, "C3 Utah teapot 640k triangles (bvh_node), Beckmann, sky_2.png sky sphere 1024x1024 @ 256 spp", 48),
    "soldier": ("soldier", {}, "C4 soldier_scene: Soilder.FBX mesh[0] (3 971 triangles, Beckmann + albedo image) + sky4.jpg env "
                               "map + MERL-material floor (synthetic table) + glass 1920x1080 @ 1024 spp", 48),
    "tessellated": ("tessellated", {"n_side": 512}, "C5-like tessellated terrain (524 288 triangles) + constant_medium "
                                                    "3840x2160 @ 256 spp", 48),
    "tessellated_10m": ("tessellated", {"n_side": 2237}, "C5 tessellated terrain (10 008 338 triangles) + "
                                                         "constant_medium 3840x2160 @ 256 spp", 48),
}
# the roof SURVEY.md 8d names per configuration: C1-C4 are L2-resident (FP32 issue + L2); C5 at 10 M triangles
# (2.8 GB on the device) is HBM-resident
HBM_BOUND = {"tessellated_10m"}


def config_of(name):
    from simple_raytracing_render_b200 import scenes
    scene_name, kw, label, prim_bytes = CONFIGS[name]
    kw = dict(kw)
    if name == "tessellated" and os.environ.get("SRT_TESS_NSIDE"):
        kw["n_side"] = int(os.environ["SRT_TESS_NSIDE"])
        label = f"C5 tessellated terrain (2 x {kw['n_side']}^2 triangles) + constant_medium 3840x2160 @ 256 spp"
    if name == "soldier" and not scenes.real_assets():
        label = ("C4-like textured Beckmann mesh + env map + MERL-material floor + glass 1920x1080 @ 1024 spp "
                 "(stand-in assets: assets_cache/ is empty)")
    if name == "teapot" and not scenes.real_assets():
        label += " (synthetic sky image: assets_cache/ is empty)"
    nx, ny, ns = scenes.CONFIG_SIZES[scene_name]
    return scene_name, kw, nx, ny, ns, label, prim_bytes


def build_scene(name):
    from simple_raytracing_render_b200 import scenes
    scene_name, kw, *_ = config_of(name)
    return scenes.by_name(scene_name, **kw)


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
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


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


REAL_STDOUT = None


def emit(line):
    """The ONE JSON line, on the process's original stdout (see main)."""
    data = (json.dumps(line) + "\n").encode()
    if REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(REAL_STDOUT, data)


def log(msg):
    sys.stderr.write(f"[bench {time.time() - T_START:7.1f}s] {msg}\n")
    sys.stderr.flush()


# ---- algorithmic work and rooflines -----------------------------------------------------
PRIM_FLOPS = {20: 6, 16: 27, 48: 46}  # rect / sphere / triangle test, keyed by the primitive's fetch size

KERNEL_OF_MODE = {0: "k_render", 1: "k_wf_trace + k_wf_shade", 2: "k_wf_persist", 3: "k_wf_sorted"}


def algorithmic_work(counters, n_pixels, prim_bytes):
    """SURVEY.md §8d per-unit figures x the launch's own counters (the instrumented build of the kernel that
    was timed): per ray 144 B of ray + hit + path state written and read once, 128 B per wide-node visit,
    the primitive's record per primitive test, 24 B per pixel of accumulation; flops per ray
    12 + 18 per box (4 per wide node) + 6 / 27 / 46 per rect / sphere / triangle test, 130 per scattering vertex."""
    rays, nodes, prims = counters["rays"], counters["node_visits"], counters["prim_tests"]
    state = 144 * rays + 24 * n_pixels
    geometry = 128 * nodes + prim_bytes * prims
    fl = 12 * rays + 18 * 4 * nodes + PRIM_FLOPS.get(prim_bytes, 27) * prims + 130 * max(rays - counters["paths"], 0)
    return state + geometry, fl, state


def measured_traffic(config, kernel, spp):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE single-GPU launch of the dominant kernel, from the
    committed ncu --set full capture of this configuration (profiles/traffic.json, tools/ncu_traffic.py).  A capture
    taken at fewer samples per pixel (`captured_spp`: short replays for the long frames) is scaled to the frame's
    spp: this kernel's DRAM traffic is proportional to the paths it runs."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path) as f:
            rec = json.load(f).get(config)
    except (OSError, ValueError):
        return None
    if not rec or rec.get("kernel") != kernel:
        return None
    b = rec.get("dram_bytes_per_launch")
    if b is not None and rec.get("captured_spp"):
        b = b * spp / float(rec["captured_spp"])
    return b


def make_roofline(name, cnt, n_pixels, prim_bytes, kernel_s, kernel_name, shard_frac, hw, spp):
    abytes, aflops, state_bytes = algorithmic_work(cnt, n_pixels, prim_bytes)
    traffic = measured_traffic(name, kernel_name, spp)
    if traffic is not None and shard_frac < 1.0:
        traffic = traffic * shard_frac  # a rank renders shard_frac of the frame's samples: traffic of ITS launch
    fp32 = {"achieved": aflops / kernel_s / 1e12, "peak": hw["fp32_tflops"], "unit": "TFLOP/s",
            "frac": (aflops / kernel_s / 1e12) / hw["fp32_tflops"] if hw["fp32_tflops"] else None,
            "peak_source": "measured FMA-chain peak (srt_measure_fp32_peak)",
            "nominal_peak": FP32_LANES * hw["sm_max_mhz"] * 1e6 / 1e12, "algorithmic_flops_per_launch": int(aflops)}
    l2 = {"achieved": state_bytes / kernel_s / 1e9, "peak": hw["l2_gbps"], "unit": "GB/s",
          "frac": (state_bytes / kernel_s / 1e9) / hw["l2_gbps"] if hw["l2_gbps"] else None,
          "peak_source": "measured L2-resident streaming read (srt_measure_l2_peak)",
          "bytes_per_launch": int(state_bytes),
          "what": "ray + hit + path state (144 B per ray) and the accumulation (24 B per pixel): the bytes that "
                  "move through L2 between the phases of the wavefront; node / primitive fetches are L1 hits"}
    hbm = {"achieved": abytes / kernel_s / 1e9, "peak": hw["hbm_gbps"], "unit": "GB/s",
           "frac": (abytes / kernel_s / 1e9) / hw["hbm_gbps"], "peak_source": hw["hbm_source"],
           "algorithmic_bytes_per_launch": int(abytes),
           "measured_dram_gbps": traffic / kernel_s / 1e9 if traffic is not None else None,
           "measured_dram_frac": traffic / kernel_s / 1e9 / hw["hbm_gbps"] if traffic is not None else None}
    if name in HBM_BOUND:
        top = {"bound": "hbm", "achieved": hbm["achieved"], "peak": hbm["peak"], "unit": "GB/s", "frac": hbm["frac"]}
        note = ("scene exceeds L2 (HBM-resident): algorithmic bytes (state + node + triangle fetches) / kernel time "
                "against the measured HBM copy bandwidth; `traffic` is the measured DRAM traffic of one launch")
    else:
        top = {"bound": "fp32", "achieved": fp32["achieved"], "peak": fp32["peak"], "unit": "TFLOP/s", "frac": fp32["frac"]}
        note = ("scene and path state are L2 / L1 resident (SURVEY.md 8d): the roof is FP32 issue + L2; algorithmic "
                "flops / kernel time against the measured FMA-chain peak, `l2` beside it; `hbm` is informational "
                "(algorithmic bytes are mostly L1 hits; measured DRAM traffic in `traffic`)")
    top.update({"traffic": traffic, "kernel": kernel_name, "kernel_ms": kernel_s * 1e3, "fp32": fp32, "l2": l2, "hbm": hbm,
                "note": note})
    return top


# ---- the reference's own renderer on the host cores ------------------------------------------------------
_REF_SCENES = {}


def ref_scene(name):
    from oracle import ref
    if name not in _REF_SCENES:
        _REF_SCENES[name] = ref.RefBuilder().realize(*build_scene(name))
    return _REF_SCENES[name]


def cpu_baseline(name, reps=3, single_thread=True):
    """oracle/_ref (the unmodified reference renderer: renderthread pool, std::thread x host cores, wall clock
    around the pool as at Raytracing_n.cpp:926-946) on a bounded sample of the configuration."""
    from oracle import ref
    if not ref.available():
        return {"value": None, "unit": "paths/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    _, _, nx, ny, ns, _, _ = config_of(name)
    sx, sy, ss = cpu_sample(nx, ny, ns)
    cores = os.cpu_count() or 1
    with stdout_to_stderr():
        t0 = time.time()
        rs = ref_scene(name)
        build_s = time.time() - t0
        rs.render_timed(sx, sy, ss, cores)
        secs = [rs.render_timed(sx, sy, ss, cores) for _ in range(reps)]
        one = rs.render_timed(sx, sy, ss, 1) if single_thread else None
        _, nrays = rs.render_float(max(sx // 4, 8), max(sy // 4, 8), ss, mod_ny=sx == sy)
    paths = sx * sy * ss
    rpp = nrays / float(max(sx // 4, 8) * max(sy // 4, 8) * ss)
    out = {"value": paths / float(np.mean(secs)), "unit": "paths/s", "cores": cores, "kind": "reference",
           "sample": f"{sx}x{sy} @ {ss} spp of the same scene/camera ({paths} paths), reference renderthread pool",
           "rays_per_path": rpp, "mrays_per_s": paths / float(np.mean(secs)) * rpp / 1e6,
           "scene_build_s": build_s}
    if one is not None:
        out["single_thread_value"] = paths / one
    return out


# ---- one configuration on the GPUs -------------------------------------------------------------------------
class Runtime:
    """Per-process state shared by every configuration: device, stream, context, process group, peaks."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        from simple_raytracing_render_b200 import render
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; this backend has no CPU path")
        torch.cuda.set_device(self.local_rank)
        self.dev = torch.device("cuda", self.local_rank)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.stream = torch.cuda.Stream(device=self.dev)
        self.ctx = render.Context(self.local_rank, self.stream.cuda_stream)
        self.flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=self.dev)
        hbm, src, sm_max = peaks()
        with torch.cuda.stream(self.stream):
            self.hw = {"hbm_gbps": hbm, "hbm_source": src, "sm_max_mhz": sm_max,
                       "fp32_tflops": self.ctx.measure_fp32_peak(), "l2_gbps": self.ctx.measure_l2_peak()}

    def sync_all(self):
        self.torch.cuda.synchronize(self.dev)
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        self.ctx.close()
        if self.world > 1:
            self.dist.destroy_process_group()


def measure_config(rt, name, steps, warmup, want_cpu, want_clocks=False):
    """value / kernel / counters / e2e / roofline (/ cpu_baseline) of one configuration on rt.world GPUs."""
    torch, dist = rt.torch, rt.dist
    from simple_raytracing_render_b200 import multigpu, render, scene as S
    scene_name, kw, nx, ny, ns, label, prim_bytes = config_of(name)
    n_pixels = nx * ny
    t0 = time.time()
    flat = S.flatten(*build_scene(name))
    host_scene_s = time.time() - t0
    t0 = time.time()
    sc = render.Scene(rt.ctx, flat)
    first_create_s = time.time() - t0
    info = sc.info()
    log(f"{name}: host scene {host_scene_s:.1f} s, srt_scene_create {first_create_s:.2f} s, "
        f"{info['n_prims']} prims, {info['device_bytes'] / 1e6:.1f} MB")
    s0, s1 = multigpu.sample_shard(ns, rt.rank, rt.world)
    params = render.render_params(nx, ny, ns, s0, s1, seed=SEED)
    # N = 1: the kernels add into `fixed`, srt_resolve_fixed resolves and clears it.  N > 1: the tiled exchange
    # (multigpu.exchange_tiles): reduce-scatter of the int64 buffers, resolve per tile, all-gather of the 8-bit tiles
    ex = multigpu.TileExchange(n_pixels, rt.world, rt.dev)
    fixed = ex.fixed
    rgb8 = torch.zeros(n_pixels * 3, dtype=torch.uint8, device=rt.dev)

    def resolve_tile(tile, n_tile_pixels, n_samples, rgb_tile):
        rt.ctx.resolve_fixed(tile, n_tile_pixels, n_samples, rgb_tile, None, clear=False)

    # N > 1, default: the exchange as ONE kernel over NVLink peer memory (multigpu.PeerExchange); SRT_PEER_EXCHANGE=0, or a
    # box without peer mappings, falls back to the NCCL form (reduce-scatter + per-tile resolve + all-gather)
    peer = None
    if rt.world > 1 and os.environ.get("SRT_PEER_EXCHANGE", "1") != "0":
        try:
            with torch.cuda.stream(rt.stream):
                peer = multigpu.PeerExchange(rt.ctx, n_pixels)
        except Exception as e:  # noqa: BLE001 — every rank fails alike (same box), so the fallback is collective too
            log(f"{name}: peer-memory exchange unavailable ({e}); NCCL tiles instead")
            peer = None
    exchange_kind = "one kernel over NVLink peer memory" if peer is not None else \
        ("NCCL int64 reduce-scatter + per-tile resolve + uint8 all-gather" if rt.world > 1 else "none (one GPU)")

    def render_into(scene):
        scene.render_fixed(params, peer.fixed() if peer is not None else fixed)

    def exchange(out, host=None):
        if peer is not None:
            peer.run(ns, host)
        elif rt.world > 1:
            multigpu.exchange_tiles(ex, ns, resolve_tile)
        else:
            rt.ctx.resolve_fixed(fixed, n_pixels, ns, out, None, clear=True)

    def step(out):
        # the whole frame: this rank's shard into the int64 buffer, the exchange step, the resolved image
        render_into(sc)
        exchange(out)

    clk = None
    with torch.cuda.stream(rt.stream):
        for _ in range(warmup):
            step(rgb8)
        rt.sync_all()
        launches0 = rt.ctx.counters()["kernel_launches"]
        clocks = ClockSampler(rt.local_rank) if (want_clocks and rt.rank == 0) else None
        if clocks:
            clocks.start()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for k in range(steps):
            rt.flush.fill_(k & 0xFF)  # L2 flush between timed iterations (untimed)
            if rt.world > 1:
                dist.barrier()
            ev[k][0].record(rt.stream)
            kev[k][0].record(rt.stream)
            render_into(sc)
            kev[k][1].record(rt.stream)
            exchange(rgb8)
            ev[k][1].record(rt.stream)
        rt.sync_all()
        if clocks:
            clk = clocks.stop()
        launches = rt.ctx.counters()["kernel_launches"] - launches0
    ms_per_step = rt.max_over_ranks(sum(a.elapsed_time(b) for a, b in ev)) / steps
    kernel_ms = sum(a.elapsed_time(b) for a, b in kev) / steps
    paths_per_step = n_pixels * ns
    value = paths_per_step / (ms_per_step * 1e-3)
    kernel_name = KERNEL_OF_MODE.get(int(sc.info()["render_variant"]) // 100, "k_render")

    # ---- untimed instrumented replay of the same launch: the timed kernel's own work counters ----------
    with torch.cuda.stream(rt.stream):
        scratch = torch.zeros(n_pixels * 3, dtype=torch.float32, device=rt.dev)
        cnt = sc.render_counted(params, scratch)
        del scratch
    rpp = cnt["rays"] / max(cnt["paths"], 1)

    # ---- e2e: HOST scene description in, 8-bit image in pinned HOST memory out, every step ----------------
    host_rgb = torch.empty(n_pixels * 3, dtype=torch.uint8).pin_memory()
    parts = {"scene_create_ms": 0.0, "render_ms": 0.0}

    def e2e_step():
        ta = time.perf_counter()
        sc2 = render.Scene(rt.ctx, flat)  # host scene description -> device (h2d), acceleration structures
        tb = time.perf_counter()
        parts["scene_create_ms"] += (tb - ta) * 1e3
        if rt.world > 1:
            render_into(sc2)
            if peer is not None:
                exchange(None, host_rgb.numpy() if rt.rank == 0 else None)  # d2h of the finished image + sync inside
            else:
                exchange(None)
                if rt.rank == 0:
                    host_rgb.copy_(ex.image(), non_blocking=True)  # d2h of the gathered 8-bit image
            torch.cuda.synchronize(rt.dev)
        else:
            sc2.render_image(params, rgb8=host_rgb.numpy())  # d2h + sync inside
        parts["render_ms"] += (time.perf_counter() - tb) * 1e3
        sc2.close()

    slow = first_create_s > 1.0 or ms_per_step > 500.0
    e2e_warm, e2e_steps = (1, 2) if slow else (2, max(3, min(steps, 10)))
    with torch.cuda.stream(rt.stream):
        for _ in range(e2e_warm):
            e2e_step()
        rt.sync_all()
        parts = {k: 0.0 for k in parts}
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        rt.sync_all()
        e2e_s = (time.perf_counter() - t0) / e2e_steps
    e2e_s = rt.max_over_ranks(e2e_s)
    overflow = rt.ctx.counters()["stack_overflows"]
    sc.close()
    if peer is not None:
        peer.close()
    del fixed, rgb8, ex

    res = None
    if rt.rank == 0:
        shard_frac = (s1 - s0) / float(ns)
        res = {
            "config": name, "workload": label, "nx": nx, "ny": ny, "spp": ns, "paths_per_step": paths_per_step,
            "value": value, "unit": "paths/s", "ms_per_step": ms_per_step, "steps": steps, "warmup": warmup,
            "mrays_per_s": value * rpp / 1e6, "rays_per_path": rpp,
            "kernel": kernel_name, "kernel_ms": kernel_ms, "kernel_share_of_step": kernel_ms / ms_per_step,
            "e2e": {"value": paths_per_step / e2e_s, "unit": "paths/s", "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
                    "h2d_bytes_per_step": int(info["device_bytes"]), "d2h_bytes_per_step": int(n_pixels * 3),
                    "parts_ms": {k: v / e2e_steps for k, v in parts.items()}},
            "roofline": make_roofline(name, cnt, n_pixels, prim_bytes, kernel_ms * 1e-3, kernel_name, shard_frac, rt.hw, ns),
            "counters": {k: cnt[k] for k in ("paths", "rays", "culled_rays", "probe_rays", "node_visits", "prim_tests",
                                             "reject_iters")},
            "counters_from": "instrumented build of " + kernel_name + " (srt_render_counted), rank 0's shard",
            "stack_overflows": int(overflow),
            "scene": {"prims": int(info["n_prims"]), "device_bytes": int(info["device_bytes"]),
                      "bvh_nodes": int(info["bvh_nodes"]), "bvh_max_depth": int(info["bvh_max_depth"]),
                      "bvh_builder": "device LBVH" if info.get("bvh_builder") else "host binned SAH",
                      "host_scene_s": host_scene_s, "scene_create_s": first_create_s},
            "gpu_launches": int(launches), "exchange": exchange_kind,
        }
        if clk is not None:
            res["clocks"] = clk
        if want_cpu:
            t0 = time.time()
            res["cpu_baseline"] = cpu_baseline(name, reps=3 if name != "tessellated_10m" else 1,
                                               single_thread=name != "tessellated_10m")
            log(f"{name}: cpu baseline {res['cpu_baseline'].get('value')} paths/s in {time.time() - t0:.1f} s")
            v = res["cpu_baseline"].get("value")
            if v:
                res["x_cpu_reference"] = {"value": value / v, "e2e": res["e2e"]["value"] / v}
    return res


def run_reference(args):
    """The reference's own renderer on the host cores: bounded sample of the same workload (the timed
    steps), ONE full-size frame of the headline configuration, and the sample of every other configuration."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import ref
    if not ref.available():
        emit({"impl": "reference", "unavailable": "oracle/_ref/libsrt_ref.so was not built"})
        return 0
    _, _, nx, ny, ns, label, _ = config_of(args.config)
    cores = os.cpu_count() or 1
    sx, sy, ss = cpu_sample(nx, ny, ns)
    paths = sx * sy * ss
    with stdout_to_stderr():
        rs = ref_scene(args.config)
        for _ in range(max(args.warmup, 1)):
            rs.render_timed(sx, sy, ss, cores)
        times = [rs.render_timed(sx, sy, ss, cores) for _ in range(args.steps)]
        _, nrays = rs.render_float(max(sx // 4, 8), max(sy // 4, 8), ss, mod_ny=sx == sy)
    rpp = nrays / float(max(sx // 4, 8) * max(sy // 4, 8) * ss)
    sec = float(np.mean(times))
    v = paths / sec
    sample = f"{sx}x{sy} @ {ss} spp of the same scene/camera ({paths} paths per step), {cores} std::threads"
    line = {
        "impl": "reference", "metric": "paths/s", "value": v, "unit": "paths/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": label, "sample": sample},
        "mrays_per_s": v * rpp / 1e6, "rays_per_path": rpp,
        "cpu_baseline": {"value": v, "unit": "paths/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": "paths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    # one frame of the headline configuration at its FULL size (the rate of the sample above is quoted as
    # resolution-independent: this is the check), bounded to configurations of a minute or so
    if not args.no_full_frame and nx * ny * ns / v <= 90.0:
        with stdout_to_stderr():
            t = rs.render_timed(nx, ny, ns, cores)
        line["full_frame"] = {"nx": nx, "ny": ny, "spp": ns, "paths": nx * ny * ns, "seconds": t,
                              "value": nx * ny * ns / t, "unit": "paths/s", "cores": cores,
                              "ratio_to_sample_rate": (nx * ny * ns / t) / v}
    if not args.no_per_config:
        per = []
        for name in ("cornell", "spheres", "teapot", "soldier", "tessellated"):
            if name == args.config:
                continue
            if time.time() - T_START > 420:
                break
            b = cpu_baseline(name, reps=2, single_thread=False)
            b["config"] = name
            per.append(b)
            log(f"reference {name}: {b['value']:.3e} paths/s")
        line["per_config"] = per
    emit(line)
    return 0


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
    ap.add_argument("--config", default="spheres", choices=list(CONFIGS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-per-config", action="store_true")
    ap.add_argument("--no-full-frame", action="store_true")
    ap.add_argument("--per-config", default=None, help="comma-separated configurations for per_config")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    rt = Runtime()
    want_cpu = rt.world == 1 and not args.no_cpu_baseline
    head = measure_config(rt, args.config, args.steps, args.warmup, want_cpu, want_clocks=True)
    per = []
    if not args.no_per_config:
        if args.per_config is not None:
            names = [n for n in args.per_config.split(",") if n]
        elif rt.world == 1:
            names = ["cornell", "spheres", "teapot", "soldier", "tessellated", "tessellated_10m"]
        else:  # the configurations BASELINE.json names for the 8-GPU box
            names = ["soldier", "tessellated_10m"]
        for name in names:
            if name == args.config:
                continue
            # every rank must take the same decision: rank 0's clock, broadcast
            elapsed = time.time() - T_START
            if rt.world > 1:
                t = rt.torch.tensor([elapsed], dtype=rt.torch.float64, device=rt.dev)
                rt.dist.broadcast(t, src=0)
                elapsed = float(t.item())
            go = elapsed < 600
            if not go:
                log(f"per_config: {name} skipped (time budget)")
                continue
            big = name in ("soldier", "tessellated", "tessellated_10m")
            r = measure_config(rt, name, 3 if big else 5, 3, want_cpu and (name != "tessellated_10m" or
                                                                          time.time() - T_START < 360))
            if r is not None:
                per.append(r)

    if rt.rank == 0:
        line = {
            "metric": "paths/s", "value": head["value"], "unit": "paths/s", "n_gpus": rt.world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": head["workload"], "nx": head["nx"], "ny": head["ny"], "spp": head["spp"],
                       "max_depth": 50, "parallelism": (f"sample-range x{rt.world}; exchange: {head['exchange']}" if rt.world > 1 else "one GPU"),
                       "l2": "flushed between timed steps (256 MiB write)",
                       "scene_device_bytes": head["scene"]["device_bytes"], "bvh_nodes": head["scene"]["bvh_nodes"]},
            "mrays_per_s": head["mrays_per_s"], "rays_per_path": head["rays_per_path"],
            "kernel_ms": head["kernel_ms"], "kernel_share_of_step": head["kernel_share_of_step"],
            "roofline": head["roofline"], "counters": head["counters"], "counters_from": head["counters_from"],
            "e2e": head["e2e"], "gpu_launches": head["gpu_launches"], "clocks": head.get("clocks"),
            "stack_overflows": head["stack_overflows"],
            "measured_peaks": rt.hw,
        }
        if "cpu_baseline" in head:
            line["cpu_baseline"] = head["cpu_baseline"]
            line["x_cpu_reference"] = head.get("x_cpu_reference")
        if per:
            line["per_config"] = per
        emit(line)
    rt.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
