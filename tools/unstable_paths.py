"""tools/unstable_paths.py <config> : finds the (pixel, Sobol index) pairs whose radiance is not the same in two
renders of the same scene (tools/determinism_check.py found such pixels) and prints their bounce traces from
the counting build of the fused kernel, three times each."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from simple_raytracing_render_b200 import render, scene as S

name = sys.argv[1]
scene_name, kw, nx, ny, ns, label, _ = bench.config_of(name)
ctx = render.Context(0)
sc = render.Scene(ctx, S.flatten(*bench.build_scene(name)))


def frame(a, b, env, counted=False):
    for k, v in env.items():
        os.environ[k] = v
    dt = torch.float32 if counted else torch.int64
    buf = torch.zeros(nx * ny * 3, dtype=dt, device="cuda")
    torch.cuda.synchronize()
    p = render.render_params(nx, ny, ns, a, b)
    if counted:
        sc.render_counted(p, buf)
    else:
        sc.render_fixed(p, buf)
    ctx.synchronize()
    torch.cuda.synchronize()
    for k in env:
        del os.environ[k]
    return buf


unstable = []
for k in range(ns):
    a = frame(k, k + 1, {})
    b = frame(k, k + 1, {"SRT_WF_SLOTS_PER_SM": "2048"})
    c = frame(k, k + 1, {"SRT_WAVEFRONT": "0"})
    d = frame(k, k + 1, {"SRT_WAVEFRONT": "0", "SRT_GRAB": "32"})
    bad = ((a != b) | (c != d)).view(-1, 3).any(1).nonzero().flatten().tolist()
    for px in bad:
        print(f"sample {k} pixel {px} (x {px % nx}, y {px // nx}): wavefront {a.view(-1, 3)[px].tolist()} / "
              f"{b.view(-1, 3)[px].tolist()}  fused {c.view(-1, 3)[px].tolist()} / {d.view(-1, 3)[px].tolist()}", flush=True)
        unstable.append((px, k))
print(len(unstable), "unstable (pixel, sample) pairs", flush=True)
for px, k in unstable[:6]:
    for rep, grab in enumerate(("128", "32", "64")):
        print(f"--- pixel {px} sample {k}, replay {rep}", flush=True)
        frame(k, k + 1, {"SRT_WAVEFRONT": "0", "SRT_DEBUG_PIXEL": str(px), "SRT_DEBUG_SAMPLE": str(k), "SRT_GRAB": grab},
              counted=True)
