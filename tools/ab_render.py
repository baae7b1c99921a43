"""tools/ab_render.py <config> [fixed|float] [peaks] : kernel time of one configuration through a minimal call
sequence (no bench.py machinery), for A/B of library builds (SRT_B200_LIB) and call paths."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from simple_raytracing_render_b200 import render, scene as S
name = sys.argv[1]
mode = sys.argv[2] if len(sys.argv) > 2 else "float"
scene_name, kw, nx, ny, ns, label, _ = bench.config_of(name)
ns = int(os.environ.get("AB_NS", ns))  # fewer samples per pixel for a profiler run
reps = int(os.environ.get("AB_REPS", 4))
stream = torch.cuda.Stream()
ctx = render.Context(0, stream.cuda_stream)
if "peaks" in sys.argv:
    with torch.cuda.stream(stream):
        print("fp32", ctx.measure_fp32_peak(), "l2", ctx.measure_l2_peak())
sc = render.Scene(ctx, S.flatten(*bench.build_scene(name)))
params = render.render_params(nx, ny, ns)
buf = torch.zeros(nx * ny * 3, dtype=torch.int64 if mode == "fixed" else torch.float32, device="cuda")
with torch.cuda.stream(stream):
    for k in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        (sc.render_fixed if mode == "fixed" else sc.render)(params, buf)
        e1.record(stream)
        torch.cuda.synchronize()
        print(name, mode, os.environ.get("SRT_B200_LIB", "default"), "ms", e0.elapsed_time(e1), flush=True)
if mode == "fixed":  # integer sums: builds that render the same paths print the same number
    print(name, "checksum", int(buf.sum().item()), int((buf * torch.arange(buf.numel(), device="cuda") % 1000003).sum().item()), flush=True)
