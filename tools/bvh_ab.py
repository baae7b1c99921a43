"""tools/bvh_ab.py <config> : host SAH builder vs device builder (srt_lbvh.cuh) on one configuration:
srt_scene_create time, wide nodes, depth, node visits / primitive tests per ray of a counted render, kernel time."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from simple_raytracing_render_b200 import render, scene as S
name = sys.argv[1]
scene_name, kw, nx, ny, ns, label, _ = bench.config_of(name)
flat = S.flatten(*bench.build_scene(name))
stream = torch.cuda.Stream()
ctx = render.Context(0, stream.cuda_stream)
spp = min(ns, 16)
for dev in ("0", "1"):
    os.environ["SRT_DEVICE_BVH"] = dev
    ts = []
    for rep in range(3):
        if rep == 2:
            os.environ["SRT_SCENE_TIMING"] = "1"
        t0 = time.perf_counter()
        sc = render.Scene(ctx, flat)
        ts.append((time.perf_counter() - t0) * 1e3)
        os.environ.pop("SRT_SCENE_TIMING", None)
        if rep < 2:
            sc.close()
    info = sc.info()
    params = render.render_params(nx, ny, spp)
    buf = torch.zeros(nx * ny * 3, dtype=torch.float32, device="cuda")
    with torch.cuda.stream(stream):
        c = sc.render_counted(params, buf)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sc.render(params, buf)
        e0.record(stream)
        sc.render(params, buf)
        e1.record(stream)
        torch.cuda.synchronize()
    print(f"{name} device_bvh={dev}: scene_create ms {['%.1f' % t for t in ts]}, builder {info['bvh_builder']}, "
          f"nodes {info['bvh_nodes']}, depth {info['bvh_max_depth']}, sah {info['sah_cost']:.2f}, "
          f"node visits/ray {c['node_visits'] / c['rays']:.3f}, prim tests/ray {c['prim_tests'] / c['rays']:.3f}, "
          f"render {spp} spp {e0.elapsed_time(e1):.2f} ms, overflows {c['stack_overflows']}", flush=True)
    sc.close()
