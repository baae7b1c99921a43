"""tools/determinism_check.py <config> [spp] : the int64 sums of one frame must not depend on how paths are grouped
into warps and CTAs.  Renders the frame with the shipped schedule, again, with two CTAs per SM, with a smaller slot
pool, and on the fused kernel; prints how many of the nx*ny*3 sums differ from the first render and where."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from simple_raytracing_render_b200 import render, scene as S

name = sys.argv[1]
scene_name, kw, nx, ny, ns, label, _ = bench.config_of(name)
if len(sys.argv) > 2:
    ns = int(sys.argv[2])
stream = torch.cuda.Stream()
ctx = render.Context(0, stream.cuda_stream)
sc = render.Scene(ctx, S.flatten(*bench.build_scene(name)))
params = render.render_params(nx, ny, ns)
runs = [("shipped", {}), ("shipped again", {}), ("2 CTAs per SM", {"SRT_WF_CTAS_PER_SM": "2"}),
        ("4096 slots per SM", {"SRT_WF_SLOTS_PER_SM": "4096"}), ("1024 slots per SM", {"SRT_WF_SLOTS_PER_SM": "1024"})]
if "fused" in sys.argv:
    runs.append(("fused kernel", {"SRT_WAVEFRONT": "0"}))
first = None
for title, env in runs:
    for k, v in env.items():
        os.environ[k] = v
    buf = torch.zeros(nx * ny * 3, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        sc.render_fixed(params, buf)
        e1.record(stream)
    torch.cuda.synchronize()
    for k in env:
        del os.environ[k]
    if first is None:
        first = buf
    bad = (buf != first).nonzero().flatten()
    print(f"{name} @ {ns}: {title}: {e0.elapsed_time(e1):.1f} ms, {bad.numel()} of {buf.numel()} sums differ", flush=True)
    for i in bad[:6].tolist():
        print(f"   pixel {i // 3} (x {i // 3 % nx}, y {i // 3 // nx}) channel {i % 3}: {int(first[i])} vs {int(buf[i])}"
              f" (difference {(int(buf[i]) - int(first[i])) / 2 ** 31:.6g})", flush=True)
# the scene built a second time (flattener, tree builders, upload): same sums again
sc2 = render.Scene(ctx, S.flatten(*bench.build_scene(name)))
buf = torch.zeros(nx * ny * 3, dtype=torch.int64, device="cuda")
torch.cuda.synchronize()
sc2.render_fixed(params, buf)
torch.cuda.synchronize()
bad = (buf != first).nonzero().flatten()
print(f"{name} @ {ns}: scene built again: {bad.numel()} of {buf.numel()} sums differ", flush=True)
for i in bad[:6].tolist():
    print(f"   pixel {i // 3} (x {i // 3 % nx}, y {i // 3 // nx}) channel {i % 3}: {int(first[i])} vs {int(buf[i])}", flush=True)
i1, i2 = sc.info(), sc2.info()
print("info equal:", i1 == i2, {k: (i1[k], i2[k]) for k in i1 if i1[k] != i2[k]})
print("counters", ctx.counters())
