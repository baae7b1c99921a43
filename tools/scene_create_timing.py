import os, sys, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
os.environ["SRT_SCENE_TIMING"] = "1"
from simple_raytracing_render_b200 import render, scenes, scene as S
name = sys.argv[1]
t = time.perf_counter(); triple = scenes.by_name(name); t1 = time.perf_counter(); flat = S.flatten(*triple); t2 = time.perf_counter()
print(f"{name}: python scene {t1-t:.2f}s flatten {t2-t1:.2f}s", file=sys.stderr)
ctx = render.Context(0)
for i in range(2):
    t = time.perf_counter(); sc = render.Scene(ctx, flat); dt = time.perf_counter() - t
    print(f"Scene() {dt*1e3:.1f} ms", file=sys.stderr); sc.close()
