"""NEE + MIS against BSDF sampling alone, one light kind / one material at a time (corrected-transport mode)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from simple_raytracing_render_b200 import ffi, render, scenes
ctx = render.Context(0)
def run(kw, sampled, ns):
    sc = render.Scene.from_objects(ctx, *scenes.corrected_probe(1.0, sampled=sampled, **kw))
    buf = torch.zeros(40 * 40 * 3, dtype=torch.float32, device="cuda")
    sc.render(render.render_params(40, 40, ns, flags=ffi.RENDER_CORRECTED), buf)
    ctx.synchronize(); sc.close()
    return buf.cpu().numpy().reshape(40, 40, 3) / ns
cases = {"rect+lambert": dict(lights=("rect",), materials=()), "sphere+lambert": dict(lights=("sphere",), materials=()),
         "tri+lambert": dict(lights=("tri",), materials=()), "rect+merl": dict(lights=("rect",), materials=("merl",)),
         "rect+oren": dict(lights=("rect",), materials=("oren",)), "rect+beckmann": dict(lights=("rect",), materials=("beckmann",)),
         "rect+glass": dict(lights=("rect",), materials=("glass",)), "all": {}}
for name, kw in cases.items():
    a, b = run(kw, True, 4096), run(kw, False, 32768)
    rel = float(np.mean((a - b) ** 2 / (b * b + 1e-2)))
    print(f"{name:16s} NEE {a.mean():.5f}  BSDF {b.mean():.5f}  ratio {a.mean() / b.mean():.4f}  relMSE {rel:.2e}", flush=True)
