python tools/ab_render.py tessellated float
python tools/ab_render.py teapot float
python tools/ab_render.py spheres float
SRT_B200_LIB=$PWD/simple_raytracing_render_b200/csrc/build/libsrt_tiny.so python - <<'PY'
import sys; sys.path.insert(0,'.')
import torch, bench
from simple_raytracing_render_b200 import render, scene as S
ctx = render.Context(0)
sc = render.Scene(ctx, S.flatten(*bench.build_scene("teapot")))
buf = torch.zeros(256*256*3, dtype=torch.float32, device="cuda")
sc.render(render.render_params(256,256,4), buf); ctx.synchronize()
print("tiny stack (4 entries): stack_overflows =", ctx.counters()["stack_overflows"])
PY
