"""srt_render_image_static (include/srt_ffi.h): the allocation-free form of srt_render_image — k_render on the
library's file-scope device scratch.  It has no counterpart in the reference,
so the check is against the existing entry point: same scene, same params, same bytes (the images of
this backend are bit-reproducible: integer accumulation).  Collected last on
purpose."""
import numpy as np
import pytest

from conftest import get_scene
from simple_raytracing_render_b200 import ffi, render

pytestmark = pytest.mark.gpu


def _static_image(sc, params):
    import torch
    n = params.nx * params.ny
    # pre-filled: every byte of the frame has to be written by the call
    out = torch.full((n, 3), 0x5A, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()  # torch's stream is not the context's
    sc.render_image_static(params, out)
    sc.ctx.synchronize()
    return out.cpu().numpy()


def test_static_scratch_render_matches_render_image_on_calls_in_a_row(gpu_ctx):
    """Calls in a row share one scratch array, which keeps the previous call's sums: a large frame, then a small
    one of another scene (its sums land on top of the first frame's), then the first again."""
    frames = [("cornell", 200, 160, 8, 0), ("spheres", 128, 96, 4, 0), ("cornell", 200, 160, 8, 0),
              ("spheres", 160, 120, 6, 2)]  # the last one: a shard [2, 6) of 6 samples
    scs = {}
    for name in ("cornell", "spheres"):
        world, hlist, cam = get_scene(name)
        scs[name] = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    want, got = [], []
    for name, nx, ny, ns, s0 in frames:
        p = render.render_params(nx, ny, ns, s_begin=s0, s_end=ns, max_depth=50)
        got.append(_static_image(scs[name], p))
    for name, nx, ny, ns, s0 in frames:
        p = render.render_params(nx, ny, ns, s_begin=s0, s_end=ns, max_depth=50)
        want.append(scs[name].render_image(p))
    for (name, nx, ny, ns, s0), g, w in zip(frames, got, want):
        assert g.shape == w.shape == (nx * ny, 3)
        assert len(np.unique(w)) > 16, "the comparison image is not a picture"
        assert np.array_equal(g, w), f"{name} {nx}x{ny}@{ns}: {int((g != w).sum())} bytes differ"
    assert np.array_equal(got[0], got[2])
    for sc in scs.values():
        sc.close()


def test_static_scratch_limits_are_refused_with_a_message(gpu_ctx):
    import torch
    world, hlist, cam = get_scene("cornell")
    sc = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    out = torch.zeros(3 * (ffi.STATIC_MAX_PIXELS + 1024), dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    with pytest.raises(ffi.BackendError, match="SRT_STATIC_MAX_PIXELS"):
        sc.render_image_static(render.render_params(1024, 257, 1), out)
    with pytest.raises(ffi.BackendError, match="SRT_STATIC_MAX_SAMPLES"):
        sc.render_image_static(render.render_params(8, 8, ffi.STATIC_MAX_SAMPLES + 1), out)
    with pytest.raises(ffi.BackendError, match="not device memory"):
        sc.render_image_static(render.render_params(8, 8, 1), np.zeros(8 * 8 * 3, np.uint8))
    # the largest frame the scratch holds is accepted (1 sample: the test is about the size)
    p = render.render_params(512, 512, 1)
    sc.render_image_static(p, out)
    gpu_ctx.synchronize()
    assert np.array_equal(out[: 3 * 512 * 512].cpu().numpy().reshape(-1, 3), sc.render_image(p))
    sc.close()
