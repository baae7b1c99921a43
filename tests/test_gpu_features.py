"""Isolated GPU checks of the device code of single reference classes (through the C ABI), against vectors
made from the unmodified reference (tests/golden/features.npz, made by make_golden.py) and, where
oracle/_ref/libsrt_ref.so travelled with the snapshot, against the live reference:

  * camera::get_ray on the device (camera.h:51-59; the rays srt_render starts its paths with),
  * constant_medium::hit free-flight sampling (constant_medium.h:19-50) on >= 1e6 rays,
  * noise_texture + perlin, checker_texture, image_texture through rect and sphere uv (texture.h, perlin.h),
  * orennayar / onrennayar_pdf alone (material.h:127-149, pdf.h:61-116).
"""
import os
import sys

import numpy as np
import pytest

from conftest import GOLDEN, get_scene, record_parity
from simple_raytracing_render_b200 import ffi, render, scenes, scene as S

sys.path.insert(0, GOLDEN)
from make_golden_rays import medium_rays  # noqa: E402  (the seeded ray set shared with make_golden.py)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def feat():
    return np.load(os.path.join(GOLDEN, "features.npz"))


# ---- a3: camera::get_ray on the device ----------------------------------------------------------------
@pytest.mark.parametrize("name", ["cornell", "teapot", "tessellated", "final", "zoo_env"])
def test_device_camera_rays_equal_reference_rays(gpu_ctx, name):
    """lens_radius == 0: the device's origin and direction equal the reference's camera::get_ray bit for bit
    on the golden primary rays (same pixel order, same Sobol jitter, same float expressions).  The ray TIME
    is time0 + drand48() * (time1 - time0) from the path's own Philox stream: only its range is checked,
    and bit-equality where time0 == time1."""
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    nx, ny, ns = int(g["nx"]), int(g["ny"]), int(g["ns_rays"])
    world, hlist, cam = get_scene(name)
    assert float(cam.lens_radius) == 0.0
    sc = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    rays = sc.primary_rays(render.render_params(nx, ny, ns))
    sc.close()
    want = g["rays"]
    assert rays.shape == want.shape
    assert np.array_equal(rays[:, :6].view(np.uint32), want[:, :6].view(np.uint32))
    t0, t1 = float(cam.time0), float(cam.time1)
    if t0 == t1:
        assert np.array_equal(rays[:, 6], want[:, 6])
    else:
        assert rays[:, 6].min() >= t0 and rays[:, 6].max() <= t1
        assert abs(rays[:, 6].mean() - 0.5 * (t0 + t1)) < 0.02 * (t1 - t0)
    record_parity(f"camera_rays/{name}", {"rays": int(len(rays)), "origin_direction_bit_equal": True})


@pytest.mark.parametrize("name", ["spheres", "soldier", "zoo"])
def test_device_camera_lens_branch(gpu_ctx, name):
    """lens_radius != 0 (camera.h:8-14, :53-57): the lens sample is random, so the device rays are checked
    through what the reference's arithmetic guarantees for ANY sample: the origin lies on the lens disk
    (in the u-v plane through the camera origin), the direction is unit, the ray passes through the pixel's
    point on the focus plane, and the offsets are uniform on the disk (mean 0, E[r^2] = R^2 / 2)."""
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    nx, ny, ns = int(g["nx"]), int(g["ny"]), int(g["ns_rays"])
    world, hlist, cam = get_scene(name)
    R = float(cam.lens_radius)
    assert R > 0
    sc = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    p = render.render_params(nx, ny, ns)
    rays = sc.primary_rays(p).astype(np.float64)
    sc.close()
    c = cam.record()
    origin, u, v, w = (np.asarray(c[k], np.float64).reshape(3) for k in ("origin", "u", "v", "w"))
    llc, hor, ver = (np.asarray(c[k], np.float64).reshape(3) for k in ("lower_left_corner", "horizontal", "vertical"))
    off = rays[:, :3] - origin
    scale = max(1.0, float(np.abs(origin).max()))
    assert np.abs(off @ w).max() <= 1e-5 * scale            # in the lens plane
    r2 = (off @ u) ** 2 + (off @ v) ** 2
    assert r2.max() <= (R * (1 + 1e-4)) ** 2 + 1e-6 * scale  # on the disk
    assert abs(np.linalg.norm(rays[:, 3:6], axis=1) - 1.0).max() <= 1e-5
    # the pixel's focus-plane point: lower_left + s * horizontal + t * vertical (Raytracing_n.cpp:834-835)
    sob = np.zeros((max(ns, 2), 2), np.uint32)
    ffi.load_backend().srt_sobol_points_u32(len(sob), sob.ctypes.data)
    sob = sob.astype(np.float64) / 4294967296.0
    pix = np.repeat(np.arange(nx * ny), ns)
    smp = np.tile(np.arange(ns), nx * ny)
    i = pix % (ny if nx == ny else nx)
    j = ny - 1 - pix // nx
    s_ = (sob[smp, 0] + i) / nx
    t_ = (sob[smp, 1] + j) / ny
    target = llc + s_[:, None] * hor + t_[:, None] * ver
    to = target - rays[:, :3]
    miss = np.linalg.norm(np.cross(to, rays[:, 3:6]), axis=1) / np.linalg.norm(to, axis=1)
    assert miss.max() <= 2e-5
    n = len(rays)
    assert abs((off @ u).mean()) <= 5 * R / np.sqrt(n) and abs((off @ v).mean()) <= 5 * R / np.sqrt(n)
    assert abs(r2.mean() - 0.5 * R * R) <= 5 * R * R / np.sqrt(12 * n) + 1e-6
    record_parity(f"camera_rays/{name}", {"rays": int(n), "lens_radius": R, "max_focus_miss": float(miss.max()),
                                          "mean_r2_over_R2": float(r2.mean() / (R * R))})


# ---- a12: constant_medium --------------------------------------------------------------------------------
def _chords(kind, rays):
    """(t_enter, t_exit) of each ray with the boundary in float64; no hit -> t_exit < t_enter."""
    o, d = rays[:, :3].astype(np.float64), rays[:, 3:6].astype(np.float64)
    if kind == "sphere":
        a = (d * d).sum(1)
        b = (o * d).sum(1)
        c = (o * o).sum(1) - 1.0
        disc = b * b - a * c
        ok = disc > 0
        sq = np.sqrt(np.where(ok, disc, 0.0))
        return np.where(ok, (-b - sq) / a, 1.0), np.where(ok, (-b + sq) / a, 0.0)
    lo, hi = np.array([-1.0, -0.6, -0.8]), np.array([1.0, 0.6, 0.8])
    with np.errstate(divide="ignore", invalid="ignore"):
        ta, tb = (lo - o) / d, (hi - o) / d
    return np.minimum(ta, tb).max(1), np.maximum(ta, tb).min(1)


@pytest.mark.parametrize("kind", ["sphere", "box"])
def test_constant_medium_transmittance(gpu_ctx, feat, kind):
    """constant_medium::hit on 1e6 rays through a homogeneous volume of density 0.7: the scatter
    probability of every ray is 1 - exp(-density * |d| * (t_exit - max(t_enter, t_min))) and the scatter
    distance is exponential inside the chord.  The device result (Philox) is compared with the analytic
    expectation and with the reference's own statistics on the same rays (drand48), both within 5 sigma."""
    density = 0.7
    world, hlist, cam = scenes.medium_probe(kind, density)
    sc = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    rays = medium_rays(kind, 1_000_000, 21)
    ids, t, _ = sc.trace_closest(rays, want_rec=False)
    n_prims = sc.n_prims
    sc.close()
    assert set(np.unique(ids).tolist()) <= {-1, n_prims}
    assert np.array_equal(np.unique(ids), feat[f"medium_{kind}_ids"])
    hit = ids == n_prims
    n = len(rays)
    # analytic: chord clipped to (t_min, inf), in distance units
    t_in, t_out = _chords(kind, rays)
    t1 = np.maximum(t_in, 0.001)
    L = np.maximum(t_out - t1, 0.0) * np.linalg.norm(rays[:, 3:6].astype(np.float64), axis=1)
    L[t_out <= t_in] = 0.0
    p = 1.0 - np.exp(-density * L)
    sigma = np.sqrt((p * (1 - p)).sum()) / n
    frac = hit.mean()
    assert abs(frac - p.mean()) <= 5 * sigma + 1e-6, (frac, p.mean(), sigma)
    ref_frac = float(feat[f"medium_{kind}_scatter_frac"])
    assert abs(frac - ref_frac) <= 5 * np.sqrt(2) * sigma + 1e-6, (frac, ref_frac)
    # scatter points lie inside the chord
    assert (t[hit] >= t1[hit] - 1e-4).all() and (t[hit] <= t_out[hit] + 1e-4).all()
    # mean scatter parameter against the reference's (same rays, independent random numbers)
    mt, ref_mt = t[hit].astype(np.float64).mean(), float(feat[f"medium_{kind}_mean_t"])
    ref_var = float(feat[f"medium_{kind}_mean_t2"]) - ref_mt ** 2
    se = np.sqrt(2 * ref_var / hit.sum())
    assert abs(mt - ref_mt) <= 5 * se, (mt, ref_mt, se)
    # the distance travelled inside the volume is exponential with rate `density`, truncated at the chord:
    # E[s | scatter] = 1/density - L e^{-density L} / (1 - e^{-density L})
    dlen = np.linalg.norm(rays[:, 3:6].astype(np.float64), axis=1)
    s_got = ((t - t1) * dlen)[hit]
    Lh = L[hit]
    s_exp = 1.0 / density - Lh * np.exp(-density * Lh) / np.maximum(1 - np.exp(-density * Lh), 1e-12)
    assert abs(s_got.mean() - s_exp.mean()) <= 5 * s_got.std() / np.sqrt(hit.sum())
    record_parity(f"constant_medium/{kind}", {"rays": n, "scatter_fraction_gpu": float(frac),
                                              "scatter_fraction_reference": ref_frac,
                                              "scatter_fraction_analytic": float(p.mean()), "sigma": float(sigma),
                                              "mean_t_gpu": float(mt), "mean_t_reference": ref_mt})


# ---- a19 / a13: textures ---------------------------------------------------------------------------------
@pytest.mark.parametrize("kind", ["noise", "checker", "image", "image_sphere"])
def test_texture_probe_images(gpu_ctx, feat, kind):
    """An emissive surface fills the view, so a pixel is the mean of texture::value at its 4 primary hits
    (same Sobol jitter on both sides, no random numbers involved): the images agree to float rounding of
    the texture arithmetic, except where a sample sits on a checker / texel edge."""
    world, hlist, cam = scenes.texture_probe(kind)
    sc = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    _, accum = sc.render_image(render.render_params(64, 64, 4), want_accum=True)
    sc.close()
    x, r = accum / 4.0, feat[f"tex_{kind}"]
    diff = np.abs(x - r)
    close = diff <= 2e-3 + 1e-3 * np.abs(r)
    print(f"texture {kind}: max abs diff {diff.max():.3e}, values within tolerance {close.mean():.5f}")
    record_parity(f"texture_probe/{kind}", {"max_abs_diff": float(diff.max()), "median_abs_diff": float(np.median(diff)),
                                            "within_tolerance": float(close.mean())})
    assert r.std() > 0.05  # the probe sees texture variation, not a constant
    assert close.mean() >= (0.999 if kind == "noise" else 0.99)
    assert abs(x.mean() - r.mean()) <= 2e-3 * max(r.mean(), 1e-3)


# ---- a14 / a17: orennayar + onrennayar_pdf ------------------------------------------------------------------
@pytest.mark.parametrize("scheduler", ["auto", "sorted"])
def test_orennayar_only_scene(gpu_ctx, feat, scheduler, monkeypatch):
    if scheduler == "sorted":
        monkeypatch.setenv("SRT_WAVEFRONT", "3")
    sc = render.Scene.from_objects(gpu_ctx, *scenes.oren_only())
    ns = 4 * int(feat["oren_ns"])
    _, accum = sc.render_image(render.render_params(32, 32, ns), want_accum=True)
    c = gpu_ctx.counters()
    sc.close()
    x, r = (accum / float(ns)).astype(np.float64), feat["oren_img"].astype(np.float64)
    err = float(np.mean((x - r) ** 2 / (r ** 2 + 1e-2)))
    print(f"orennayar-only: relMSE {err:.3e}, means {x.mean():.4f} / {r.mean():.4f}")
    record_parity(f"image_relmse/oren_only/{scheduler}", {"relmse": err, "relmse_untrimmed": err, "trim": 0.0, "nx": 32,
                                                          "ny": 32, "gpu_spp": ns, "reference_spp": int(feat["oren_ns"]),
                                                          "gpu_mean": float(x.mean()), "reference_mean": float(r.mean())})
    assert c["stack_overflows"] == 0
    assert abs(x.mean() - r.mean()) <= 0.01 * r.mean()
    assert err <= 1e-3


# ---- device-side builds of srt_scene_create (large scenes) -------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 7, 8, 9, 31, 100, 1023, 1025, 32768, 40001, 999999])
def test_device_reference_tree_equals_host_recursion(gpu_ctx, n):
    """k_ref_tree_level (one launch per level, node indices from the level tables) writes the same records as the
    host recursion that follows bvh.h:96-119 step by step: indices, children, split ranks and the box bits."""
    rng = np.random.default_rng(1000 + n)
    lo = rng.uniform(-50, 50, size=(n, 3)).astype(np.float32)
    boxes = np.concatenate([lo, lo + rng.uniform(0, 3, size=(n, 3)).astype(np.float32)], axis=1)
    host = gpu_ctx.ref_tree_boxes(boxes, on_device=False)
    dev = gpu_ctx.ref_tree_boxes(boxes, on_device=True)
    assert host.shape == dev.shape and host.shape[0] >= 1
    assert np.array_equal(host, dev)
    # root box = union of the item boxes
    root = host[0].view(np.float32)
    assert np.array_equal(root[0:3], boxes[:, :3].min(axis=0)) and np.array_equal(root[4:7], boxes[:, 3:].max(axis=0))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["teapot", "tessellated"])
def test_device_scene_build_renders_the_same_image(gpu_ctx, name, monkeypatch):
    """SRT_DEVICE_BVH=1 (device LBVH + per-cluster SAH, device reference tree, device primitive records) and =0 (host
    builders) are two search orders over the same primitives: closest hits, hence every pixel, are identical."""
    import torch
    world, hlist, cam = get_scene(name)
    flat = S.flatten(world, hlist, cam)
    nx, ny, ns = 96, 64, 4
    imgs, infos = [], []
    for dev in ("1", "0"):
        monkeypatch.setenv("SRT_DEVICE_BVH", dev)
        sc = render.Scene(gpu_ctx, flat)
        infos.append(sc.info())
        buf = torch.zeros(nx * ny * 3, dtype=torch.int64, device="cuda")
        sc.render_fixed(render.render_params(nx, ny, ns), buf)
        gpu_ctx.synchronize()
        imgs.append(buf.cpu().numpy().copy())
        sc.close()
    assert infos[0]["bvh_builder"] == 1 and infos[1]["bvh_builder"] == 0
    assert np.abs(imgs[0]).max() > 0
    assert np.array_equal(imgs[0], imgs[1])
    assert gpu_ctx.counters()["stack_overflows"] == 0
