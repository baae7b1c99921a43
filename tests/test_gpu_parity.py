"""Parity of the CUDA path (through the C ABI, libsrt_b200.so) with the reference.

Three gates (BASELINE.json:north_star): primary-hit primitive ids bit-exact against the reference's
world->hit on the same camera rays; hit distance t within 1e-5 relative; converged image within a
stated relMSE of the reference's CPU render.  The committed golden vectors (tests/golden/*.npz,
made from the UNMODIFIED reference by tests/golden/make_golden.py) are the fixed anchors; where
oracle/_ref/libsrt_ref.so travelled to the GPU box the same checks are repeated live at larger
sizes.  Nothing here reads /root/reference.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, get_ref_scene, get_scene, record_parity, trace_name
from simple_raytracing_render_b200 import ffi, render, scenes, scene as S

pytestmark = pytest.mark.gpu

T_REL_TOL = 1e-5       # north_star: "hit distance t within 1e-5 relative"
RELMSE_TOL = 1e-3      # north_star: "converged relMSE <= 1e-3 versus the reference"


def rel_mse(x, r):
    """mean over pixels/channels of (x - r)^2 / (r^2 + 1e-2) on linear means."""
    x, r = np.asarray(x, np.float64), np.asarray(r, np.float64)
    return float(np.mean((x - r) ** 2 / (r ** 2 + 1e-2)))


# C1-C5 of BASELINE.json:configs, then the parity-only scenes that instantiate every reference class those
# leave out (scenes.py: book_final, feature_zoo, zoo_env)
ALL_SCENES = ["cornell", "spheres", "teapot", "soldier", "soldier_wood", "tessellated", "final", "zoo", "zoo_env"]


@pytest.fixture(scope="module", params=ALL_SCENES)
def golden_scene(request, gpu_ctx):
    name = request.param
    scene = get_scene(name)  # skips when the scene needs assets that are not cached
    g = np.load(os.path.join(GOLDEN, f"{name}.npz"))
    sc = render.Scene.from_objects(gpu_ctx, *scene)
    yield name, g, sc
    sc.close()


@pytest.fixture(scope="module")
def trace_scene(golden_scene, gpu_ctx):
    """The scene the closest-hit vectors were made on: the same one, or — when it holds a
    constant_medium, whose hits are random (constant_medium.h:19-50) — its twin without the medium."""
    name, g, sc = golden_scene
    if trace_name(name) == name:
        yield name, g, sc
        return
    sc2 = render.Scene.from_objects(gpu_ctx, *get_scene(trace_name(name)))
    yield name, g, sc2
    sc2.close()


def _check_hits(gid, gt, rid, rt):
    assert np.array_equal(gid, rid), f"{int((gid != rid).sum())} of {len(rid)} primitive ids differ"
    hit = rid >= 0
    rel = np.abs(gt[hit] - rt[hit]) / np.maximum(np.abs(rt[hit]), 1e-30)
    assert rel.size == 0 or rel.max() <= T_REL_TOL, f"max relative t error {rel.max():.3e}"
    return float(rel.max()) if rel.size else 0.0


def test_primary_hits_match_golden(trace_scene):
    name, g, sc = trace_scene
    gid, gt, grec = sc.trace_closest(g["rays"])
    worst = _check_hits(gid, gt, g["ids"], g["t"])
    # the whole hit_record (u, v, p, normal) of hitable.h:17-25
    hit = g["ids"] >= 0
    assert np.allclose(grec[hit], g["rec"][hit], rtol=1e-5, atol=1e-4)
    record_parity(f"primary_hits/{name}", {"rays": int(len(gid)), "id_mismatches": 0, "max_rel_t_err": worst,
                                           "t_bit_equal": float(np.mean(gt[hit] == g["t"][hit])) if hit.any() else 1.0})
    assert sc.ctx.counters()["stack_overflows"] == 0


def test_secondary_hits_match_golden(trace_scene):
    """Un-normalised directions from surface points (what hitable_pdf samples produce); random ray times
    for the scenes with moving spheres."""
    name, g, sc = trace_scene
    gid, gt, grec = sc.trace_closest(g["rays2"])
    worst = _check_hits(gid, gt, g["ids2"], g["t2"])
    hit = g["ids2"] >= 0
    # env_sphere::hit with an un-normalised direction lands far outside the sphere: asin(y > 1) = NaN in v, on
    # both sides (env_sphere.h:27-38, hitable.h:10-15)
    assert np.array_equal(np.isnan(grec[hit]), np.isnan(g["rec2"][hit]))
    assert np.allclose(grec[hit], g["rec2"][hit], rtol=1e-5, atol=1e-4, equal_nan=True)
    record_parity(f"secondary_hits/{name}", {"rays": int(len(gid)), "id_mismatches": 0, "max_rel_t_err": worst})
    assert sc.ctx.counters()["stack_overflows"] == 0


def _check_light_pdf(got, want):
    """hitable_list::pdf_value: the zero / NaN pattern (which light shapes the ray reaches) is exact; values
    within 1e-5 relative (the density arithmetic after the exact hit test)."""
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = ~np.isnan(want)
    assert np.array_equal(got[ok] == 0, want[ok] == 0)
    assert np.allclose(got[ok], want[ok], rtol=1e-5, atol=0)
    nz = ok & (want != 0)
    return float(np.max(np.abs(got[nz] - want[nz]) / np.abs(want[nz]))) if nz.any() else 0.0


def test_light_pdf_matches_golden(trace_scene):
    name, g, sc = trace_scene
    worst = _check_light_pdf(sc.light_pdf(g["light_od"]), g["light_pdf"])
    rec = {"random_directions": int(len(g["light_pdf"])), "max_rel_err": worst,
           "nonzero": int(np.count_nonzero(np.nan_to_num(g["light_pdf"])))}
    if "light_aim_od" in g.files:
        # directions drawn by the reference's own hitable_list::random (sphere.h:80-86, triangle.h:89-94,
        # env_sphere.h:40-46, aarect.h:57-60): they reach the shapes, so every density formula is exercised
        worst = _check_light_pdf(sc.light_pdf(g["light_aim_od"]), g["light_aim_pdf"])
        rec.update({"aimed_directions": int(len(g["light_aim_pdf"])), "aimed_max_rel_err": worst,
                    "aimed_nonzero": int(np.count_nonzero(np.nan_to_num(g["light_aim_pdf"])))})
    record_parity(f"light_pdf/{name}", rec)


def test_device_pointers_are_accepted(trace_scene):
    import torch
    name, g, sc = trace_scene
    rays = torch.from_numpy(g["rays"][:4096].copy()).cuda()
    ids = torch.empty(len(rays), dtype=torch.int32, device="cuda")
    t = torch.empty(len(rays), dtype=torch.float32, device="cuda")
    ffi.check(sc.lib, sc.lib.srt_trace_closest(sc.h, rays.data_ptr(), len(rays), ids.data_ptr(), t.data_ptr(), None))
    sc.ctx.synchronize()
    assert np.array_equal(ids.cpu().numpy(), g["ids"][:4096])


# Scenes with a Beckmann material: beckmann_pdf::generate evaluates G with the WORLD-space incoming
# direction (pdf.h:145), so Lambda grows like 1 / |d.z| for rays nearly perpendicular to the world z
# axis and single samples weigh 1e6 and more — on both sides (the estimator's variance is infinite; a
# traced example is in profiles/r01_notes.md, capture C).  One such sample in one pixel dominates an
# untrimmed mean for any feasible spp, so for these scenes the gate drops the largest 0.2 % of the
# per-value terms (8 of 3 888); the untrimmed figure is printed.
TRIM = {"teapot": 0.002, "soldier": 0.002, "soldier_wood": 0.002}


@pytest.mark.parametrize("scheduler", ["auto", "sorted"])
def test_converged_image_relmse_vs_golden(golden_scene, scheduler, monkeypatch):
    """GPU image at 4x the golden's spp against the reference's single-thread float render; once with the
    scheduler the backend picks for this frame size (the fused kernel) and once through the sorted
    persistent wavefront (k_wf_sorted), which is what full-size frames of BVH scenes run on."""
    name, g, sc = golden_scene
    nx, ny, ns_ref = int(g["img_nx"]), int(g["img_ny"]), int(g["img_ns"])
    ns = 4 * ns_ref
    if scheduler == "sorted":
        monkeypatch.setenv("SRT_WAVEFRONT", "3")
    _, accum = sc.render_image(render.render_params(nx, ny, ns), want_accum=True)
    if scheduler == "sorted":
        assert sc.info()["render_variant"] // 100 == 3
    x = (accum / float(ns)).astype(np.float64)
    r = g["img"].astype(np.float64)
    terms = ((x - r) ** 2 / (r ** 2 + 1e-2)).reshape(-1)
    order = np.argsort(terms)
    keep = order[: int(round(len(order) * (1.0 - TRIM.get(name, 0.0))))]
    err, err_all = float(terms[keep].mean()), float(terms.mean())
    xm, rm = x.reshape(-1)[keep].mean(), r.reshape(-1)[keep].mean()
    print(f"{name}: relMSE {err:.3e} (untrimmed {err_all:.3e}; gpu {ns} spp vs reference {ns_ref} spp), "
          f"means {xm:.4f} / {rm:.4f}")
    record_parity(f"image_relmse/{name}/{scheduler}",
                  {"relmse": err, "relmse_untrimmed": err_all, "trim": TRIM.get(name, 0.0), "nx": nx, "ny": ny,
                   "gpu_spp": ns, "reference_spp": ns_ref, "gpu_mean": float(xm), "reference_mean": float(rm),
                   "kernel_variant": int(sc.info()["render_variant"])})
    assert sc.ctx.counters()["stack_overflows"] == 0
    assert abs(xm - rm) <= 0.02 * rm
    assert err <= RELMSE_TOL


@pytest.mark.parametrize("scheduler", ["auto", "sorted"])
def test_render_is_deterministic_and_shards_add_up(golden_scene, scheduler, monkeypatch):
    """Sample-range shards: [0, ns) == sum of disjoint ranges up to float summation
    order; the same range twice is bit-identical (fixed-point accumulation: for any schedule, including
    the wavefront's nondeterministic queue order)."""
    import torch
    name, g, sc = golden_scene
    if scheduler == "sorted":
        monkeypatch.setenv("SRT_WAVEFRONT", "3")
    nx, ny, ns = 96, 64, 32
    flags = 0

    def run(a, b):
        buf = torch.zeros(nx * ny * 3, dtype=torch.float32, device="cuda")
        sc.render(render.render_params(nx, ny, ns, a, b, flags=flags), buf)
        sc.ctx.synchronize()
        return buf.cpu().numpy()

    whole, again = run(0, ns), run(0, ns)
    assert np.array_equal(whole, again)
    parts = run(0, 8) + run(8, 16) + run(16, 32)
    assert np.allclose(parts, whole, rtol=1e-4, atol=1e-4)
    assert np.isfinite(whole).all() and whole.mean() > 0


def test_fixed_point_shards_are_bit_identical(golden_scene):
    """srt_render_fixed / srt_resolve_fixed (the multi-GPU form): the int64 sums of disjoint
    sample ranges add up to the unsharded buffer EXACTLY (integer addition), and resolving them gives the very
    bytes srt_render_image produces."""
    import torch
    name, g, sc = golden_scene
    nx, ny, ns = 96, 64, 32

    def run(ranges):
        buf = torch.zeros(nx * ny * 3, dtype=torch.int64, device="cuda")
        for a, b in ranges:
            sc.render_fixed(render.render_params(nx, ny, ns, a, b), buf)
        sc.ctx.synchronize()
        return buf

    whole = run([(0, ns)])
    parts = run([(0, 5), (5, 16), (16, 32)])
    assert torch.equal(whole, parts)
    rgb = np.zeros((nx * ny, 3), np.uint8)
    acc = torch.zeros(nx * ny * 3, dtype=torch.float32, device="cuda")
    sc.ctx.resolve_fixed(parts, nx * ny, ns, rgb, acc, clear=True)
    assert int(parts.abs().max()) == 0  # cleared for the next frame
    want_rgb, want_acc = sc.render_image(render.render_params(nx, ny, ns), want_accum=True)
    assert np.array_equal(rgb, want_rgb)
    assert np.array_equal(acc.cpu().numpy().reshape(-1, 3), want_acc)


def test_resolve_matches_reference_quantisation(gpu_ctx):
    """mean -> sqrt -> int(255.99 c) -> clamp (Raytracing_n.cpp:848-862)."""
    import torch
    rng = np.random.default_rng(1)
    n, ns = 4099, 16
    acc = (rng.uniform(0, 2.5, size=(n, 3)) * ns).astype(np.float32)
    acc[0] = 0
    acc[1] = 1e9
    d = torch.from_numpy(acc).cuda()
    out = np.zeros((n, 3), np.uint8)
    gpu_ctx.resolve(d, n, ns, out)
    k = np.float32(1.0 / ns)
    c = np.sqrt(acc * k).astype(np.float32)
    want = np.clip((255.99 * c.astype(np.float64)).astype(np.int64), 0, 255).astype(np.uint8)
    assert np.array_equal(out, want)


def test_counters_are_consistent(golden_scene, monkeypatch):
    """srt_render_counted replays a launch on the instrumented build of the scheduler the plain call runs:
    k_render for small frames, k_wf_sorted (the kernel bench.py times) when the sorted wavefront is
    selected.  Both walk the same paths (same Philox counters), so their tallies agree."""
    import torch
    name, g, sc = golden_scene
    nx, ny, ns = 64, 64, 8
    buf = torch.zeros(nx * ny * 3, dtype=torch.float32, device="cuda")
    c = sc.render_counted(render.render_params(nx, ny, ns), buf)
    assert c["paths"] == nx * ny * ns
    assert c["rays"] >= c["paths"] and c["prim_tests"] > 0
    assert c["stack_overflows"] == 0
    # the reference's own rays/path for this scene (golden img_rays) within a few percent
    ref_rpp = float(g["img_rays"]) / (int(g["img_nx"]) * int(g["img_ny"]) * int(g["img_ns"]))
    # The backend stops a path whose throughput is exactly zero (it cannot change the sample); the
    # reference keeps bouncing it.  culled_rays counts the first skipped ray of each such path, so
    # rays + culled is a lower bound of the reference's count and not far below it.
    got = (c["rays"] + c["culled_rays"]) / c["paths"]
    # `final`: a culled path would go on for many bounces inside the scene-wide constant_medium (every scatter
    # there is specular and keeps the zero throughput alive up to depth 50), so the bound is looser
    assert (0.6 if name == "final" else 0.85) * ref_rpp <= got <= 1.02 * ref_rpp
    monkeypatch.setenv("SRT_WAVEFRONT", "3")
    buf2 = torch.zeros_like(buf)
    w = sc.render_counted(render.render_params(nx, ny, ns), buf2)
    assert w["paths"] == c["paths"]
    for k in ("rays", "culled_rays", "probe_rays", "node_visits", "prim_tests", "reject_iters"):
        assert abs(w[k] - c[k]) <= 0.01 * max(c[k], 100), (k, w[k], c[k])
    a, b = buf.cpu().numpy(), buf2.cpu().numpy()
    assert np.isclose(a, b, rtol=1e-3, atol=1e-4).mean() >= 0.995
    record_parity(f"counters/{name}", {"fused": c, "sorted_wavefront": w, "reference_rays_per_path": ref_rpp,
                                       "gpu_rays_plus_culled_per_path": got})


# ---- live oracle (oracle/_ref/libsrt_ref.so travelled with the snapshot) ---------------------------
@pytest.mark.parametrize("name,nx,ny", [("cornell", 512, 512), ("spheres", 600, 400), ("teapot", 512, 512)])
def test_primary_hits_match_live_reference_full_size(oracle, gpu_ctx, name, nx, ny):
    rs = get_ref_scene(oracle, name)
    sc = render.Scene.from_desc_address(gpu_ctx, rs.desc_address())  # the C++ flattener's output
    rays = rs.primary_rays(nx, ny, 1, seed=7, mod_ny=nx == ny)
    rid, rt, _ = rs.trace(rays, want_rec=False)
    gid, gt, _ = sc.trace_closest(rays, want_rec=False)
    worst = _check_hits(gid, gt, rid, rt)
    print(f"{name} {nx}x{ny}: {len(rid)} rays, ids exact, max rel t err {worst:.2e}")
    sc.close()


def test_10m_triangle_scene_hits_match_golden(gpu_ctx):
    """BASELINE.json:configs[4] at its full size: 10 008 338 triangles in one bvh_node (2.8 GB on the device,
    HBM-resident).  The reference's closest hits on 100 000 of its own camera rays and 100 000 random rays
    (tests/golden/tessellated_10m.npz: ids and t only; made by make_golden.py tessellated_10m) against the
    wide-BVH traversal: ids exact, t within 1e-5, and no traversal-stack overflow in the deep tree."""
    import sys
    sys.path.insert(0, GOLDEN)
    from make_golden_rays import tess10m_random_rays
    g = np.load(os.path.join(GOLDEN, "tessellated_10m.npz"))
    world, hlist, cam = scenes.by_name("tessellated", n_side=int(g["n_side"]), fog=False)
    sc = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    info = sc.info()
    gid, gt, _ = sc.trace_closest(g["rays"], want_rec=False)
    worst1 = _check_hits(gid, gt, g["ids"], g["t"])
    gid2, gt2, _ = sc.trace_closest(tess10m_random_rays(len(g["ids2"]), int(g["seed2"])), want_rec=False)
    worst2 = _check_hits(gid2, gt2, g["ids2"], g["t2"])
    # a short render through the sorted wavefront exercises the same tree with incoherent bounce rays
    import torch
    buf = torch.zeros(480 * 270 * 3, dtype=torch.float32, device="cuda")
    sc.render(render.render_params(480, 270, 16, flags=ffi.RENDER_WAVEFRONT), buf)
    gpu_ctx.synchronize()
    c = gpu_ctx.counters()
    sc.close()
    print(f"10M triangles: {info['bvh_nodes']} wide nodes, depth {info['bvh_max_depth']}, ids exact on "
          f"{len(gid) + len(gid2)} rays, max rel t err {max(worst1, worst2):.2e}, stack overflows {c['stack_overflows']}")
    record_parity("primary_hits/tessellated_10m", {"rays": int(len(gid)), "id_mismatches": 0, "max_rel_t_err": worst1,
                                                   "triangles": int(info["n_prims"]), "bvh_max_depth": int(info["bvh_max_depth"])})
    record_parity("secondary_hits/tessellated_10m", {"rays": int(len(gid2)), "id_mismatches": 0, "max_rel_t_err": worst2})
    assert c["stack_overflows"] == 0
    assert torch.isfinite(buf).all() and float(buf.mean()) > 0


def _read_p3(path):
    tok = open(path).read().split()
    assert tok[0] == "P3"
    nx, ny, mx = int(tok[1]), int(tok[2]), int(tok[3])
    return nx, ny, mx, np.array(tok[4:], np.int64).reshape(-1, 3)


def test_python_crender_writes_the_reference_p3(gpu_ctx, tmp_path):
    """render.CRender (mirror of Render.h:16-51): the P3 file of Raytracing_n.cpp:886, :873-876 holds
    exactly what srt_render_image returns for the same parameters."""
    world, hlist, cam = scenes.by_name("cornell")
    path = str(tmp_path / "out.ppm")
    r = render.CRender(50, 64, 64, 8, 8, path)
    r.set_scene(world, hlist, cam)
    img = r.Run()
    nx, ny, mx, px = _read_p3(path)
    assert (nx, ny, mx) == (64, 64, 255) and r.isfinished and r.finishedPixel == 64 * 64
    assert np.array_equal(px, img.reshape(-1, 3).astype(np.int64))
    sc = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    again = sc.render_image(render.render_params(64, 64, 8))
    sc.close()
    assert np.array_equal(again, img)


def test_cpp_crender_run_through_reference_headers(oracle, gpu_ctx, tmp_path):
    """host/crender.hpp compiled against the UNMODIFIED reference headers (inside the oracle harness TU):
    CRender(...).Run() on real reference objects -> C++ flattener -> libsrt_b200.so -> P3 file, equal
    to the image the Python host renders from the mirrored scene (same flat scene, same seed)."""
    world, hlist, cam = scenes.by_name("cornell")
    rs = oracle.RefBuilder().realize(world, hlist, cam)
    path = str(tmp_path / "cpp.ppm")
    rs.crender_run(50, 64, 64, 8, ffi.BACKEND_PATH, path)
    nx, ny, mx, px = _read_p3(path)
    assert (nx, ny, mx) == (64, 64, 255)
    sc = render.Scene.from_objects(gpu_ctx, world, hlist, cam)
    img = sc.render_image(render.render_params(64, 64, 8))
    sc.close()
    assert np.array_equal(px, img.reshape(-1, 3).astype(np.int64))


def test_merl_lookup_matches_golden(gpu_ctx):
    """srt_merl_lookup (fp64 half/diff-angle index arithmetic of brdf.h:17-61, :106-153, :190-214 on the
    device) against the reference's lookup_brdf_val.  The table fetch is an exact copy, so values are
    bit-equal whenever the three bin indices agree; device libm may round an angle across a bin edge."""
    g = np.load(os.path.join(GOLDEN, "merl.npz"))
    sc = render.Scene.from_objects(gpu_ctx, *get_scene("soldier"))
    got = sc.merl_lookup(0, g["angles"])
    sc.close()
    same = (got == g["rgb"]).all(axis=1)
    # rows with (theta_out, phi_out) == (theta_in, phi_in) — what brdfmaterial::scatter looks up
    # (material.h:231) — have diff == (0, 0, 1) up to rounding, so phi_diff = atan2(+-1e-17, +-1e-17) is
    # decided by the last bit of libm's sin / cos: any phi bin may come out, on either side
    degenerate = (g["angles"][:, :2] == g["angles"][:, 2:]).all(axis=1)
    print(f"merl: {int(same[~degenerate].sum())} of {int((~degenerate).sum())} generic lookups bit-equal, "
          f"{int(same[degenerate].sum())} of {int(degenerate.sum())} degenerate ones")
    assert same[~degenerate].mean() >= 0.999
    assert np.allclose(got, g["rgb"], rtol=0.2)  # a neighbouring bin / another phi bin of the smooth table at worst


def test_peer_exchange_kernel_single_rank(gpu_ctx):
    """srt_exchange_* (the multi-GPU exchange step as one kernel over peer memory) with ONE rank: the
    kernel's own arithmetic — exact int64 sum of the tile, resolve, image store, alternating accumulation buffers that
    come back zero — against srt_resolve_fixed, over several frames and an image size that is not a multiple of 4.
    The N > 1 form (CUDA IPC, peer loads / stores, the in-kernel barrier) is exercised by tools/peer_exchange_check.py
    under torchrun and by the multi-GPU test below."""
    import ctypes as C
    import torch
    lib = gpu_ctx.lib
    sc = render.Scene.from_objects(gpu_ctx, *get_scene("cornell"))
    nx, ny, ns = 37, 23, 8
    n = nx * ny
    h = C.c_void_p()
    ffi.check(lib, lib.srt_exchange_create(gpu_ctx.h, n, 1, 0, C.byref(h)))
    handle = (C.c_uint8 * ffi.EXCHANGE_HANDLE_BYTES)()
    ffi.check(lib, lib.srt_exchange_handle(h, handle))
    ffi.check(lib, lib.srt_exchange_connect(h, bytes(handle)))
    fixed = torch.zeros(n * 3, dtype=torch.int64, device="cuda")
    want = torch.zeros(n * 3, dtype=torch.uint8, device="cuda")
    seen = set()
    for frame in range(4):
        p = render.render_params(nx, ny, ns, seed=100 + frame)
        buf = int(lib.srt_exchange_fixed(h))
        seen.add(buf)
        sc.render_fixed(p, buf)
        got = np.zeros(n * 3, np.uint8)
        ffi.check(lib, lib.srt_exchange_run(h, ns, got.ctypes.data))
        sc.render_fixed(p, fixed)
        gpu_ctx.resolve_fixed(fixed, n, ns, want, None, clear=True)
        gpu_ctx.synchronize()
        assert np.array_equal(got, want.cpu().numpy()), f"frame {frame}"
    assert len(seen) == 2  # the two accumulation buffers alternate
    lib.srt_exchange_destroy(h)
    sc.close()


def test_peer_exchange_on_two_gpus():
    """The N > 1 form, when the box has at least two GPUs: tools/peer_exchange_check.py under torchrun (peer-memory
    exchange == NCCL tiled exchange == unsharded render, bit for bit, on every rank)."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29531", os.path.join(root, "tools", "peer_exchange_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
