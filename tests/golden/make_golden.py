"""Generates tests/golden/*.npz from the REAL reference (oracle/_ref/libsrt_ref.so, i.e. the
unmodified /root/reference translation unit).  Run in the build container:

    python tests/golden/make_golden.py

The vectors pin (a) the reference's generators and closed-form helpers, (b) its closest-hit results
on its own camera rays for the C1 / C2 scenes, (c) its light pdf, (d) converged float images.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

sys.path.insert(0, os.path.dirname(HERE))
from conftest import get_scene, trace_name  # noqa: E402
from make_golden_rays import medium_rays, tess10m_random_rays  # noqa: E402
from oracle import ref  # noqa: E402
from simple_raytracing_render_b200 import scenes  # noqa: E402


def kats():
    L = ref.lib()
    f, u = ref.kat_pcg32(16)
    xs = np.linspace(-3, 3, 25).astype(np.float32)
    ys = np.linspace(-0.999, 0.999, 25).astype(np.float32)
    cam = scenes.cornell_box_classic(1.0)[2]
    np.savez(os.path.join(HERE, "kats.npz"),
             drand48_seed1=ref.kat_drand48(1, 16), pcg_float=f, pcg_u32=u,
             sobol_8=ref.kat_sobol(8), sobol_1024=ref.kat_sobol(1024),
             erf_x=xs, erf=np.array([L.ref_kat_erf(float(x)) for x in xs], np.float32),
             erfinv_x=ys, erfinv=np.array([L.ref_kat_erfinv(float(y)) for y in ys], np.float32),
             schlick=np.array([L.ref_kat_schlick(c, 1.5) for c in (0.0, 0.25, 0.5, 1.0)], np.float32),
             r2a=np.array([L.ref_kat_roughness_to_alpha(r) for r in (0.0005, 0.01, 0.05, 0.3, 1.0)], np.float32))


def merl_angles(n=4096, seed=11):
    rng = np.random.default_rng(seed)
    a = np.stack([rng.uniform(0, np.pi / 2, n), rng.uniform(0, 2 * np.pi, n),
                  rng.uniform(0, np.pi / 2, n), rng.uniform(0, 2 * np.pi, n)], 1)
    a[: n // 8, 2:] = a[: n // 8, :2]  # brdfmaterial::scatter looks up (theta_in, phi_in) twice (material.h:231)
    return a


def merl_kat():
    """brdf::lookup_brdf_val (brdf.h:190-214) of the reference on the synthetic MERL-layout table."""
    from simple_raytracing_render_b200 import scene as S
    m = S.brdfmaterial(scenes.synthetic_merl_table(), S.vec3(0.7, 0.7, 0.7))
    a = merl_angles()
    np.savez_compressed(os.path.join(HERE, "merl.npz"), angles=a, rgb=ref.merl_lookup(m, a))


def zero_unset_uv(rs, ids, rec):
    """moving_sphere::hit never writes rec.u / rec.v (moving_sphere.h:32-48): what the oracle reports there
    is whatever its stack held.  The vectors carry 0 for those hits."""
    from simple_raytracing_render_b200 import ffi
    prims = rs.flat().prims
    if len(prims) == 0:
        return
    moving = np.zeros(len(ids), bool)
    ok = (ids >= 0) & (ids < len(prims))
    moving[ok] = prims["type"][ids[ok]] == ffi.PRIM_MOVING_SPHERE
    rec[moving, 0:2] = 0.0


def _render_chunk(args):
    name, nx, ny, ns, s0, s1, seed = args
    rs = ref.RefBuilder().realize(*get_scene(name))
    return rs.render_float(nx, ny, ns, s0, s1, seed=seed, mod_ny=nx == ny)


def render_float_parallel(name, nx, ny, ns, procs=6):
    """The oracle's float image at a high sample count: disjoint Sobol index ranges in separate processes
    (each with its own drand48 seed: the streams are independent), summed.  Single-threaded per process, as
    the reference's global generators demand."""
    import multiprocessing as mp
    step = (ns + procs - 1) // procs
    jobs = [(name, nx, ny, ns, s0, min(ns, s0 + step), 1 + k) for k, s0 in enumerate(range(0, ns, step))]
    with mp.get_context("fork").Pool(procs) as pool:
        parts = pool.map(_render_chunk, jobs)
    return sum(p[0] for p in parts), sum(p[1] for p in parts)


def scene_vectors(name, nx, ny, ns_rays, ns_img, img_nx, img_ny, extended=False):
    world, hlist, cam = get_scene(name)
    rs_img = ref.RefBuilder().realize(world, hlist, cam)
    rs = rs_img if trace_name(name) == name else ref.RefBuilder().realize(*get_scene(trace_name(name)))
    mod_ny = nx == ny
    rays = rs.primary_rays(nx, ny, ns_rays, seed=1, mod_ny=mod_ny)
    ids, t, rec = rs.trace(rays)
    zero_unset_uv(rs, ids, rec)
    rng = np.random.default_rng(7)
    # secondary-like rays: from hit points towards random directions (un-normalised, as light samples are)
    hit = ids >= 0
    o = rec[hit][:, 2:5][:4096]
    d = rng.normal(size=o.shape).astype(np.float32) * rng.uniform(0.2, 300, size=(len(o), 1)).astype(np.float32)
    tm = rng.uniform(0, 1, size=(len(o), 1)) if extended else np.zeros((len(o), 1))  # moving_sphere.h:19-21
    rays2 = np.concatenate([o, d, tm], 1).astype(np.float32)
    ids2, t2, rec2 = rs.trace(rays2)
    zero_unset_uv(rs, ids2, rec2)
    od = np.concatenate([o, d], 1).astype(np.float32)
    lp = rs.light_pdf(od)
    extra = {}
    if extended:  # directions drawn by the reference's own hitable_list::random: they do reach the light shapes
        aim = rs.light_random(o, seed=3)
        extra["light_aim_od"] = np.concatenate([o, aim], 1).astype(np.float32)
        extra["light_aim_pdf"] = rs.light_pdf(extra["light_aim_od"])
    if ns_img > 8192:  # a high-variance scene: the golden's own noise must stay well below the 1e-3 gate
        img, nrays = render_float_parallel(name, img_nx, img_ny, ns_img)
    else:
        img, nrays = rs_img.render_float(img_nx, img_ny, ns_img, mod_ny=img_nx == img_ny)
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), nx=nx, ny=ny, ns_rays=ns_rays,
                        rays=rays, ids=ids, t=t, rec=rec, rays2=rays2, ids2=ids2, t2=t2, rec2=rec2,
                        light_od=od, light_pdf=lp, img=img / ns_img, img_nx=img_nx, img_ny=img_ny, img_ns=ns_img,
                        img_rays=nrays, **extra)
    print(name, "hits", int(hit.sum()), "of", len(ids), "img mean", float(img.mean() / ns_img), "rays/path",
          nrays / (img_nx * img_ny * ns_img))


def tessellated_10m():
    """tests/golden/tessellated_10m.npz: the reference's closest hits on the FULL C5 geometry (10 008 338
    triangles in one bvh_node, BASELINE.json:configs[4]; the constant_medium is left out, its hits are random):
    400 x 250 primary rays and 100 000 random rays.  Takes ~3 minutes and ~4 GB here (10 M `triangle`
    objects of 208 B and the reference's O(n log^2 n) bvh_node constructor)."""
    world, hlist, cam = scenes.by_name("tessellated", n_side=2237, fog=False)
    rs = ref.RefBuilder().realize(world, hlist, cam)
    nx, ny = 400, 250
    rays = rs.primary_rays(nx, ny, 1, seed=1)
    ids, t, _ = rs.trace(rays, want_rec=False)
    ids2, t2, _ = rs.trace(tess10m_random_rays(), want_rec=False)
    np.savez_compressed(os.path.join(HERE, "tessellated_10m.npz"), n_side=2237, nx=nx, ny=ny, rays=rays, ids=ids, t=t,
                        seed2=9, ids2=ids2, t2=t2)


def feature_vectors():
    """tests/golden/features.npz: texture-probe images (every pixel is texture::value at the primary hit),
    the Oren-Nayar-only scene, and the constant_medium transmittance statistics of the reference."""
    out = {}
    for kind in ("noise", "checker", "image", "image_sphere"):
        rs = ref.RefBuilder().realize(*scenes.texture_probe(kind))
        img, _ = rs.render_float(64, 64, 4, mod_ny=True)
        out[f"tex_{kind}"] = img / 4.0
    rs = ref.RefBuilder().realize(*scenes.oren_only())
    img, nrays = rs.render_float(32, 32, 8192, mod_ny=True)
    out["oren_img"], out["oren_ns"], out["oren_rays"] = img / 8192.0, 8192, nrays
    for kind in ("sphere", "box"):
        world, hlist, cam = scenes.medium_probe(kind)
        rs = ref.RefBuilder().realize(world, hlist, cam)
        rays = medium_rays(kind, 1_000_000, 21)
        ids, t, _ = rs.trace(rays, seed=5, want_rec=False)
        sc = ids >= 0
        out[f"medium_{kind}_scatter_frac"] = float(sc.mean())
        out[f"medium_{kind}_mean_t"] = float(t[sc].astype(np.float64).mean())
        out[f"medium_{kind}_mean_t2"] = float((t[sc].astype(np.float64) ** 2).mean())
        out[f"medium_{kind}_ids"] = np.unique(ids)
        print("medium", kind, "scatter fraction", sc.mean(), "mean t", t[sc].mean(), "ids", np.unique(ids))
    np.savez_compressed(os.path.join(HERE, "features.npz"), **out)


if __name__ == "__main__":
    only = sys.argv[1:]
    if only == ["merl"]:
        merl_kat()
        sys.exit(0)
    if only == ["tessellated_10m"]:
        tessellated_10m()
        sys.exit(0)
    if only == ["features"]:
        feature_vectors()
        sys.exit(0)
    if only:
        cfg = {"cornell": (64, 64, 2, 4096, 48, 48), "spheres": (96, 64, 2, 16384, 48, 32), "teapot": (96, 96, 1, 4096, 40, 40),
               "soldier": (96, 54, 2, 4096, 48, 27), "soldier_wood": (96, 54, 2, 4096, 48, 27),
               "tessellated": (96, 54, 2, 4096, 48, 27),
               "final": (96, 96, 2, 65536, 24, 24, True), "zoo": (96, 64, 2, 4096, 48, 32, True),
               "zoo_env": (96, 64, 2, 4096, 48, 32, True)}
        for name in only:
            scene_vectors(name, *cfg[name])
        sys.exit(0)
    kats()
    merl_kat()
    scene_vectors("cornell", 64, 64, 2, 4096, 48, 48)
    scene_vectors("spheres", 96, 64, 2, 16384, 48, 32)
    scene_vectors("teapot", 96, 96, 1, 4096, 40, 40)
    scene_vectors("soldier", 96, 54, 2, 4096, 48, 27)
    scene_vectors("soldier_wood", 96, 54, 2, 4096, 48, 27)
    scene_vectors("tessellated", 96, 54, 2, 4096, 48, 27)
    scene_vectors("final", 96, 96, 2, 65536, 24, 24, True)
    scene_vectors("zoo", 96, 64, 2, 4096, 48, 32, True)
    scene_vectors("zoo_env", 96, 64, 2, 4096, 48, 32, True)
    feature_vectors()
