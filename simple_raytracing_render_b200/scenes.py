"""The benchmark / parity scenes of BASELINE.json:configs, written against the mirror of the
reference's scene API (scene.py) the way the reference's own scene functions are written
(Raytracing_n.cpp:108-711).  Every function returns the (world, light list, camera) triple a
reference scene function hands back; `oracle.ref.RefBuilder().realize(*triple)` builds the same
scene out of REAL reference objects.

C1 cornell_box_classic   512x512 @ 16     rect walls + 2 rotated boxes + diffuse_light, hitable_pdf light
C2 random_spheres        1200x800 @ 64    ~490 lambertian / metal / dielectric spheres in one bvh_node
C3 teapot_beckmann       1024x1024 @ 256  640 000-triangle teapot (teapot.h), Beckmann, sky sphere
C4 soldier_scene         1920x1080 @ 1024 Soilder.FBX mesh + sky4.jpg env map + floor with brdfmaterial (MERL);
                                          soldier_like: the same layout on stand-in assets when the cache is missing
C5 tessellated_medium    3840x2160 @ 256  10 M triangles + constant_medium
"""
from __future__ import annotations

import math

import os

import numpy as np

from . import assets
from . import scene as S
from .scene import (beckmann, box, brdfmaterial, bvh_node, camera, checker_texture, constant_medium, constant_texture, dielectric,
                    diffuse_light, env_sphere, flip_normals, hitable_list, image_texture, lambertian, metal, moving_sphere,
                    noise_texture, orennayar, rotate_x, rotate_y, sphere, teapot_bvh, translate, triangle, triangle_soup,
                    vec3, xy_rect, xz_rect, yz_rect)

CONFIG_SIZES = {
    "cornell": (512, 512, 16),
    "spheres": (1200, 800, 64),
    "teapot": (1024, 1024, 256),
    "soldier": (1920, 1080, 1024),
    "tessellated": (3840, 2160, 256),
    "soldier_wood": (1920, 1080, 1024),  # soldier_scene literally (Oren-Nayar wood floor): parity only
    # parity-only scenes (no BASELINE.json config): every reference class C1-C5 leave out
    "final": (500, 500, 64),
    "zoo": (480, 320, 64),
    "zoo_env": (480, 320, 64),
}


def cornell_box_classic(aspect: float = 1.0):
    """C1: the classic Cornell box the reference keeps commented at
    Raytracing_n.cpp:258-265, with the light sampled through hitable_pdf."""
    red = lambertian(constant_texture(vec3(0.65, 0.05, 0.05)))
    white = lambertian(constant_texture(vec3(0.73, 0.73, 0.73)))
    green = lambertian(constant_texture(vec3(0.12, 0.45, 0.15)))
    light = diffuse_light(constant_texture(vec3(15, 15, 15)))
    lst = [
        flip_normals(yz_rect(0, 555, 0, 555, 555, green)),
        yz_rect(0, 555, 0, 555, 0, red),
        flip_normals(xz_rect(213, 343, 227, 332, 554, light)),
        flip_normals(xz_rect(0, 555, 0, 555, 555, white)),
        xz_rect(0, 555, 0, 555, 0, white),
        flip_normals(xy_rect(0, 555, 0, 555, 555, white)),
        translate(rotate_y(box(vec3(0, 0, 0), vec3(165, 165, 165), white), -18), vec3(130, 0, 65)),
        translate(rotate_y(box(vec3(0, 0, 0), vec3(165, 330, 165), white), 15), vec3(265, 0, 295)),
    ]
    world = hitable_list(lst, len(lst))
    hlist = hitable_list([flip_normals(xz_rect(213, 343, 227, 332, 554, None))], 1)
    cam = camera(vec3(278, 278, -800), vec3(278, 278, 0), vec3(0, 1, 0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0)
    return world, hlist, cam


def random_spheres(aspect: float = 1.5, seed: int = 1, grid: int = 11):
    """C2: the random-spheres recipe of random_scene (Raytracing_n.cpp:117-147) with static spheres in
    one bvh_node and a sky rectangle as the sampled light."""
    S.set_seed(seed)
    checker = checker_texture(constant_texture(vec3(0.2, 0.3, 0.1)), constant_texture(vec3(0.9, 0.9, 0.9)))
    lst = [sphere(vec3(0, -1000, 0), 1000, lambertian(checker))]
    for a in range(-grid, grid):
        for b in range(-grid, grid):
            choose_mat = np.float32(S.drand48())
            cx = np.float32(a + 0.9 * S.drand48())
            cz = np.float32(b + S.drand48())
            center = vec3(cx, 0.2, cz)
            if float(S._length(center - vec3(4, 0.2, 0))) > 0.9:
                if choose_mat < 0.8:
                    alb = vec3(S.drand48() * S.drand48(), S.drand48() * S.drand48(), S.drand48() * S.drand48())
                    lst.append(sphere(center, 0.2, lambertian(constant_texture(alb))))
                elif choose_mat < 0.95:
                    alb = vec3(0.5 * (1 + S.drand48()), 0.5 * (1 + S.drand48()), 0.5 * (1 + S.drand48()))
                    lst.append(sphere(center, 0.2, metal(alb, 0.5 * S.drand48())))
                else:
                    lst.append(sphere(center, 0.2, dielectric(1.5)))
    lst.append(sphere(vec3(0, 1, 0), 1.0, dielectric(1.5)))
    lst.append(sphere(vec3(-4, 1, 0), 1.0, lambertian(constant_texture(vec3(0.4, 0.2, 0.1)))))
    lst.append(sphere(vec3(4, 1, 0), 1.0, metal(vec3(0.7, 0.6, 0.5), 0.0)))
    sky = diffuse_light(constant_texture(vec3(3, 3, 3)))
    world = hitable_list([bvh_node(lst, len(lst), 0.0, 1.0),
                          flip_normals(xz_rect(-50, 50, -50, 50, 60, sky))], 2)
    hlist = hitable_list([xz_rect(-50, 50, -50, 50, 60, None)], 1)
    cam = camera(vec3(13, 2, 3), vec3(0, 0, 0), vec3(0, 1, 0), 20.0, aspect, 0.1, 10.0, 0.0, 1.0)
    return world, hlist, cam


def synthetic_sky(nx: int = 1024, ny: int = 512) -> np.ndarray:
    """Stand-in for contents/environment_map/sky_2.png (2000x1000 RGB8; the asset cannot travel to the
    GPU box): a deterministic RGB8 sky — vertical gradient, a soft sun lobe and low-frequency bands — so
    that image_texture::value (texture.h:58-70) has real texel variation to fetch."""
    v = (np.arange(ny, dtype=np.float64) + 0.5) / ny  # row 0 = top of the image = v -> 1
    u = (np.arange(nx, dtype=np.float64) + 0.5) / nx
    up = 1.0 - v[:, None]
    base = np.stack([0.35 + 0.45 * (1 - up) + 0 * u[None, :], 0.55 + 0.30 * (1 - up) + 0 * u[None, :],
                     0.95 - 0.15 * (1 - up) + 0 * u[None, :]], -1)
    sun = np.exp(-(((u[None, :] - 0.3) * 2.0) ** 2 + ((v[:, None] - 0.25) * 4.0) ** 2) * 18.0)
    bands = 0.06 * np.sin(u[None, :] * 2 * np.pi * 5.0) * np.sin(v[:, None] * np.pi * 3.0)
    img = base + sun[..., None] * np.array([0.6, 0.5, 0.3]) + bands[..., None]
    ground = v[:, None] > 0.55
    img = np.where(ground[..., None], img * np.array([0.45, 0.42, 0.38]), img)
    return np.clip(img * 255.0 + 0.5, 0, 255).astype(np.uint8)


def teapot_beckmann(aspect: float = 1.0, divs: int = 100, seed: int = 1):
    """C3: teapot_scene (Raytracing_n.cpp:306-377) reduced to what BASELINE.json names —
    the procedural Utah teapot (teapot.h, 32 x divs x divs x 2 triangles; 640 000 at divs = 100) with the
    reference's gold Beckmann material in a bvh_node under rotate_x / translate, the lambertian floor, the
    sampled flipped xz_rect light and the image-textured sky sphere around the camera."""
    S.set_seed(seed)
    lookfrom = vec3(100, 800, -400)
    gold = beckmann(constant_texture(vec3(0.945, 0.75, 0.336)), 0.01, 0.05)
    brown = lambertian(constant_texture(vec3(0.426, 0.3, 0.254)))
    light = diffuse_light(constant_texture(vec3(40, 40, 40)))
    # contents/environment_map/sky_2.png (2000 x 1000), decoded by the reference's stb_image (tools/make_assets.py)
    sky_img = assets.load_cached("sky_2.npy") if assets.have_cached("sky_2.npy") else synthetic_sky()
    sky = diffuse_light(image_texture(sky_img))
    lst = [
        flip_normals(xz_rect(3, 153, 217, 343, 800, light)),
        xz_rect(0, 555, 0, 555, 0, brown),
        flip_normals(sphere(lookfrom, 10000, sky)),
        translate(rotate_x(teapot_bvh(40, gold, divs), 90), vec3(200, 0, 250)),
    ]
    world = hitable_list(lst, len(lst))
    hlist = hitable_list([flip_normals(xz_rect(3, 153, 217, 343, 800, None))], 1)
    cam = camera(lookfrom, vec3(300, 278, 200), vec3(0, 1, 0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0)
    return world, hlist, cam


def constant_merl_table(albedo: float = 0.5) -> np.ndarray:
    """A MERL-layout table whose looked-up value (brdf.h:190-214, channel scales 1/1500, 1.15/1500, 1.66/1500) is
    albedo / pi for every angle: a brdfmaterial with it IS a lambertian of that albedo wherever the table value is
    applied (corrected-transport mode)."""
    n = 90 * 90 * 180
    f = albedo / np.pi
    return np.concatenate([np.full(n, f * 1500.0), np.full(n, f * 1500.0 / 1.15), np.full(n, f * 1500.0 / 1.66)]).astype(np.float64)


def synthetic_merl_table() -> np.ndarray:
    """Stand-in for contents/brdfs/aluminium.binary (absent from the reference tree,
    .MISSING_LARGE_BLOBS): a smooth metallic lobe in the MERL layout of brdf.h:171-184 —
    [channel 3][theta_half 90][theta_diff 90][phi_diff 180] doubles, pre-scale (brdf.h:12-14)."""
    th = (np.arange(90, dtype=np.float64) + 0.5) ** 2 / 90.0 / 90.0 * (np.pi / 2)  # theta_half_index is sqrt-spaced
    td = (np.arange(90, dtype=np.float64) + 0.5) / 90.0 * (np.pi / 2)
    pd = (np.arange(180, dtype=np.float64) + 0.5) / 180.0 * np.pi
    lobe = np.exp(-(th[:, None, None] / 0.12) ** 2) * 40.0 + 0.35
    fres = 1.0 + 2.5 * (1.0 - np.cos(td))[None, :, None] ** 5
    aniso = 1.0 + 0.05 * np.cos(2 * pd)[None, None, :]
    base = lobe * fres * aniso
    tab = np.stack([base * 1500.0 / 1.0 * 0.91, base * 1500.0 / 1.15 * 0.92, base * 1500.0 / 1.66 * 0.93])
    return np.ascontiguousarray(tab.reshape(-1), np.float64)


def _ellipsoid(center, radii, nu, nv, uv_lo, uv_hi):
    """UV-sphere tessellation: positions [n][9], smooth normals [n][9], uvs [n][6]."""
    cx, cy, cz = center
    rx, ry, rz = radii
    u = np.linspace(0.0, 2 * np.pi, nu + 1)
    v = np.linspace(0.0, np.pi, nv + 1)
    uu, vv = np.meshgrid(u, v, indexing="ij")
    n = np.stack([np.cos(uu) * np.sin(vv), np.cos(vv), np.sin(uu) * np.sin(vv)], -1)
    p = n * np.array([rx, ry, rz]) + np.array([cx, cy, cz])
    nn = n / np.array([rx, ry, rz])
    nn = nn / np.maximum(np.linalg.norm(nn, axis=-1, keepdims=True), 1e-20)
    tu = uv_lo[0] + (uv_hi[0] - uv_lo[0]) * (uu / (2 * np.pi))
    tv = uv_lo[1] + (uv_hi[1] - uv_lo[1]) * (vv / np.pi)
    t = np.stack([tu, tv], -1)
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    i, j = i.reshape(-1), j.reshape(-1)
    quads = [(i, j), (i + 1, j), (i + 1, j + 1), (i, j + 1)]
    tris = [(0, 1, 2), (0, 2, 3)]
    pos, nrm, uv = [], [], []
    for a, b, c in tris:
        pos.append(np.concatenate([p[quads[k]] for k in (a, b, c)], -1))
        nrm.append(np.concatenate([nn[quads[k]] for k in (a, b, c)], -1))
        uv.append(np.concatenate([t[quads[k]] for k in (a, b, c)], -1))
    pos, nrm, uv = np.concatenate(pos), np.concatenate(nrm), np.concatenate(uv)
    # drop the degenerate pole triangles (zero area: the reference's det test rejects them anyway)
    e1, e2 = pos[:, 3:6] - pos[:, 0:3], pos[:, 6:9] - pos[:, 0:3]
    keep = np.linalg.norm(np.cross(e1, e2), axis=1) > 1e-9
    return pos[keep].astype(np.float32), nrm[keep].astype(np.float32), uv[keep].astype(np.float32)


def figure_mesh(detail: int = 12):
    """Stand-in for mesh[0] of contents/models/Soilder.FBX (~3.3 k triangles; assimp is a Windows binary,
    SURVEY.md §8c): a standing figure of tessellated ellipsoids with smooth normals and atlas uvs."""
    parts = [  # centre, radii, uv window
        ((0, 62, 0), (5.0, 6.0, 5.0), (0.0, 0.0), (0.5, 0.25)),      # head
        ((0, 42, 0), (9.0, 14.0, 6.0), (0.0, 0.25), (1.0, 0.75)),    # torso
        ((-12, 44, 0), (3.0, 13.0, 3.0), (0.5, 0.0), (0.75, 0.25)),  # arms
        ((12, 44, 0), (3.0, 13.0, 3.0), (0.75, 0.0), (1.0, 0.25)),
        ((-5, 14, 0), (4.0, 15.0, 4.0), (0.0, 0.75), (0.5, 1.0)),    # legs
        ((5, 14, 0), (4.0, 15.0, 4.0), (0.5, 0.75), (1.0, 1.0)),
    ]
    out = [_ellipsoid(c, r, 2 * detail, detail, lo, hi) for c, r, lo, hi in parts]
    return tuple(np.concatenate([o[k] for o in out]) for k in range(3))


def synthetic_albedo(n: int = 512) -> np.ndarray:
    """Stand-in for contents/textures/NPC_YuanChengBing_A.png (1024x1024 RGB8)."""
    y, x = np.mgrid[0:n, 0:n].astype(np.float64) / n
    r = 0.45 + 0.35 * np.sin(x * 17.0) * np.cos(y * 11.0)
    g = 0.40 + 0.30 * np.cos(x * 7.0 + y * 13.0)
    b = 0.35 + 0.25 * np.sin((x + y) * 23.0)
    return np.clip(np.stack([r, g, b], -1) * 255.0 + 0.5, 0, 255).astype(np.uint8)


def soldier_like(aspect: float = 16.0 / 9.0, detail: int = 12, seed: int = 1):
    """C4: soldier_scene (Raytracing_n.cpp:585-657) — textured Beckmann mesh in a
    bvh_node under rotate_y / translate, sky-sphere environment map, glass slab, sampled xz_rect light —
    with the floor given the MERL measured-BRDF material (brdfmaterial) as BASELINE.json:configs[3]
    asks.  Assets that cannot travel are deterministic stand-ins of the same kind and layout."""
    S.set_seed(seed)
    lookfrom = vec3(300, 500, -800)
    light_1 = diffuse_light(constant_texture(vec3(35, 35, 35)))
    floor = brdfmaterial(synthetic_merl_table(), vec3(0.7, 0.7, 0.72))
    glass = dielectric(1.4)
    sky = diffuse_light(image_texture(synthetic_sky(2048, 1024)))
    skin = beckmann(image_texture(synthetic_albedo()), 0.9, 0.85)
    pos, nrm, uv = figure_mesh(detail)
    pos = (pos * np.float32(8.0)).astype(np.float32)  # model(..., vec3(8, 8, 8))
    lst = [
        flip_normals(xz_rect(203, 353, 17, 167, 800, light_1)),
        box(vec3(0, -0.1, 0), vec3(600, 0.1, 600), floor),
        box(vec3(0, -1, 0), vec3(600, 1, 600), glass),
        flip_normals(sphere(lookfrom, 10000, sky)),
        translate(rotate_y(triangle_soup(pos, skin, uv=uv, nrm=nrm), 180), vec3(250, 0, 300)),
    ]
    world = hitable_list(lst, len(lst))
    hlist = hitable_list([flip_normals(xz_rect(203, 353, 17, 167, 800, None))], 1)
    cam = camera(lookfrom, vec3(300, 278, 200), vec3(0, 1, 0), 40.0, aspect, 10.0, 1000.0, 0.0, 1.0)
    return world, hlist, cam


SOLDIER_ASSETS = ("soldier_mesh.npz", "sky4.npy", "soldier_albedo.npy", "wood_albedo.npy")


def real_assets() -> bool:
    """True when tools/make_assets.py has filled assets_cache/: the mesh / image scenes then use the reference's own
    blobs (Soilder.FBX, sky4.jpg, sky_2.png, the albedo PNGs); otherwise deterministic stand-ins of the same kind."""
    return assets.have_cached(*SOLDIER_ASSETS, "sky_2.npy")


def soldier_scene(aspect: float = 16.0 / 9.0, floor: str = "merl", seed: int = 1):
    """C4: soldier_scene (Raytracing_n.cpp:585-657) on the reference's own assets — mesh[0] of Soilder.FBX
    (FlipWindingOrder, scale 8: model.h:26-57, geometry.h:61-77) with beckmann(image_texture(NPC_YuanChengBing_A.png),
    0.9, 0.85) in a bvh_node under rotate_y 180 / translate, the sky4.jpg sky sphere around the camera, the glass
    slab, the sampled flipped xz_rect light, thin-lens camera with aperture 10.
    floor = "merl": the floor box carries the MERL measured-BRDF material (brdfmaterial) BASELINE.json:configs[3]
    names — the MERL binaries are not in the reference tree, so the table is synthetic in the MERL layout;
    floor = "wood": the scene function literally — orennayar(image_texture(Wood_Wenge albedo), 0.5), :620."""
    S.set_seed(seed)
    lookfrom = vec3(300, 500, -800)
    light_1 = diffuse_light(constant_texture(vec3(35, 35, 35)))
    if floor == "wood":
        floor_mat = orennayar(image_texture(assets.load_cached("wood_albedo.npy")), 0.5)
    else:
        floor_mat = brdfmaterial(synthetic_merl_table(), vec3(0.7, 0.7, 0.72))
    glass = dielectric(1.4)
    sky = diffuse_light(image_texture(assets.load_cached("sky4.npy")))
    skin = beckmann(image_texture(assets.load_cached("soldier_albedo.npy")), 0.9, 0.85)
    mesh = np.load(assets.cache_path("soldier_mesh.npz"))
    pos = (mesh["pos"] * np.float32(8.0)).astype(np.float32)  # geometry.h:66: p * scale, in float
    lst = [
        flip_normals(xz_rect(203, 353, 17, 167, 800, light_1)),
        box(vec3(0, -0.1, 0), vec3(600, 0.1, 600), floor_mat),
        box(vec3(0, -1, 0), vec3(600, 1, 600), glass),
        flip_normals(sphere(lookfrom, 10000, sky)),
        translate(rotate_y(triangle_soup(pos, skin, uv=mesh["uv"], nrm=mesh["nrm"]), 180), vec3(250, 0, 300)),
    ]
    world = hitable_list(lst, len(lst))
    hlist = hitable_list([flip_normals(xz_rect(203, 353, 17, 167, 800, None))], 1)
    cam = camera(lookfrom, vec3(300, 278, 200), vec3(0, 1, 0), 40.0, aspect, 10.0, 1000.0, 0.0, 1.0)
    return world, hlist, cam


def soldier(aspect: float = 16.0 / 9.0, **kw):
    """C4 on the reference's assets when they are cached, on the stand-ins otherwise (real_assets())."""
    if assets.have_cached(*SOLDIER_ASSETS):
        return soldier_scene(aspect, **kw)
    kw.pop("floor", None)
    return soldier_like(aspect, **kw)


def soldier_wood(aspect: float = 16.0 / 9.0, **kw):
    return soldier_scene(aspect, floor="wood", **kw)


def terrain_triangles(n_side: int) -> np.ndarray:
    """2 * n_side^2 triangles of a tessellated height field over [0, 555]^2 (positions [n][9])."""
    g = np.linspace(0.0, 555.0, n_side + 1)
    x, z = np.meshgrid(g, g, indexing="ij")
    y = (40.0 * np.sin(x * 0.021) * np.cos(z * 0.017) + 12.0 * np.sin(x * 0.13 + 1.0) * np.sin(z * 0.11)
         + 3.0 * np.sin(x * 0.9) * np.cos(z * 0.7) + 60.0)
    p = np.stack([x, y, z], -1).astype(np.float32)
    a, b, c, d = p[:-1, :-1], p[1:, :-1], p[1:, 1:], p[:-1, 1:]
    t0 = np.concatenate([a, c, b], -1).reshape(-1, 9)  # counter-clockwise seen from +y
    t1 = np.concatenate([a, d, c], -1).reshape(-1, 9)
    return np.concatenate([t0, t1]).astype(np.float32)


def tessellated_medium(aspect: float = 16.0 / 9.0, n_side: int = 512, seed: int = 1, fog: bool = True):
    """C5: a tessellated triangle scene (2 * n_side^2 triangles in one bvh_node;
    n_side = 2237 gives the 10 M of BASELINE.json:configs[4]) plus a constant_medium volume, lit by a
    sampled xz_rect light and a constant sky."""
    S.set_seed(seed)
    ground = lambertian(constant_texture(vec3(0.48, 0.63, 0.36)))
    light = diffuse_light(constant_texture(vec3(25, 25, 25)))
    sky = diffuse_light(constant_texture(vec3(0.35, 0.45, 0.7)))
    lookfrom = vec3(278, 420, -600)
    fog_boundary = sphere(vec3(300, 150, 280), 110, dielectric(1.5))
    lst = [
        flip_normals(xz_rect(123, 423, 147, 412, 700, light)),
        flip_normals(sphere(lookfrom, 10000, sky)),
        triangle_soup(terrain_triangles(n_side), ground),
        fog_boundary,
    ]
    if fog:  # constant_medium::hit draws its free-flight distance at random: closest-hit parity is checked without it
        lst.append(constant_medium(fog_boundary, 0.01, constant_texture(vec3(0.9, 0.9, 0.95))))
    world = hitable_list(lst, len(lst))
    hlist = hitable_list([flip_normals(xz_rect(123, 423, 147, 412, 700, None))], 1)
    cam = camera(lookfrom, vec3(278, 120, 278), vec3(0, 1, 0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0)
    return world, hlist, cam


def synthetic_earth(nx: int = 512, ny: int = 256) -> np.ndarray:
    """Stand-in for contents/textures/earthmap.jpg (cannot travel to the GPU box): continents of
    low-frequency noise over an ocean gradient, RGB8."""
    v, u = np.mgrid[0:ny, 0:nx].astype(np.float64)
    u, v = u / nx, v / ny
    h = (np.sin(u * 2 * np.pi * 3.0 + 1.3) * np.cos(v * np.pi * 4.0) + 0.6 * np.sin(u * 2 * np.pi * 7.0 + v * 9.0)
         + 0.3 * np.cos(u * 2 * np.pi * 13.0 - v * 17.0))
    land = h > 0.25
    ocean = np.stack([0.05 + 0.1 * v, 0.2 + 0.2 * v, 0.5 + 0.3 * (1 - v)], -1)
    ground = np.stack([0.2 + 0.3 * np.clip(h, 0, 1), 0.45 - 0.15 * np.clip(h, 0, 1), 0.12 + 0 * h], -1)
    img = np.where(land[..., None], ground, ocean)
    ice = (v < 0.08) | (v > 0.93)
    img = np.where(ice[..., None], 0.92, img)
    return np.clip(img * 255.0 + 0.5, 0, 255).astype(np.uint8)


def book_final(aspect: float = 1.0, seed: int = 1, nb: int = 20, ns: int = 1000, media: bool = True):
    """The reference's own `final` scene function (Raytracing_n.cpp:475-533), line by line: a bvh_node of
    `box` leaves (a generic bvh: every leaf is a hitable_list of six rects), a moving_sphere, glass and
    fuzzy-metal spheres, two constant_medium volumes (one inside a glass boundary, one around the whole
    scene), an image-textured and a noise-textured sphere, and a rotated / translated bvh_node of 1000
    spheres.  earthmap.jpg is replaced by a synthetic RGB8 map (the asset cannot travel)."""
    S.set_seed(seed)
    white = lambertian(constant_texture(vec3(0.73, 0.73, 0.73)))
    ground = lambertian(constant_texture(vec3(0.48, 0.83, 0.53)))
    boxlist = []
    f32 = np.float32
    for i in range(nb):
        for j in range(nb):
            w = f32(100)
            x0 = f32(-1000 + i * w)
            z0 = f32(-1000 + j * w)
            y0 = f32(0)
            x1 = f32(x0 + w)
            y1 = f32(100 * (S.drand48() + 0.01))
            z1 = f32(z0 + w)
            boxlist.append(box(vec3(x0, y0, z0), vec3(x1, y1, z1), ground))
    lst = [bvh_node(boxlist, len(boxlist), 0, 1)]
    light = diffuse_light(constant_texture(vec3(7, 7, 7)))
    lst.append(flip_normals(xz_rect(123, 423, 147, 412, 554, light)))
    center = vec3(400, 400, 200)
    lst.append(moving_sphere(center, center + vec3(30, 0, 0), 0, 1, 50,
                             lambertian(constant_texture(vec3(0.7, 0.3, 0.1)))))
    lst.append(sphere(vec3(260, 150, 45), 50, dielectric(1.5)))
    lst.append(sphere(vec3(0, 150, 145), 50, metal(vec3(0.8, 0.8, 0.9), 1.0)))
    boundary = sphere(vec3(360, 150, 145), 70, dielectric(1.5))
    lst.append(boundary)
    if media:  # constant_medium::hit is random: closest-hit parity is checked on the twin without the volumes
        lst.append(constant_medium(boundary, 0.2, constant_texture(vec3(0.2, 0.4, 0.9))))
        boundary = sphere(vec3(0, 0, 0), 5000, dielectric(1.5))
        lst.append(constant_medium(boundary, 0.0001, constant_texture(vec3(1.0, 1.0, 1.0))))
    emat = lambertian(image_texture(synthetic_earth()))
    lst.append(sphere(vec3(400, 200, 400), 100, emat))
    pertext = noise_texture(0.1)
    lst.append(sphere(vec3(220, 280, 300), 80, lambertian(pertext)))
    boxlist2 = [sphere(vec3(165 * S.drand48(), 165 * S.drand48(), 165 * S.drand48()), 10, white) for _ in range(ns)]
    lst.append(translate(rotate_y(bvh_node(boxlist2, ns, 0.0, 1.0), 15), vec3(-100, 270, 395)))
    world = hitable_list(lst, len(lst))
    cam = camera(vec3(478, 278, -600), vec3(278, 278, 0), vec3(0, 1, 0), 40.0, aspect, 0.0, 10.0, 0.0, 1.0)
    hlist = hitable_list([flip_normals(xz_rect(123, 423, 147, 412, 554, None))], 1)
    return world, hlist, cam


def feature_zoo(aspect: float = 1.5, seed: int = 1, media: bool = True):
    """Parity scene for the reference classes no BASELINE.json config instantiates: orennayar +
    onrennayar_pdf (material.h:127-149, pdf.h:61-116), moving_sphere alone and inside a generic bvh_node
    (moving_sphere.h:24-58), noise_texture + perlin (texture.h:35-46, perlin.h:7-58), a sphere and a
    triangle in the light list (sphere.h:69-86, triangle.h:70-94), triangles inside a hitable_list (they
    ignore t_max and overwrite nearer hits, triangle.h:117-187), a checker over an image texture, and a
    constant_medium inside a `box` boundary (constant_medium.h:19-50)."""
    S.set_seed(seed)
    noise = lambertian(noise_texture(1.5))
    oren = orennayar(constant_texture(vec3(0.8, 0.45, 0.3)), 25.0)
    oren_noise = orennayar(noise_texture(4.0), 60.0)
    mover = lambertian(constant_texture(vec3(0.2, 0.4, 0.8)))
    lamp = diffuse_light(constant_texture(vec3(9, 9, 8)))
    lamp2 = diffuse_light(constant_texture(vec3(6, 7, 9)))
    chk = lambertian(checker_texture(image_texture(synthetic_albedo(128)), constant_texture(vec3(0.9, 0.9, 0.9))))
    # triangle light: counter-clockwise seen from below, so that its normal points down (emission is
    # one-sided, material.h:348-354)
    t_light = (vec3(-2.5, 4.5, 2.0), vec3(0.5, 4.5, -1.5), vec3(2.5, 4.5, 2.0))
    s_light = (vec3(3.2, 2.6, -1.0), 0.55)
    small = []
    for k in range(24):  # a generic bvh: spheres, moving spheres and boxes as leaves
        cx, cz = -4.5 + 9.0 * S.drand48(), 1.0 + 3.5 * S.drand48()
        r = 0.12 + 0.18 * S.drand48()
        m = [oren, mover, chk, metal(vec3(0.8, 0.8, 0.7), 0.3)][k % 4]
        if k % 3 == 0:
            small.append(moving_sphere(vec3(cx, r, cz), vec3(cx, r + 0.4, cz), 0.0, 1.0, r, m))
        elif k % 3 == 1:
            small.append(sphere(vec3(cx, r, cz), r, m))
        else:
            small.append(box(vec3(cx - r, 0, cz - r), vec3(cx + r, 2 * r, cz + r), m))
    lst = [
        triangle(*t_light, lamp),
        xz_rect(-8, 8, -6, 8, 0, noise),
        flip_normals(xy_rect(-8, 8, 0, 8, 8, chk)),
        sphere(vec3(0, 1, 0), 1.0, oren),
        sphere(vec3(-2.2, 0.8, -1.2), 0.8, oren_noise),
        moving_sphere(vec3(2.0, 0.6, -1.5), vec3(2.0, 1.2, -1.5), 0.0, 1.0, 0.6, mover),
        sphere(s_light[0], s_light[1], lamp2),
        bvh_node(small, len(small), 0.0, 1.0),
        triangle(vec3(-5, 0.2, 3), vec3(-4, 2.2, 3), vec3(-3, 0.2, 3), oren),  # late in the list: seen through everything
    ]
    if media:
        fog = box(vec3(-1.5, 0.0, -3.5), vec3(1.0, 1.2, -2.5), dielectric(1.5))
        lst.append(constant_medium(fog, 0.8, constant_texture(vec3(0.9, 0.8, 0.7))))
    world = hitable_list(lst, len(lst))
    hlist = hitable_list([sphere(s_light[0], s_light[1], None), triangle(*t_light, None)], 2)
    cam = camera(vec3(0, 3.0, -10), vec3(0, 1.2, 0), vec3(0, 1, 0), 38.0, aspect, 0.05, 10.0, 0.0, 1.0)
    return world, hlist, cam


def zoo_env(aspect: float = 1.5, seed: int = 1):
    """Parity scene for env_sphere as a world object and as a light (env_sphere.h:27-50): its hit() ignores
    (t_min, t_max) and always reports a hit, so it stands FIRST in the world list (anything later that is
    nearer replaces it); in the light list its pdf_value is 1 whatever the direction, which ends the
    `while (pdf_val == 0)` loop at once, and its random() points away from the centre (NaN from inside).
    Also a light-list element without pdf_value / random (an xy_rect: hitable.h:31-32)."""
    S.set_seed(seed)
    sky = diffuse_light(image_texture(synthetic_sky(256, 128)))
    oren = orennayar(constant_texture(vec3(0.7, 0.7, 0.6)), 40.0)
    lamp = diffuse_light(constant_texture(vec3(12, 12, 12)))
    lst = [
        env_sphere(vec3(0, 1, 0), 60.0, sky),
        xz_rect(-12, 12, -12, 12, 0, oren),
        sphere(vec3(-1.6, 1, 0), 1.0, metal(vec3(0.9, 0.8, 0.6), 0.05)),
        sphere(vec3(1.6, 1, 0), 1.0, lambertian(constant_texture(vec3(0.3, 0.6, 0.4)))),
        flip_normals(xz_rect(-2, 2, -2, 2, 6, lamp)),
    ]
    world = hitable_list(lst, len(lst))
    hlist = hitable_list([flip_normals(xz_rect(-2, 2, -2, 2, 6, None)), env_sphere(vec3(0, 1, 0), 60.0, None),
                          xy_rect(-1, 1, 0, 2, 5, None)], 3)
    cam = camera(vec3(0, 2.5, -9), vec3(0, 1, 0), vec3(0, 1, 0), 35.0, aspect, 0.0, 10.0, 0.0, 0.0)
    return world, hlist, cam


def texture_probe(kind: str = "noise", aspect: float = 1.0):
    """One emissive surface filling the view: every pixel's radiance is texture::value at the primary hit
    (diffuse_light::emitted, material.h:348-354), so an image compares the texture arithmetic directly —
    noise_texture + perlin (texture.h:35-46, perlin.h:7-58), checker_texture (texture.h:9-23) and
    image_texture through rect (aarect.h) and sphere uv (hitable.h:10-15, texture.h:58-70)."""
    if kind == "noise":
        tex = noise_texture(2.0)
    elif kind == "checker":
        tex = checker_texture(noise_texture(0.7), constant_texture(vec3(0.9, 0.2, 0.1)))
    elif kind in ("image", "image_sphere"):
        tex = image_texture(synthetic_albedo(64))
    else:
        raise KeyError(kind)
    lamp = diffuse_light(tex)
    if kind == "image_sphere":
        lst = [sphere(vec3(0, 0, 0), 2.0, lamp)]
        cam = camera(vec3(0, 1.5, -6), vec3(0, 0, 0), vec3(0, 1, 0), 45.0, aspect, 0.0, 10.0, 0.0, 0.0)
    else:
        lst = [xz_rect(-5, 5, -5, 5, 0, lamp)]
        cam = camera(vec3(0.3, 7, -2), vec3(0, 0, 0), vec3(0, 1, 0), 50.0, aspect, 0.0, 10.0, 0.0, 0.0)
    return hitable_list(lst, len(lst)), hitable_list([], 0), cam


def oren_only(aspect: float = 1.0, sigma: float = 45.0):
    """orennayar + onrennayar_pdf (material.h:127-149, pdf.h:61-116) and nothing else that scatters: a rough
    floor and a rough sphere under one sampled rect light."""
    oren = orennayar(constant_texture(vec3(0.75, 0.6, 0.5)), sigma)
    lamp = diffuse_light(constant_texture(vec3(10, 10, 10)))
    lst = [xz_rect(-6, 6, -6, 6, 0, oren), sphere(vec3(0, 1, 0), 1.0, oren),
           flip_normals(xz_rect(-1.5, 1.5, -1.5, 1.5, 5, lamp))]
    hlist = hitable_list([flip_normals(xz_rect(-1.5, 1.5, -1.5, 1.5, 5, None))], 1)
    cam = camera(vec3(0, 3, -7), vec3(0, 0.8, 0), vec3(0, 1, 0), 40.0, aspect, 0.0, 10.0, 0.0, 0.0)
    return hitable_list(lst, len(lst)), hlist, cam


def medium_probe(boundary_kind: str = "sphere", density: float = 0.7):
    """A constant_medium and nothing else (constant_medium.h:19-50): rays either scatter inside it (id =
    n_prims + 0) or miss.  Boundaries: a sphere, or a `box` (hitable_list of six flipped / unflipped rects)."""
    glass = dielectric(1.5)
    if boundary_kind == "sphere":
        b = sphere(vec3(0, 0, 0), 1.0, glass)
    else:
        b = box(vec3(-1.0, -0.6, -0.8), vec3(1.0, 0.6, 0.8), glass)
    world = hitable_list([constant_medium(b, density, constant_texture(vec3(1, 1, 1)))], 1)
    cam = camera(vec3(0, 0, -5), vec3(0, 0, 0), vec3(0, 1, 0), 40.0, 1.0, 0.0, 10.0, 0.0, 0.0)
    return world, hitable_list([], 0), cam


def furnace(aspect: float = 1.0, sampled: bool = True, albedo: float = 0.5):
    """White furnace for the corrected-transport mode (SRT_RENDER_CORRECTED): a closed box of six one-sided emitters
    of radiance 1 around a lambertian sphere.  Every direction from the (convex) sphere sees radiance 1, so the
    sphere's pixels converge to exactly `albedo` and every other pixel to 1 — with the floor and ceiling in the
    light list (next-event estimation + MIS) and without (BSDF sampling alone)."""
    one = diffuse_light(constant_texture(vec3(1, 1, 1)))
    a = 10.0
    lst = [
        flip_normals(yz_rect(-a, a, -a, a, a, one)), yz_rect(-a, a, -a, a, -a, one),
        flip_normals(xz_rect(-a, a, -a, a, a, one)), xz_rect(-a, a, -a, a, -a, one),
        flip_normals(xy_rect(-a, a, -a, a, a, one)), xy_rect(-a, a, -a, a, -a, one),
        sphere(vec3(0, 0, 0), 2.0, lambertian(constant_texture(vec3(albedo, albedo, albedo)))),
    ]
    world = hitable_list(lst, len(lst))
    lights = [flip_normals(xz_rect(-a, a, -a, a, a, None)), xz_rect(-a, a, -a, a, -a, None)] if sampled else []
    cam = camera(vec3(0, 0, -8), vec3(0, 0, 0), vec3(0, 1, 0), 40.0, aspect, 0.0, 8.0, 0.0, 0.0)
    return world, hitable_list(lights, len(lights)), cam


def corrected_probe(aspect: float = 1.0, sampled: bool = True, seed: int = 1, lights=("rect", "sphere", "tri"),
                    materials=("merl", "oren", "beckmann", "glass"), merl_table=None):
    """A Cornell-like room for the corrected-transport mode with every non-specular material — lambertian walls,
    an Oren-Nayar sphere, a Beckmann sphere, a floor with the MERL measured-BRDF material — a glass sphere, and
    three kinds of light shape (rect, sphere, triangle).  `sampled`: the lights are in the light list (next-event
    estimation + MIS) or not (BSDF sampling alone); both estimate the same image."""
    S.set_seed(seed)
    red = lambertian(constant_texture(vec3(0.65, 0.05, 0.05)))
    white = lambertian(constant_texture(vec3(0.73, 0.73, 0.73)))
    green = lambertian(constant_texture(vec3(0.12, 0.45, 0.15)))
    light = diffuse_light(constant_texture(vec3(12, 12, 12)))
    glow = diffuse_light(constant_texture(vec3(6, 5, 3)))
    floor = white
    if "merl" in materials:
        floor = brdfmaterial(synthetic_merl_table() if merl_table is None else merl_table, vec3(0.7, 0.7, 0.7))
    elif "floor50" in materials:
        floor = lambertian(constant_texture(vec3(0.5, 0.5, 0.5)))
    oren = orennayar(constant_texture(vec3(0.3, 0.45, 0.8)), 40.0) if "oren" in materials else white
    gold = beckmann(constant_texture(vec3(0.945, 0.75, 0.336)), 0.3, 0.15) if "beckmann" in materials else white
    glass = dielectric(1.5) if "glass" in materials else white
    t0, t1, t2 = vec3(40, 300, 554), vec3(120, 450, 554), vec3(200, 300, 554)  # normal -z: faces the room
    lst = [
        flip_normals(yz_rect(0, 555, 0, 555, 555, green)), yz_rect(0, 555, 0, 555, 0, red),
        flip_normals(xz_rect(0, 555, 0, 555, 555, white)), xz_rect(0, 555, 0, 555, 0, floor),
        flip_normals(xy_rect(0, 555, 0, 555, 555, white)),
        sphere(vec3(150, 90, 190), 90, oren),
        sphere(vec3(390, 110, 350), 110, gold),
        sphere(vec3(270, 60, 100), 60, glass),
    ]
    shapes = []
    if "rect" in lights:
        lst.append(flip_normals(xz_rect(163, 393, 177, 382, 554, light)))
        shapes.append(flip_normals(xz_rect(163, 393, 177, 382, 554, None)))
    if "sphere" in lights:
        lst.append(sphere(vec3(400, 60, 150), 25, glow))
        shapes.append(sphere(vec3(400, 60, 150), 25, None))
    if "tri" in lights:
        lst.append(triangle(t0, t1, t2, glow))
        shapes.append(triangle(t0, t1, t2, None))
    world = hitable_list(lst, len(lst))
    lights = shapes if sampled else []
    cam = camera(vec3(278, 278, -800), vec3(278, 278, 0), vec3(0, 1, 0), 40.0, aspect, 0.0, 10.0, 0.0, 0.0)
    return world, hitable_list(lights, len(lights)), cam


def by_name(name: str, **kw):
    fn = {"cornell": cornell_box_classic, "spheres": random_spheres, "teapot": teapot_beckmann,
          "soldier": soldier, "soldier_wood": soldier_wood, "tessellated": tessellated_medium, "final": book_final, "zoo": feature_zoo,
          "zoo_env": zoo_env}.get(name)
    if fn is None:
        raise KeyError(name)
    nx, ny, _ = CONFIG_SIZES[name]
    if name == "tessellated" and "n_side" not in kw and os.environ.get("SRT_TESS_NSIDE"):
        kw["n_side"] = int(os.environ["SRT_TESS_NSIDE"])  # 2237: the 10 M triangles of BASELINE.json:configs[4]
    return fn(aspect=float(np.float32(nx) / np.float32(ny)), **kw)
