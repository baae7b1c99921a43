"""The benchmark / parity scenes of BASELINE.json:configs, written against the mirror of the
reference's scene API (scene.py) the way the reference's own scene functions are written
(Raytracing_n.cpp:108-711).  Every function returns the (world, light list, camera) triple a
reference scene function hands back; `oracle.ref.RefBuilder().realize(*triple)` builds the same
scene out of REAL reference objects.

C1 cornell_box_classic   512x512 @ 16     rect walls + 2 rotated boxes + diffuse_light, hitable_pdf light
C2 random_spheres        1200x800 @ 64    ~490 lambertian / metal / dielectric spheres in one bvh_node
C3 teapot_beckmann       1024x1024 @ 256  640 000-triangle teapot (teapot.h), Beckmann, sky sphere
C4 soldier_like          1920x1080 @ 1024 mesh + sky4-like env map + floor with brdfmaterial (MERL)
C5 tessellated_medium    3840x2160 @ 256  10 M triangles + constant_medium
"""
from __future__ import annotations

import math

import numpy as np

from . import scene as S
from .scene import (box, bvh_node, camera, checker_texture, constant_medium, constant_texture, dielectric,
                    diffuse_light, flip_normals, hitable_list, image_texture, lambertian, metal, rotate_y,
                    sphere, translate, vec3, xy_rect, xz_rect, yz_rect)

CONFIG_SIZES = {
    "cornell": (512, 512, 16),
    "spheres": (1200, 800, 64),
    "teapot": (1024, 1024, 256),
    "soldier": (1920, 1080, 1024),
    "tessellated": (3840, 2160, 256),
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


def by_name(name: str, **kw):
    fn = {"cornell": cornell_box_classic, "spheres": random_spheres}.get(name)
    if fn is None:
        raise KeyError(name)
    nx, ny, _ = CONFIG_SIZES[name]
    return fn(aspect=float(np.float32(nx) / np.float32(ny)), **kw)
