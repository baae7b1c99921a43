"""Host logic: the Python mirror of the reference's scene API flattens to the same bytes as the C++
flattener walking REAL reference objects (host/srt_flatten.hpp through the oracle harness)."""
import numpy as np
import pytest

from simple_raytracing_render_b200 import ffi, scene as S, scenes


@pytest.mark.parametrize("name,kw", [("cornell", {}), ("spheres", {}), ("teapot", {"divs": 12}),
                                     ("soldier", {}), ("tessellated", {"n_side": 16}),
                                     ("final", {}), ("zoo", {}), ("zoo_env", {})])
def test_python_and_cpp_flatteners_agree(oracle, name, kw):
    world, hlist, cam = scenes.by_name(name, **kw)
    mine = S.flatten(world, hlist, cam)
    theirs = oracle.RefBuilder().realize(world, hlist, cam).flat()
    assert mine.equal(theirs) == []


def test_cornell_layout():
    fs = S.flatten(*scenes.cornell_box_classic(1.0))
    assert len(fs.entries) == 8 and len(fs.prims) == 18 and len(fs.lights) == 1
    # two instanced boxes: translate then rotate_y, outermost first
    e = fs.entries[6]
    assert e["op_count"] == 2 and fs.ops[e["first_op"]]["kind"] == ffi.OP_TRANSLATE
    assert fs.ops[e["first_op"] + 1]["kind"] == ffi.OP_ROTATE_Y
    assert fs.groups[e["geom_index"]]["prim_count"] == 6
    # flip_normals folded into the prim flag
    assert fs.prims[0]["flags"] == ffi.PRIM_FLIP and fs.prims[1]["flags"] == 0
    assert fs.lights[0]["type"] == ffi.PRIM_XZ_RECT


def test_bvh_leaf_order_reproduces_reference_qsort(oracle):
    """bvh.h:96-119: random axis from the global drand48 stream, qsort with a comparator that never
    returns 0, median split — including ties (equal box minima)."""
    S.set_seed(12345)
    mat = S.lambertian(S.constant_texture(S.vec3(.5, .5, .5)))
    rng = np.random.default_rng(3)
    items = []
    for i in range(257):
        c = rng.integers(0, 6, size=3).astype(np.float32)  # many exact ties
        items.append(S.sphere(S.vec3(*c), 0.4, mat))
    world = S.hitable_list([S.bvh_node(items, len(items), 0, 1)], 1)
    cam = S.camera(S.vec3(0, 0, -10), S.vec3(0, 0, 0), S.vec3(0, 1, 0), 40, 1, 0, 10, 0, 1)
    mine = S.flatten(world, None, cam)
    theirs = oracle.RefBuilder().realize(world, None, cam).flat()
    assert mine.equal(theirs) == []


def test_flat_scene_roundtrip(tmp_path):
    fs = S.flatten(*scenes.by_name("spheres"))
    p = str(tmp_path / "s.npz")
    fs.save(p)
    assert fs.equal(ffi.FlatScene.load(p)) == []
    d, keep = fs.to_desc()
    assert ffi.FlatScene.from_desc(d).equal(fs) == []


def test_unsupported_graphs_are_rejected():
    mat = S.lambertian(S.constant_texture(S.vec3(1, 1, 1)))
    cam = S.camera(S.vec3(0, 0, -10), S.vec3(0, 0, 0), S.vec3(0, 1, 0), 40, 1, 0, 10, 0, 1)
    inner = S.translate(S.sphere(S.vec3(0, 0, 0), 1, mat), S.vec3(1, 0, 0))
    with pytest.raises(S.FlattenError):
        S.flatten(S.hitable_list([S.bvh_node([inner, inner], 2)], 1), None, cam)
    with pytest.raises(S.FlattenError):
        S.flatten(S.hitable_list([S.sphere(S.vec3(0, 0, 0), 1, mat)], 1), S.sphere(S.vec3(0, 0, 0), 1, None), cam)
