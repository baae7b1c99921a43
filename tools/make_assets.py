#!/usr/bin/env python
"""Decode the reference's asset blobs once, on the machine that has the reference tree, into assets_cache/
(git-ignored; travels to the GPU box with the repo snapshot, like oracle/_ref and the built .so files).

  * images: contents/environment_map/sky4.jpg, contents/textures/NPC_YuanChengBing_A.png,
    contents/textures/TexturesCom_Wood_Wenge_1K_albedo.png (soldier_scene, Raytracing_n.cpp:614-637) and
    contents/environment_map/sky_2.png (teapot_scene, :336) -> uint8 [ny][nx][3] .npy, decoded by the reference's
    OWN stb_image (`stbi_load(path, &x, &y, &n, 0)` exported by oracle/_ref/libsrt_ref.so, which compiles the
    unmodified translation unit): the bytes are what the reference's image_texture would hold.  image_texture
    indexes 3 bytes per texel (texture.h:66-68), so 4-channel files keep their first three channels.
  * mesh: contents/models/Soilder.FBX through simple_raytracing_render_b200.assets (the FBX-binary subset
    loader), FlipWindingOrder as soldier_scene asks (:640), geometry 0 only (model::genhitablemodel returns
    geometries_[0], model.h:73-90) -> soldier_mesh.npz (pos [n][9], nrm [n][9], uv [n][6], unscaled).

Called by __graft_entry__.build() when /root/reference exists; a no-op when the cache is up to date."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simple_raytracing_render_b200 import assets  # noqa: E402

REFERENCE = os.environ.get("SRT_REFERENCE", "/root/reference")
IMAGES = {
    "sky4.npy": "contents/environment_map/sky4.jpg",
    "soldier_albedo.npy": "contents/textures/NPC_YuanChengBing_A.png",
    "wood_albedo.npy": "contents/textures/TexturesCom_Wood_Wenge_1K_albedo.png",
    "sky_2.npy": "contents/environment_map/sky_2.png",
}
MESH = ("soldier_mesh.npz", "contents/models/Soilder.FBX")


def stbi_decode(lib, path: str) -> np.ndarray:
    x, y, n = C.c_int(0), C.c_int(0), C.c_int(0)
    p = lib.stbi_load(path.encode(), C.byref(x), C.byref(y), C.byref(n), 0)
    if not p:
        raise RuntimeError(f"stbi_load failed on {path}")
    try:
        a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), shape=(y.value, x.value, n.value)).copy()
    finally:
        lib.stbi_image_free(p)
    if a.shape[2] < 3:
        a = np.repeat(a[:, :, :1], 3, axis=2)
    # image_texture::value reads data[3*i + 3*nx*j + c] (texture.h:66-68): a 4-channel file would be MIS-indexed by
    # the reference; every file used here is checked to be 3-channel so that the question does not arise
    if a.shape[2] != 3:
        raise RuntimeError(f"{path}: {a.shape[2]} channels; the reference's image_texture assumes 3")
    return np.ascontiguousarray(a)


def main(force: bool = False) -> int:
    if not os.path.isdir(REFERENCE):
        print(f"make_assets: {REFERENCE} not present; nothing to do")
        return 0
    os.makedirs(assets.CACHE_DIR, exist_ok=True)
    lib = None
    for out, rel in IMAGES.items():
        dst, src = assets.cache_path(out), os.path.join(REFERENCE, rel)
        if not force and os.path.exists(dst) and os.path.getmtime(dst) >= os.path.getmtime(src):
            continue
        if lib is None:
            so = os.path.join(ROOT, "oracle", "_ref", "libsrt_ref.so")
            lib = C.CDLL(so)
            lib.stbi_load.restype = C.c_void_p
            lib.stbi_load.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
            lib.stbi_image_free.argtypes = [C.c_void_p]
        img = stbi_decode(lib, src)
        np.save(dst, img)
        print(f"make_assets: {rel} -> {out} {img.shape}")
    dst, src = assets.cache_path(MESH[0]), os.path.join(REFERENCE, MESH[1])
    if force or not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
        geo = assets.fbx_geometries(src)[0]
        pos, nrm, uv = assets.soup_from_geometry(geo, flip_winding=True, flip_uvs=False)
        np.savez(dst, pos=pos, nrm=nrm, uv=uv)
        print(f"make_assets: {MESH[1]} geometry 0 -> {MESH[0]} {pos.shape[0]} triangles")
    return 0


if __name__ == "__main__":
    sys.exit(main(force="--force" in sys.argv))
