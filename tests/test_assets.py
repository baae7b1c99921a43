"""Host logic of the asset ingestion (simple_raytracing_render_b200/assets.py): the ASCII-PLY loader,
the FBX-binary subset (on a file this test writes, raw and zlib arrays, and on the reference's own blobs where the
reference tree is present), and the MERL-layout reader of brdfmaterial (brdf.h:156-188)."""
import os
import struct
import zlib

import numpy as np
import pytest

from simple_raytracing_render_b200 import assets, ffi, scene as S, scenes

REF = "/root/reference/contents"


def test_ply_ascii_quad_and_triangle(tmp_path):
    p = tmp_path / "m.ply"
    p.write_text("ply\nformat ascii 1.0\ncomment made by a test\nelement vertex 5\nproperty float x\nproperty float y\n"
                 "property float z\nproperty float nx\nproperty float ny\nproperty float nz\nproperty float s\nproperty float t\n"
                 "element face 2\nproperty list uchar int vertex_indices\nend_header\n"
                 "0 0 0 0 0 1 0 0\n1 0 0 0 0 1 1 0\n1 1 0 0 0 1 1 1\n0 1 0 0 0 1 0 1\n2 2 2 0 1 0 0.5 0.5\n"
                 "4 0 1 2 3\n3 0 2 4\n")
    ply = assets.load_ply_ascii(str(p))
    assert ply["vertices"].shape == (5, 3) and ply["normals"].shape == (5, 3) and ply["uvs"].shape == (5, 2)
    assert ply["faces"] == [[0, 1, 2, 3], [0, 2, 4]]
    pos, nrm, uv = assets.soup_from_ply(ply)
    assert pos.shape == (3, 9) and nrm.shape == (3, 9) and uv.shape == (3, 6)  # the quad fans into two triangles
    assert np.array_equal(pos[0], [0, 0, 0, 1, 0, 0, 1, 1, 0]) and np.array_equal(pos[1], [0, 0, 0, 1, 1, 0, 0, 1, 0])
    flipped, _, _ = assets.soup_from_ply(ply, flip_winding=True)  # aiProcess_FlipWindingOrder: corners reversed
    assert np.array_equal(flipped[0], [1, 1, 0, 1, 0, 0, 0, 0, 0])
    with pytest.raises(assets.AssetError):
        (tmp_path / "bad.ply").write_text("plx\n")
        assets.load_ply_ascii(str(tmp_path / "bad.ply"))


def _fbx_node(name: bytes, props: bytes, n_props: int, children: bytes, at: int) -> bytes:
    """One FBX 7.x (< 7500) node record placed at file offset `at`: end offset, property count, property bytes, name."""
    body = props + children + (b"\x00" * 13 if children else b"")
    head_len = 13 + len(name)
    end = at + head_len + len(body)
    return struct.pack("<IIIB", end, n_props, len(props), len(name)) + name + body


def _fbx_array(code: bytes, a: np.ndarray, compress: bool) -> bytes:
    raw = a.tobytes()
    data = zlib.compress(raw) if compress else raw
    return code + struct.pack("<III", a.size, 1 if compress else 0, len(data)) + data


@pytest.mark.parametrize("compress", [False, True])
def test_fbx_binary_subset_roundtrip(tmp_path, compress):
    """A minimal FBX 7200 file: Objects / Geometry with Vertices, PolygonVertexIndex (a quad and a triangle),
    LayerElementNormal (ByPolygonVertex, Direct) and LayerElementUV (ByPolygonVertex, IndexToDirect)."""
    verts = np.array([0, 0, 0, 1, 0, 0, 1, 1, 0, 0, 1, 0, 2, 2, 2], np.float64)
    pvi = np.array([0, 1, 2, ~3, 0, 2, ~4], np.int32)
    normals = np.tile(np.array([0, 0, 1], np.float64), 7)
    uvs = np.array([0, 0, 1, 0, 1, 1, 0, 1, .5, .5], np.float64)
    uvi = np.array([0, 1, 2, 3, 0, 2, 4], np.int32)

    def build(at0):
        # children are laid out back to back; offsets inside are absolute, so build twice (sizes do not depend on `at`)
        def node(name, props, n, kids_fn, at):
            head = 13 + len(name) + len(props)
            kids = kids_fn(at + head) if kids_fn else b""
            return _fbx_node(name, props, n, kids, at)

        def string(sv):
            return b"S" + struct.pack("<I", len(sv)) + sv

        def leaf(name, props, n):
            return lambda at: node(name, props, n, None, at)

        def seq(*makers):
            def f(at):
                out = b""
                for m in makers:
                    out += m(at + len(out))
                return out
            return f

        ln = lambda at: node(b"LayerElementNormal", b"I" + struct.pack("<i", 0), 1,
                             seq(leaf(b"MappingInformationType", string(b"ByPolygonVertex"), 1),
                                 leaf(b"ReferenceInformationType", string(b"Direct"), 1),
                                 leaf(b"Normals", _fbx_array(b"d", normals, compress), 1)), at)
        lu = lambda at: node(b"LayerElementUV", b"I" + struct.pack("<i", 0), 1,
                             seq(leaf(b"MappingInformationType", string(b"ByPolygonVertex"), 1),
                                 leaf(b"ReferenceInformationType", string(b"IndexToDirect"), 1),
                                 leaf(b"UV", _fbx_array(b"d", uvs, compress), 1),
                                 leaf(b"UVIndex", _fbx_array(b"i", uvi, compress), 1)), at)
        geo = lambda at: node(b"Geometry", b"L" + struct.pack("<q", 1) + string(b"m\x00\x01Geometry") + string(b"Mesh"), 3,
                              seq(leaf(b"Vertices", _fbx_array(b"d", verts, compress), 1),
                                  leaf(b"PolygonVertexIndex", _fbx_array(b"i", pvi, compress), 1), ln, lu), at)
        return node(b"Objects", b"", 0, seq(geo), at0)

    header = b"Kaydara FBX Binary  \x00\x1a\x00" + struct.pack("<I", 7200)
    body = build(len(header))
    path = tmp_path / "m.fbx"
    path.write_bytes(header + body + b"\x00" * 13)
    geos = assets.fbx_geometries(str(path))
    assert len(geos) == 1 and geos[0]["vertices"].shape == (5, 3) and geos[0]["poly_index"].tolist() == pvi.tolist()
    pos, nrm, uv = assets.soup_from_geometry(geos[0], flip_winding=True, flip_uvs=True)
    assert pos.shape == (3, 9) and nrm.shape == (3, 9) and uv.shape == (3, 6)
    assert np.array_equal(pos[0], [1, 1, 0, 1, 0, 0, 0, 0, 0])            # quad's first fan triangle, corners reversed
    assert np.array_equal(pos[2], [2, 2, 2, 1, 1, 0, 0, 0, 0])            # the triangle, reversed
    assert np.allclose(uv[2], [.5, .5, 1, 0, 0, 1])                       # FlipUVs: v -> 1 - v
    assert np.array_equal(nrm[0], [0, 0, 1] * 3)


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference tree")
def test_reference_blobs_load():
    geos = assets.fbx_geometries(os.path.join(REF, "models", "Soilder.FBX"))
    # geometry A 2 267 vertices / 10 053 polygon-vertex indices, geometry B 202 / 696
    assert [(g["vertices"].shape[0], g["poly_index"].shape[0]) for g in geos] == [(2267, 10053), (202, 696)]
    pos, nrm, uv = assets.soup_from_geometry(geos[0], flip_winding=True)
    assert pos.shape[1] == 9 and nrm.shape == pos.shape and uv.shape == (pos.shape[0], 6) and np.isfinite(pos).all()
    ply = assets.load_ply_ascii(os.path.join(REF, "models", "bunny.ply"))
    assert ply["vertices"].shape[1] == 3 and len(ply["faces"]) > 1000 and all(len(f) == 3 for f in ply["faces"][:100])
    if scenes.real_assets():  # the cache tools/make_assets.py wrote is this mesh, FlipWindingOrder applied
        cached = np.load(assets.cache_path("soldier_mesh.npz"))
        assert np.array_equal(cached["pos"], pos) and np.array_equal(cached["uv"], uv)
        assert assets.load_cached("sky4.npy").shape == (2048, 4096, 3)


def test_merl_layout_file_reads_like_read_brdf(tmp_path):
    """brdfmaterial(path, albedo): 3 int32 dimensions (90 x 90 x 180) then 3 * 90 * 90 * 180 doubles (brdf.h:171-184);
    a file with other dimensions is refused as read_brdf refuses it ("Dimensions don't match")."""
    table = np.arange(ffi.MERL_DOUBLES, dtype=np.float64) * 1e-6
    good = tmp_path / "m.binary"
    good.write_bytes(np.array([90, 90, 180], np.int32).tobytes() + table.tobytes())
    m = S.brdfmaterial(str(good), S.vec3(0.5, 0.5, 0.5))
    assert np.array_equal(m.table, table)
    bad = tmp_path / "bad.binary"
    bad.write_bytes(np.array([90, 90, 90], np.int32).tobytes() + table.tobytes())
    with pytest.raises(ValueError):
        S.brdfmaterial(str(bad), S.vec3(0.5, 0.5, 0.5))
