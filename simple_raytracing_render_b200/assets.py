"""Mesh and image ingestion for the scenes that load assets.

The reference imports meshes through assimp (model.h:30-43: Triangulate | JoinIdenticalVertices | SortByPType, plus
FlipWindingOrder / FlipUVs on request) and expands every face into a `triangle` with per-corner position, normal and
uv (geometry.h:24-79).  assimp is a Windows binary in the reference tree, so the import is restated here for the two
formats its scenes use:

  * ASCII PLY (`bunny.ply`): `element vertex` with x y z (+ optional nx ny nz, s t / u v), `element face` with a
    vertex-index list.
  * binary FBX, the subset the soldier / hero files need: `Objects/Geometry` records with `Vertices`,
    `PolygonVertexIndex`, `LayerElementNormal` (ByPolygonVertex or ByVertice, Direct or IndexToDirect) and
    `LayerElementUV` (ByPolygonVertex, IndexToDirect or Direct); arrays raw or zlib-compressed; 32-bit record
    offsets (versions < 7500) and 64-bit ones (>= 7500).

The result is a triangle SOUP (positions [n][9], normals [n][9], uvs [n][6]) in the order geometry.h builds its
triangles: mesh after mesh, face after face.  JoinIdenticalVertices only re-indexes, so it does not change the soup;
polygons with more than three corners are fan-triangulated from their first corner (assimp does the same for convex
quads; the reference's own models are triangle meshes).  What assimp does beyond this — which geometry becomes
mesh[0], node transforms — is NOT pinned by any reference test: the oracle and the GPU are fed this
same soup.

Images are decoded by the reference's own stb_image.h (oracle/stb_decode.c, built with the oracle) on the machine
that has the reference tree, and cached as .npy under assets_cache/ (git-ignored, travels with the snapshot like
oracle/_ref): tools/make_assets.py."""
import os
import struct
import zlib
from typing import Dict, List, Optional, Tuple

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CACHE_DIR = os.environ.get("SRT_ASSETS_CACHE", os.path.join(ROOT, "assets_cache"))


class AssetError(RuntimeError):
    pass


# ---- ASCII PLY ------------------------------------------------------------------------------------------------------
def load_ply_ascii(path: str):
    """-> dict(vertices [nv][3] f32, normals [nv][3] | None, uvs [nv][2] | None, faces list of index lists)."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.find(b"end_header")
    if not data.startswith(b"ply") or end < 0:
        raise AssetError(f"{path}: not a PLY file")
    header = data[:end].decode("ascii", "replace").splitlines()
    body = data[data.index(b"\n", end) + 1:]
    if not any(line.split()[:2] == ["format", "ascii"] for line in header if line.strip()):
        raise AssetError(f"{path}: only ASCII PLY is supported")
    elements = []  # (name, count, [property names])
    for line in header:
        w = line.split()
        if not w:
            continue
        if w[0] == "element":
            elements.append([w[1], int(w[2]), []])
        elif w[0] == "property" and elements:
            elements[-1][2].append(w[-1] if w[1] != "list" else "list:" + w[-1])
    tokens = body.split()
    at = 0
    out = {"vertices": None, "normals": None, "uvs": None, "faces": []}
    for name, count, props in elements:
        if name == "vertex":
            k = len(props)
            block = np.array(tokens[at:at + count * k], dtype=np.float64).reshape(count, k)
            at += count * k
            col = {p: i for i, p in enumerate(props)}
            out["vertices"] = block[:, [col["x"], col["y"], col["z"]]].astype(np.float32)
            if all(p in col for p in ("nx", "ny", "nz")):
                out["normals"] = block[:, [col["nx"], col["ny"], col["nz"]]].astype(np.float32)
            for a, b in (("s", "t"), ("u", "v"), ("texture_u", "texture_v")):
                if a in col and b in col:
                    out["uvs"] = block[:, [col[a], col[b]]].astype(np.float32)
                    break
        elif name == "face":
            faces = []
            for _ in range(count):
                n = int(tokens[at])
                faces.append([int(t) for t in tokens[at + 1:at + 1 + n]])
                at += 1 + n
            out["faces"] = faces
        else:  # an element we do not use: skip its tokens (scalar properties only)
            if any(p.startswith("list:") for p in props):
                raise AssetError(f"{path}: list property in unused element '{name}'")
            at += count * len(props)
    if out["vertices"] is None:
        raise AssetError(f"{path}: no vertex element")
    return out


# ---- binary FBX -----------------------------------------------------------------------------------------------------
_FBX_MAGIC = b"Kaydara FBX Binary  \x00\x1a\x00"
_SCALARS = {b"Y": "<h", b"C": "<?", b"I": "<i", b"F": "<f", b"D": "<d", b"L": "<q"}
_ARRAYS = {b"f": np.dtype("<f4"), b"d": np.dtype("<f8"), b"l": np.dtype("<i8"), b"i": np.dtype("<i4"), b"b": np.dtype("u1")}


class FbxNode:
    __slots__ = ("name", "props", "children")

    def __init__(self, name, props, children):
        self.name, self.props, self.children = name, props, children

    def find(self, name: str) -> Optional["FbxNode"]:
        for c in self.children:
            if c.name == name:
                return c
        return None

    def find_all(self, name: str) -> List["FbxNode"]:
        return [c for c in self.children if c.name == name]


def _read_props(buf: bytes, at: int, count: int):
    props = []
    for _ in range(count):
        t = buf[at:at + 1]
        at += 1
        if t in _SCALARS:
            fmt = _SCALARS[t]
            props.append(struct.unpack_from(fmt, buf, at)[0])
            at += struct.calcsize(fmt)
        elif t in _ARRAYS:
            n, enc, clen = struct.unpack_from("<III", buf, at)
            at += 12
            raw = buf[at:at + clen]
            at += clen
            if enc == 1:
                raw = zlib.decompress(raw)
            elif enc != 0:
                raise AssetError(f"FBX array encoding {enc}")
            props.append(np.frombuffer(raw, dtype=_ARRAYS[t], count=n))
        elif t in (b"S", b"R"):
            (n,) = struct.unpack_from("<I", buf, at)
            at += 4
            raw = buf[at:at + n]
            at += n
            props.append(raw.decode("utf-8", "replace") if t == b"S" else raw)
        else:
            raise AssetError(f"FBX property type {t!r} at offset {at - 1}")
    return props, at


def _read_node(buf: bytes, at: int, wide: bool):
    """-> (node | None for the null sentinel, offset after it)."""
    if wide:
        end, n_props, props_len = struct.unpack_from("<QQQ", buf, at)
        at += 24
    else:
        end, n_props, props_len = struct.unpack_from("<III", buf, at)
        at += 12
    name_len = buf[at]
    at += 1
    if end == 0:
        return None, at
    name = buf[at:at + name_len].decode("ascii", "replace")
    at += name_len
    props, after = _read_props(buf, at, n_props)
    if after != at + props_len:
        raise AssetError(f"FBX node '{name}': property list length mismatch")
    at = after
    children = []
    while at < end:
        child, at = _read_node(buf, at, wide)
        if child is None:
            break
        children.append(child)
    return FbxNode(name, props, children), end


def parse_fbx(path: str) -> Tuple[int, FbxNode]:
    """-> (version, root node holding the top-level records)."""
    with open(path, "rb") as f:
        buf = f.read()
    if not buf.startswith(_FBX_MAGIC):
        raise AssetError(f"{path}: not a binary FBX file")
    (version,) = struct.unpack_from("<I", buf, len(_FBX_MAGIC))
    wide = version >= 7500
    at = len(_FBX_MAGIC) + 4
    top = []
    sentinel = 25 if wide else 13
    while at + sentinel <= len(buf):
        node, at = _read_node(buf, at, wide)
        if node is None:
            break
        top.append(node)
    return version, FbxNode("", [], top)


def _layer_values(layer: Optional[FbxNode], values_name: str, index_name: str, width: int, poly_index: np.ndarray):
    """Per polygon-vertex values of a LayerElement* record: [n_polygon_vertices][width] or None."""
    if layer is None:
        return None
    vals = layer.find(values_name)
    if vals is None:
        return None
    v = np.asarray(vals.props[0], dtype=np.float64).reshape(-1, width)
    mapping = (layer.find("MappingInformationType").props[0] if layer.find("MappingInformationType") else "ByPolygonVertex")
    reference = (layer.find("ReferenceInformationType").props[0] if layer.find("ReferenceInformationType") else "Direct")
    n_pv = len(poly_index)
    control_point = np.where(poly_index < 0, ~poly_index, poly_index)
    if mapping in ("ByPolygonVertex",):
        key = np.arange(n_pv)
    elif mapping in ("ByVertice", "ByVertex", "ByControlPoint"):
        key = control_point
    elif mapping == "AllSame":
        key = np.zeros(n_pv, dtype=np.int64)
    else:
        raise AssetError(f"FBX layer mapping '{mapping}' is outside the supported subset")
    if reference in ("IndexToDirect", "Index"):
        idx_node = layer.find(index_name)
        if idx_node is None:
            raise AssetError(f"FBX layer is {reference} but has no {index_name}")
        idx = np.asarray(idx_node.props[0], dtype=np.int64)
        key = idx[key]
    elif reference != "Direct":
        raise AssetError(f"FBX layer reference '{reference}' is outside the supported subset")
    bad = key < 0  # unmapped corners (index -1): zero, as assimp leaves them
    out = v[np.where(bad, 0, key)]
    out[bad] = 0.0
    return out


def fbx_geometries(path: str) -> List[Dict[str, np.ndarray]]:
    """The mesh geometries of a binary FBX file in file order: dict(vertices [nv][3] f64, poly_index i32
    [n_polygon_vertices] (the last corner of a polygon is stored as ~index), normals / uvs per polygon-vertex or
    None, name)."""
    version, root = parse_fbx(path)
    objects = root.find("Objects")
    if objects is None:
        raise AssetError(f"{path}: no Objects record")
    out = []
    for g in objects.find_all("Geometry"):
        vn, pn = g.find("Vertices"), g.find("PolygonVertexIndex")
        if vn is None or pn is None:
            continue  # a shape / line geometry
        vertices = np.asarray(vn.props[0], dtype=np.float64).reshape(-1, 3)
        poly_index = np.asarray(pn.props[0], dtype=np.int64)
        normals = _layer_values(g.find("LayerElementNormal"), "Normals", "NormalsIndex", 3, poly_index)
        uvs = _layer_values(g.find("LayerElementUV"), "UV", "UVIndex", 2, poly_index)
        name = g.props[1] if len(g.props) > 1 and isinstance(g.props[1], str) else ""
        out.append({"vertices": vertices, "poly_index": poly_index, "normals": normals, "uvs": uvs,
                    "name": name.split("\x00")[0], "version": version})
    if not out:
        raise AssetError(f"{path}: no mesh geometry found")
    return out


def _fan_triangles(poly_index: np.ndarray) -> np.ndarray:
    """[n_tris][3] polygon-vertex positions: every polygon as a fan from its first corner."""
    ends = np.flatnonzero(poly_index < 0)
    starts = np.concatenate([[0], ends[:-1] + 1])
    sizes = ends - starts + 1
    if np.all(sizes == 3):
        return (starts[:, None] + np.arange(3)[None, :]).astype(np.int64)
    tris = []
    for s, n in zip(starts.tolist(), sizes.tolist()):
        for k in range(1, n - 1):
            tris.append((s, s + k, s + k + 1))
    return np.asarray(tris, dtype=np.int64).reshape(-1, 3)


def soup_from_geometry(geo: Dict[str, np.ndarray], flip_winding: bool = False, flip_uvs: bool = False):
    """(positions [n][9], normals [n][9] | None, uvs [n][6] | None) as float32, corner order as geometry.h:61-77 reads
    the faces assimp hands it: FlipWindingOrder reverses the corners of every face, FlipUVs maps v -> 1 - v."""
    pi = geo["poly_index"]
    tri = _fan_triangles(pi)
    if flip_winding:
        tri = tri[:, ::-1]
    cp = np.where(pi < 0, ~pi, pi)
    pos = geo["vertices"][cp[tri]].astype(np.float32).reshape(-1, 9)
    nrm = uv = None
    if geo["normals"] is not None:
        nrm = geo["normals"][tri].astype(np.float32).reshape(-1, 9)
    if geo["uvs"] is not None:
        u = geo["uvs"][tri].astype(np.float32)
        if flip_uvs:
            u = u.copy()
            u[..., 1] = np.float32(1.0) - u[..., 1]
        uv = u.reshape(-1, 6)
    return pos, nrm, uv


def soup_from_ply(ply, flip_winding: bool = False):
    """The same for a PLY mesh (faces index the vertex rows directly)."""
    faces = ply["faces"]
    tris = []
    for f in faces:
        for k in range(1, len(f) - 1):
            tris.append((f[0], f[k], f[k + 1]))
    tri = np.asarray(tris, dtype=np.int64).reshape(-1, 3)
    if flip_winding:
        tri = tri[:, ::-1]
    pos = ply["vertices"][tri].astype(np.float32).reshape(-1, 9)
    nrm = ply["normals"][tri].astype(np.float32).reshape(-1, 9) if ply["normals"] is not None else None
    uv = ply["uvs"][tri].astype(np.float32).reshape(-1, 6) if ply["uvs"] is not None else None
    return pos, nrm, uv


def load_model(path: str, flip_uvs: bool = False, flip_winding: bool = False):
    """model(filename, flipUVs, flipWindingOrder, ...) (model.h:26-57) as one triangle soup over all meshes, before the
    `scale` the caller applies (geometry.h:66)."""
    ext = os.path.splitext(path)[1].lower()
    if ext == ".ply":
        return soup_from_ply(load_ply_ascii(path), flip_winding)
    if ext == ".fbx":
        parts = [soup_from_geometry(g, flip_winding, flip_uvs) for g in fbx_geometries(path)]
        pos = np.concatenate([p[0] for p in parts])
        nrm = np.concatenate([p[1] if p[1] is not None else np.zeros_like(p[0]) for p in parts])
        uv = np.concatenate([p[2] if p[2] is not None else np.zeros((len(p[0]), 6), np.float32) for p in parts])
        return pos, nrm, uv
    raise AssetError(f"{path}: unsupported model format")


# ---- the cache that travels with the snapshot ---------------------------------------------------------------------------
def cache_path(name: str) -> str:
    return os.path.join(CACHE_DIR, name)


def have_cached(*names: str) -> bool:
    return all(os.path.exists(cache_path(n)) for n in names)


def load_cached(name: str) -> np.ndarray:
    p = cache_path(name)
    if not os.path.exists(p):
        raise AssetError(f"{p} is missing: run tools/make_assets.py on a machine that has the reference tree")
    return np.load(p)
