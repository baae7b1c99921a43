"""Host logic: the image writers behind CRender::Run (Raytracing_n.cpp:873-886 writes P3 text; the binary forms hold
the same pixels)."""
import struct
import zlib

import numpy as np

from simple_raytracing_render_b200 import render


def test_writers_hold_the_same_pixels(tmp_path):
    nx, ny = 7, 5
    img = np.random.default_rng(0).integers(0, 256, size=(ny, nx, 3), dtype=np.uint8)
    render.write_image(str(tmp_path / "a.ppm"), img, nx, ny)  # the reference's own format
    render.write_image(str(tmp_path / "a.pnm"), img, nx, ny)
    render.write_image(str(tmp_path / "a.png"), img, nx, ny)
    text = (tmp_path / "a.ppm").read_text().split("\n")
    assert text[:3] == ["P3", f"{nx} {ny}", "255"]
    assert np.array_equal(np.array([l.split() for l in text[3:3 + nx * ny]], np.uint8), img.reshape(-1, 3))
    b = (tmp_path / "a.pnm").read_bytes()
    head = f"P6\n{nx} {ny}\n255\n".encode()
    assert b.startswith(head) and b[len(head):] == img.tobytes()
    p = (tmp_path / "a.png").read_bytes()
    assert p[:8] == b"\x89PNG\r\n\x1a\n"
    w, h, depth, ctype = struct.unpack(">IIBB", p[16:26])
    assert (w, h, depth, ctype) == (nx, ny, 8, 2)
    i = p.index(b"IDAT")
    n = struct.unpack(">I", p[i - 4:i])[0]
    assert struct.unpack(">I", p[i + 4 + n:i + 8 + n])[0] == zlib.crc32(p[i:i + 4 + n]) & 0xFFFFFFFF
    rows = np.frombuffer(zlib.decompress(p[i + 4:i + 4 + n]), np.uint8).reshape(ny, 1 + 3 * nx)
    assert (rows[:, 0] == 0).all() and np.array_equal(rows[:, 1:].reshape(ny, nx, 3), img)
