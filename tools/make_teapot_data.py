#!/usr/bin/env python
"""Extracts the Utah teapot control mesh (Newell's public data set: 32 bicubic patches over 306
vertices) from the reference's teapotdata.h into simple_raytracing_render_b200/data/utah_teapot.npz.
Only the numbers are read; run in the build container (needs /root/reference)."""
import os
import re
import sys

import numpy as np

SRC = "/root/reference/Raytracing_n/teapotdata.h"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "simple_raytracing_render_b200", "data",
                   "utah_teapot.npz")


def block(text, name):
    m = re.search(name + r"\s*\[[^\]]*\]\s*\[[^\]]*\]\s*=\s*\{(.*?)\};", text, re.S)
    return [float(x) for x in re.findall(r"-?\d+\.?\d*(?:[eE][-+]?\d+)?", m.group(1))]


def main():
    text = open(SRC, errors="replace").read()
    patches = np.array(block(text, "teapotPatches"), np.int32).reshape(32, 16)
    verts = np.array(block(text, "teapotVertices"), np.float32).reshape(306, 3)
    assert patches.min() == 1 and patches.max() == 306
    np.savez_compressed(OUT, vertices=verts, patches=patches)
    print(OUT, verts.shape, patches.shape)


if __name__ == "__main__":
    sys.exit(main())
