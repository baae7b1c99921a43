"""Seeded ray sets shared by tests/golden/make_golden.py (which runs them through the reference) and the
GPU tests (which regenerate them instead of loading megabytes of fixtures)."""
import numpy as np


def medium_rays(kind, n, seed):
    """The seeded ray set of the constant_medium transmittance check (regenerated identically by the test):
    origins on a radius-3 shell and, for an eighth of the rays, inside the volume; un-normalised directions
    towards points near the volume (constant_medium.h:31 scales the chord by |direction|)."""
    rng = np.random.default_rng(seed)
    u = rng.standard_normal((n, 3))
    o = 3.0 * u / np.linalg.norm(u, axis=1, keepdims=True)
    inside = rng.uniform(size=n) < 0.125
    o[inside] = rng.uniform(-0.5, 0.5, size=(int(inside.sum()), 3))
    target = rng.uniform(-1.1, 1.1, size=(n, 3))
    d = target - o
    d *= rng.uniform(0.2, 3.0, size=(n, 1)) / np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d, np.zeros((n, 1))], 1).astype(np.float32)


def tess10m_random_rays(n=100_000, seed=9):
    """Random rays inside the bounds of the 10 M-triangle C5 scene (scenes.tessellated_medium, n_side = 2237):
    un-normalised directions, a fiftieth of them parallel to the ground plane."""
    rng = np.random.default_rng(seed)
    o = rng.uniform([0, 0, 0], [555, 180, 555], size=(n, 3))
    d = rng.standard_normal((n, 3)) * rng.uniform(0.1, 50, size=(n, 1))
    d[: n // 50, 1] = 0.0
    return np.concatenate([o, d, np.zeros((n, 1))], 1).astype(np.float32)
