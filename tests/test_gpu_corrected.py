"""Corrected-transport mode (SRT_RENDER_CORRECTED; csrc/srt_shade.cuh: shade_bounce_corrected).
It is explicitly OUTSIDE parity with the reference, so there is no oracle image; what pins it:

  * an analytic white furnace (a lambertian sphere of albedo rho in a closed box of radiance 1 renders as rho,
    exactly, with next-event estimation + MIS and without);
  * the two estimators of the same integral — light shapes sampled (NEE + balance-heuristic MIS, shadow rays) and
    not sampled (BSDF sampling alone) — agree on a scene with every non-specular material, three kinds of light
    shape, glass and occlusion;
  * both schedulers (fused kernel, sorted wavefront with its shadow-ray pass) render the same image;
  * a scene without non-specular surfaces renders exactly as in the reference-compatible mode.
"""
import numpy as np
import pytest

from simple_raytracing_render_b200 import ffi, render, scenes

pytestmark = pytest.mark.gpu


def _render(ctx, triple, nx, ny, ns, flags, monkeypatch=None, wavefront=None):
    import torch
    if monkeypatch is not None and wavefront is not None:
        monkeypatch.setenv("SRT_WAVEFRONT", wavefront)
    sc = render.Scene.from_objects(ctx, *triple)
    buf = torch.zeros(nx * ny * 3, dtype=torch.float32, device="cuda")
    sc.render(render.render_params(nx, ny, ns, flags=flags), buf)
    ctx.synchronize()
    variant = int(sc.info()["render_variant"])
    sc.close()
    return buf.cpu().numpy().reshape(ny, nx, 3) / float(ns), variant


@pytest.mark.parametrize("sampled", [True, False])
@pytest.mark.parametrize("wavefront", ["0", "3"])
def test_white_furnace(gpu_ctx, sampled, wavefront, monkeypatch):
    nx = ny = 64
    img, variant = _render(gpu_ctx, scenes.furnace(1.0, sampled=sampled, albedo=0.5), nx, ny, 1024, ffi.RENDER_CORRECTED,
                           monkeypatch, wavefront)
    assert (variant >= 100) == (wavefront == "3")
    centre = img[24:40, 24:40]  # well inside the sphere's silhouette
    corner = img[:6, :6]        # the box behind it
    print(f"furnace sampled={sampled} wavefront={wavefront}: sphere {centre.mean():.5f} box {corner.mean():.5f}")
    assert abs(centre.mean() - 0.5) < 2e-3
    assert np.abs(centre - 0.5).max() < 0.05
    assert np.allclose(corner, 1.0, atol=1e-5)


# the synthetic MERL table is a 7-degree specular lobe 100 x its diffuse floor: BSDF sampling alone (cosine-weighted)
# does not converge on it in any affordable sample count, so the measured material is pinned separately below
SMOOTH = ("oren", "beckmann", "glass")


@pytest.mark.parametrize("lights", [("rect",), ("sphere",), ("tri",), ("rect", "sphere", "tri")])
def test_light_sampling_and_bsdf_sampling_estimate_the_same_image(gpu_ctx, lights):
    nx = ny = 40
    kw = dict(lights=lights, materials=SMOOTH)
    a, _ = _render(gpu_ctx, scenes.corrected_probe(1.0, sampled=True, **kw), nx, ny, 4096, ffi.RENDER_CORRECTED)
    b, _ = _render(gpu_ctx, scenes.corrected_probe(1.0, sampled=False, **kw), nx, ny, 65536, ffi.RENDER_CORRECTED)
    rel = float(np.mean((a - b) ** 2 / (b * b + 1e-2)))
    print(f"corrected_probe {lights}: NEE+MIS mean {a.mean():.5f}, BSDF sampling mean {b.mean():.5f}, relMSE {rel:.3e}")
    assert abs(a.mean() - b.mean()) < 0.01 * b.mean()
    assert rel < 3e-3


def test_merl_value_is_applied(gpu_ctx):
    """A MERL table whose every entry looks up to 0.5 / pi is a lambertian of albedo 0.5: the two floors render the
    same image (same sampler, same random numbers), which they would not if the looked-up value were discarded as at
    material.h:231-232 — and a table of twice the value renders a brighter floor."""
    nx = ny = 40
    merl, _ = _render(gpu_ctx, scenes.corrected_probe(1.0, materials=("merl",), merl_table=scenes.constant_merl_table(0.5)),
                      nx, ny, 1024, ffi.RENDER_CORRECTED)
    lamb, _ = _render(gpu_ctx, scenes.corrected_probe(1.0, materials=("floor50",)), nx, ny, 1024, ffi.RENDER_CORRECTED)
    twice, _ = _render(gpu_ctx, scenes.corrected_probe(1.0, materials=("merl",), merl_table=scenes.constant_merl_table(1.0)),
                       nx, ny, 1024, ffi.RENDER_CORRECTED)
    rel = float(np.mean((merl - lamb) ** 2 / (lamb * lamb + 1e-2)))
    print(f"constant MERL table vs lambertian 0.5: means {merl.mean():.5f} / {lamb.mean():.5f}, relMSE {rel:.3e}; "
          f"table x 2: mean {twice.mean():.5f}")
    assert rel < 1e-4
    assert twice[30:, :].mean() > 1.5 * merl[30:, :].mean()  # the floor fills the bottom rows


def test_schedulers_agree_in_corrected_mode(gpu_ctx, monkeypatch):
    nx = ny = 40
    a, va = _render(gpu_ctx, scenes.corrected_probe(1.0, sampled=True), nx, ny, 1024, ffi.RENDER_CORRECTED, monkeypatch, "0")
    b, vb = _render(gpu_ctx, scenes.corrected_probe(1.0, sampled=True), nx, ny, 1024, ffi.RENDER_CORRECTED, monkeypatch, "3")
    assert va < 100 <= vb
    rel = float(np.mean((a - b) ** 2 / (b * b + 1e-2)))
    print(f"corrected_probe: fused mean {a.mean():.5f}, wavefront mean {b.mean():.5f}, relMSE {rel:.3e}")
    # same paths, same Philox blocks, fixed-point accumulation: the same image up to float contraction
    close = np.isclose(a, b, rtol=1e-3, atol=1e-4)
    print(f"   close {close.mean():.5f}, bit-identical {np.mean(a == b):.4f}")
    assert close.mean() >= 0.99


def test_corrected_mode_brightens_what_the_reference_estimator_loses(gpu_ctx):
    """The reference's cosine generator points half of its samples into the surface (pdf.h:47-56) and weights the
    rest twice (pdf.h:10-18): the two modes are different estimators and need not agree; this records by how much on
    the Cornell box and checks that the corrected image is finite and lit."""
    from conftest import get_scene
    nx = ny = 48
    a, _ = _render(gpu_ctx, get_scene("cornell"), nx, ny, 1024, ffi.RENDER_CORRECTED)
    b, _ = _render(gpu_ctx, get_scene("cornell"), nx, ny, 1024, 0)
    print(f"cornell: corrected mean {a.mean():.4f}, reference-compatible mean {b.mean():.4f}")
    assert np.isfinite(a).all() and a.mean() > 0.05


def test_light_list_without_a_density_is_refused(gpu_ctx):
    import torch
    from conftest import get_scene
    sc = render.Scene.from_objects(gpu_ctx, *get_scene("zoo_env"))  # an env_sphere in the light list
    buf = torch.zeros(32 * 32 * 3, dtype=torch.float32, device="cuda")
    with pytest.raises(Exception) as e:
        sc.render(render.render_params(32, 32, 4, flags=ffi.RENDER_CORRECTED), buf)
    assert "corrected-transport" in str(e.value)
    sc.close()
