import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def backend_lib():
    """libsrt_b200.so, built in-tree if missing (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g
    g.build_backend()
    from simple_raytracing_render_b200 import ffi
    return ffi.load_backend()


@pytest.fixture(scope="session")
def oracle():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref/libsrt_ref.so not built (needs /root/reference)")
    return ref


@pytest.fixture(scope="session")
def gpu_ctx(backend_lib):
    from simple_raytracing_render_b200 import render
    ctx = render.Context(0)
    yield ctx
    ctx.close()


def record_parity(key, values):
    """Measured parity figures (relMSE per scene and scheduler, t errors, ray counts ...) are merged into one
    JSON file next to the test run; the copy of a GPU-box run is committed as profiles/r02_parity.json."""
    import json
    path = os.environ.get("SRT_PARITY_JSON", os.path.join(ROOT, "gpurun_out", "parity.json"))
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        try:
            with open(path) as f:
                data = json.load(f)
        except (OSError, ValueError):
            data = {}
        data[key] = values
        with open(path, "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)
    except OSError:
        pass


_SCENES, _REF_SCENES = {}, {}

# the sizes the golden vectors were made at (tests/golden/make_golden.py); "<name>:trace" is the variant
# closest-hit vectors use when the scene holds a constant_medium (its hits are random)
SCENE_KW = {
    "tessellated": {"n_side": 128},
    "tessellated:trace": {"n_side": 128, "fog": False},
    "final:trace": {"media": False},
    "zoo:trace": {"media": False},
}


# scenes whose goldens were made on the reference's own asset blobs (tools/make_assets.py -> assets_cache/)
ASSET_SCENES = ("teapot", "soldier", "soldier_wood")


def get_scene(name):
    """(world, hlist, cam) of a named config, built once per session (the teapot takes seconds)."""
    if name not in _SCENES:
        from simple_raytracing_render_b200 import scenes
        if name.split(":")[0] in ASSET_SCENES and not scenes.real_assets():
            pytest.skip("assets_cache/ is empty: run tools/make_assets.py where /root/reference exists "
                        "(__graft_entry__.build() does); the goldens of this scene use the reference's own assets")
        _SCENES[name] = scenes.by_name(name.split(":")[0], **SCENE_KW.get(name, {}))
    return _SCENES[name]


def trace_name(name):
    return name + ":trace" if name + ":trace" in SCENE_KW else name


def get_ref_scene(ref, name):
    """The same scene as REAL reference objects (oracle), realised once per session."""
    if name not in _REF_SCENES:
        _REF_SCENES[name] = ref.RefBuilder().realize(*get_scene(name))
    return _REF_SCENES[name]
