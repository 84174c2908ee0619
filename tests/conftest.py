"""pytest configuration: `gpu` marker, repo root on sys.path, shared fixtures.

-m "not gpu": oracle vs golden vectors / reference substrate, host logic, C-ABI symbol table (no compute).
-m gpu:       the parity tests proper — CUDA path through the C ABI vs the CPU oracle on the same inputs.
Nothing here reads /root/reference at run time except the tests explicitly guarded by `reference_tree`.
"""
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """The in-tree libraries; build them when missing (CPU box) — never on the GPU path silently."""
    from volume_path_tracer_b200 import api
    from oracle import oracle_lib
    if not api.LIB_PATH.exists() or not oracle_lib.OWN_PATH.exists():
        import __graft_entry__
        __graft_entry__.build()
    return True


@pytest.fixture(scope="session")
def api(built):
    from volume_path_tracer_b200 import api as _api
    _api.lib()
    return _api


@pytest.fixture(scope="session")
def scenes(api):
    from volume_path_tracer_b200 import scenes as _s
    return _s


@pytest.fixture(scope="session")
def oracle(built):
    from oracle import oracle_lib
    return oracle_lib.Oracle("own")


@pytest.fixture(scope="session")
def oracle_ref(built):
    from oracle import oracle_lib
    if not oracle_lib.REF_PATH.exists():
        pytest.skip("oracle/_ref not built (reference tree absent on this machine)")
    return oracle_lib.Oracle("ref")


@pytest.fixture(scope="session")
def reference_tree():
    p = Path(os.environ.get("VPT_REFERENCE_DIR", "/root/reference"))
    if not (p / "scenes").is_dir():
        pytest.skip("reference tree absent")
    return p


@pytest.fixture(scope="session")
def renderer(api):
    r = api.Renderer(0)
    yield r
    r.close()


# name -> (config, output_size, waves, volume factor): small variants of BASELINE.json's configs the oracle renders in < 1 s
SMALL = {
    "fire": ("fire", (70, 100), 2, 0.25),
    "fire_lowscattering": ("fire_lowscattering", (70, 100), 2, 0.25),
    "wdas_cloud": ("wdas_cloud", (96, 54), 2, 0.05),
    "dense512": ("dense512", (96, 54), 2, 0.125),
}
_cache = {}


@pytest.fixture(scope="session")
def small_case(api, scenes):
    def get(name):
        if name not in _cache:
            cfg, size, waves, factor = SMALL[name]
            scene, spec = scenes.small_variant(cfg, size, waves, factor)
            grids = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
            _cache[name] = (scene, spec, grids)
        return _cache[name]
    return get
