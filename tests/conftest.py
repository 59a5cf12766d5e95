import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "vulkan-forge_b200", "python"), os.path.join(ROOT, "oracle"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _built():
    import importlib.util
    spec = importlib.util.spec_from_file_location("vf_build", os.path.join(ROOT, "vulkan-forge_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    if not (os.path.exists(mod.lib_path()) and os.path.exists(mod.module_path())):
        mod.build_all()
    return mod


@pytest.fixture(scope="session")
def vf_build():
    return _built()


@pytest.fixture(scope="session")
def oracle():
    import vf_oracle
    vf_oracle.lib()
    return vf_oracle


@pytest.fixture(scope="session")
def vf(vf_build):
    import vulkan_forge
    return vulkan_forge


@pytest.fixture(scope="session")
def ext(vf):
    return vf._vulkan_forge


@pytest.fixture(scope="session")
def luts():
    return {n: np.fromfile(os.path.join(ROOT, "tests", "golden", f"lut_{n}.rgba"), np.uint8)
            for n in ("viridis", "magma", "terrain")}


@pytest.fixture(scope="session")
def synthetic_dem(vf):
    from vulkan_forge._synthetic import synthetic_dem as f
    return f


def frame_diff(got, ref):
    """(fraction of pixels within 1 LSB on every channel, max abs diff)."""
    d = np.abs(got.astype(np.int16) - ref.astype(np.int16)).max(-1)
    return float((d <= 1).mean()), int(d.max())
