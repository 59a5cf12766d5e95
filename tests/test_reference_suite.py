"""The reference's own pytest files, UNMODIFIED, against the drop-in (tools/run_reference_tests.sh).

They are read in place from /root/reference when that exists, else from baseline/_ref/ (staged, git-ignored copies
that travel to the GPU box with the snapshot; `tools/run_reference_tests.sh stage`).  Neither present -> skipped.
What is asserted: no failure, no error, the tests that pin the API contract all PASS, and the only skips are the
ones the reference gates itself (env-gated tools, the dead src/grid.rs API)."""
import os
import subprocess
import sys
import xml.etree.ElementTree as ET

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

MUST_PASS = {            # file -> minimum number of passing tests (the reference's counts at v0.0.8)
    "test_api": 2, "test_api_validation": 4, "test_determinism": 2, "test_grid_generate": 6, "test_camera": 20,
    "test_t31_integration": 2, "test_t41_scene": 2, "test_colormap": 12, "test_height_upload": 8, "test_dem_stats": 1,
    "test_dem_normalization": 3, "test_tonemap": 2, "test_diagnostics": 1,
}
# the reference's own env gates, the dead src/grid.rs API, and test_colormap's "feature not built" twin of a test that runs
MAY_SKIP = {"test_backends", "test_perf", "test_grid_generator", "test_terrain_fs", "test_colormap"}


def test_reference_pytest_files_pass_unmodified(tmp_path):
    if not (os.path.isdir("/root/reference/tests") or os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "tests"))):
        pytest.skip("reference tests neither at /root/reference nor staged under baseline/_ref")
    xml = tmp_path / "ref.xml"
    env = dict(os.environ, VF_REF_JUNIT=str(xml))
    env.pop("VF_FORCE_LUT_UNORM", None)
    p = subprocess.run(["bash", os.path.join(ROOT, "tools", "run_reference_tests.sh")], env=env, capture_output=True, text=True, timeout=900)
    assert xml.exists(), p.stdout[-3000:] + p.stderr[-2000:]
    passed, skipped, bad = {}, set(), []
    for tc in ET.parse(xml).getroot().iter("testcase"):
        parts = (tc.get("classname") or tc.get("name") or "").split(".")     # tests.test_camera[.TestClass]
        mod = next((q for q in parts if q.startswith("test_") or q.endswith("_test")), parts[-1])
        if tc.find("failure") is not None or tc.find("error") is not None:
            bad.append(f"{mod}::{tc.get('name')}")
        elif tc.find("skipped") is not None:
            skipped.add(mod)
        else:
            passed[mod] = passed.get(mod, 0) + 1
    assert not bad, f"failed: {bad}\n{p.stdout[-4000:]}"
    assert p.returncode == 0, p.stdout[-3000:]
    for mod, n in MUST_PASS.items():
        assert passed.get(mod, 0) >= n, f"{mod}: {passed.get(mod, 0)} passed, expected >= {n}\n{p.stdout[-2000:]}"
    assert skipped <= MAY_SKIP, f"unexpected skips: {skipped - MAY_SKIP}"
    print("reference suite:", sum(passed.values()), "passed;", "skipped modules:", sorted(skipped), file=sys.stderr)
