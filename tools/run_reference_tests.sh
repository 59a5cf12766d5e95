#!/bin/bash
# Runs the reference's own pytest files UNMODIFIED against the drop-in.
#
# The files are never committed here (they are reference sources): on the build container they are read in
# place from /root/reference; for the GPU box, which has no /root/reference, `tools/run_reference_tests.sh stage`
# copies tests/*.py, pytest.ini and python/tools/device_diagnostics.py (which tests/test_diagnostics.py executes)
# into baseline/_ref/ -- git-ignored, but part of the gpurun snapshot, the place the task reserves for the
# reference's own files.  Then, on a GPU:
#   tools/run_reference_tests.sh [extra pytest args]      -> junit XML in gpurun_out/reference_tests.xml
# PYTHONPATH points at the drop-in package only (vulkan_forge, vshade); a top-level `_vulkan_forge` is
# deliberately NOT importable, as with the reference's maturin layout (tests/test_grid_generator.py, which
# targets the dead src/grid.rs API, skips itself on that).  The reference's own gates stay as they are:
# VF_TEST_BACKENDS / VF_TEST_PERF (tools that drive graphics backends) and VF_ENABLE_TERRAIN_TESTS (tests/test_terrain_fs.py
# compares two pixels one of which sits a pixel further down the bump than the other; with the documented normal formula,
# whose slopes are (G-1)/3 times flatter than the geometry, the colour ramp outweighs the lighting -- see DESIGN.md).
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
STAGE=$ROOT/baseline/_ref
if [ "$1" = "stage" ]; then
  SRC=${2:-/root/reference}
  rm -rf "$STAGE/tests" "$STAGE/python"
  mkdir -p "$STAGE/tests" "$STAGE/python/tools"
  cp "$SRC"/tests/*.py "$STAGE/tests/"
  cp "$SRC"/pytest.ini "$STAGE/"
  cp "$SRC"/python/tools/device_diagnostics.py "$STAGE/python/tools/"
  (cd "$SRC" && sha256sum tests/*.py pytest.ini python/tools/device_diagnostics.py) > "$STAGE/SHA256SUMS"
  echo "staged $(ls "$STAGE/tests" | wc -l) files into $STAGE"
  exit 0
fi
if [ -d /root/reference/tests ]; then SRCROOT=/root/reference; else SRCROOT=$STAGE; fi
[ -d "$SRCROOT/tests" ] || { echo "no reference tests (run: tools/run_reference_tests.sh stage)"; exit 4; }
mkdir -p "$ROOT/gpurun_out"
cd "$SRCROOT"
PYTHONDONTWRITEBYTECODE=1 PYTHONPATH="$ROOT/vulkan-forge_b200/python" \
  python -m pytest tests -c "$SRCROOT/pytest.ini" --rootdir "$SRCROOT" -p no:cacheprovider -o addopts="--tb=short -q" \
  --junitxml "${VF_REF_JUNIT:-$ROOT/gpurun_out/reference_tests.xml}" "$@"
