"""The stage-group kernels exist in two builds for the smooth-dynamics groups -- one warp per world and four worlds per warp -- and the
launch picks by world count and model size (mjb_api.cu:group_variant).  The parity suites run few worlds, i.e. the one-warp builds;
this test re-runs them in a child process with the other builds forced (MJB_SMOOTH_WIDE=0, MJB_POS_NARROW=1: read once per process)."""

import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parity_suites_with_the_packed_builds_forced():
  env = dict(os.environ, MJB_SMOOTH_WIDE="0", MJB_POS_NARROW="1")
  cmd = [sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", "-p", "no:cacheprovider",
         os.path.join(ROOT, "tests", "test_gpu_reference_golden.py"), os.path.join(ROOT, "tests", "test_gpu_parity.py"),
         os.path.join(ROOT, "tests", "test_gpu_options.py")]  # fmt: skip
  r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
  assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
