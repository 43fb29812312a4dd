import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)


def pytest_configure(config):
  config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
  """The suite needs libmjb200.so (git-ignored build output): build it once when a fresh checkout has none (nvcc cross-compiles
  without a GPU; `__graft_entry__.build()` does the same).  An existing library is left alone."""
  import shutil

  from mujoco_warp_b200 import _lib

  if not os.path.exists(os.environ.get("MJB200_LIB", _lib.LIB_PATH)) and shutil.which("nvcc"):
    _lib.build()


@pytest.fixture(scope="session")
def humanoid():
  import mujoco_warp_b200 as mjw

  return mjw.load_mjcf(os.path.join(ROOT, "benchmarks", "humanoid", "humanoid.xml"))
