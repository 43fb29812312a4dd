"""Oracle narrowphase vs golden vectors produced by executing the REFERENCE's own primitive functions
(/root/reference/mujoco_warp/_src/collision_primitive_core.py, run under tests/golden/warp_shim.py by
tests/golden/make_primitive_golden.py).  This pins the oracle's sphere / capsule / box / plane pair functions to the
reference's code: distances, positions and normals must agree to 1e-9 on 160 seeded configurations per pair type
(generic, close, axis-aligned and exactly parallel poses)."""

import ctypes
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "primitive_core.npz")
NAMES = {0: "plane", 2: "sphere", 3: "capsule", 6: "box"}
PAIRS = [(0, 2), (0, 3), (0, 6), (2, 2), (2, 3), (3, 3), (2, 6), (3, 6), (6, 6)]


@pytest.fixture(scope="module")
def lib():
  lib = O.load()
  dp = ctypes.POINTER(ctypes.c_double)
  lib.mjo_narrowphase_pair.argtypes = [ctypes.c_int, ctypes.c_int, dp, dp, dp, dp, dp, dp, ctypes.c_double, dp, dp, dp]
  lib.mjo_narrowphase_pair.restype = ctypes.c_int
  return lib


def _call(lib, g, i):
  arr = lambda a: np.ascontiguousarray(a, dtype=np.float64).ctypes.data_as(ctypes.POINTER(ctypes.c_double))  # noqa: E731
  dist, pos, frame = np.full(8, 1e10), np.zeros((8, 3)), np.zeros((8, 3, 3))
  keep = [np.ascontiguousarray(g[k][i], dtype=np.float64) for k in ("p1", "r1", "s1", "p2", "r2", "s2")]
  n = lib.mjo_narrowphase_pair(int(g["t1"][i]), int(g["t2"][i]), *[arr(a) for a in keep], float(g["margin"][i]), arr(dist), arr(pos), arr(frame))
  # arr() made temporaries for the outputs: call again with stable buffers
  d, p, f = (ctypes.c_double * 8)(), (ctypes.c_double * 24)(), (ctypes.c_double * 72)()
  n = lib.mjo_narrowphase_pair(int(g["t1"][i]), int(g["t2"][i]), *[arr(a) for a in keep], float(g["margin"][i]), d, p, f)
  return n, np.array(d[:]), np.array(p[:]).reshape(8, 3), np.array(f[:]).reshape(8, 3, 3)


@pytest.mark.parametrize("pair", PAIRS, ids=lambda p: f"{NAMES[p[0]]}-{NAMES[p[1]]}")
def test_oracle_matches_reference_primitive(lib, pair):
  g = np.load(GOLDEN)
  idx = np.nonzero((g["t1"] == pair[0]) & (g["t2"] == pair[1]))[0]
  assert idx.size >= 100
  nactive = 0
  for i in idx:
    n, dist, pos, frame = _call(lib, g, i)
    margin = float(g["margin"][i])
    assert n == int(g["ncand"][i])
    for k in range(n):
      ref_d = float(g["dist"][i, k])
      # a candidate only matters if write_contact keeps it (dist < margin, collision_core.py:199); both sides use a
      # large sentinel for unpopulated slots
      ref_on, got_on = ref_d < margin, dist[k] < margin
      assert ref_on == got_on, (i, k, ref_d, dist[k])
      if not ref_on:
        continue
      nactive += 1
      assert abs(dist[k] - ref_d) < 1e-9, (i, k)
      np.testing.assert_allclose(pos[k], g["pos"][i, k], atol=1e-9, err_msg=str((i, k)))
      np.testing.assert_allclose(frame[k, 0], g["normal"][i, k], atol=1e-9, err_msg=str((i, k)))
  assert nactive > 20
