"""Oracle GJK / EPA (the float64 build of csrc/mjb_ccd_core.h) vs golden vectors produced by executing the REFERENCE's own
`ccd` (/root/reference/mujoco_warp/_src/collision_gjk.py:2200, run under tests/golden/warp_interp.py by
tests/golden/make_ccd_golden.py): 540 seeded pairs over all 15 type combinations of sphere / capsule / ellipsoid / cylinder
/ box -- separated (early-out and distance mode), shallow and deep penetrations, with and without margin.

ncollision must be equal and dist agree to 1e-8 on every pair (both sides float64).  Witness points x1, x2 agree to 1e-7
except on a few pairs per type: (a) a stopping test that falls within rounding of its threshold costs one EPA iteration and
moves the witness by < the ccd tolerance (seen: <= 2.2e-6); (b) exact ties -- e.g. in a deep sphere-cylinder penetration the
reference zeroes the cylinder's radius (collision_gjk.py:2259 writes size2 = 0 back into a non-sphere geom2), the initial
polytope's face normal is then exactly perpendicular to the remaining segment and `sign(0 +- 1e-17)` picks the support
point (collision_gjk.py:137).  Both outcomes have the same depth; at most 3 such pairs per type are tolerated."""

import ctypes
import os

import numpy as np
import pytest

from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ccd_pairs.npz")
NAMES = {2: "sphere", 3: "capsule", 4: "ellipsoid", 5: "cylinder", 6: "box"}
TYPES = [2, 3, 4, 5, 6]
PAIRS = [(a, b) for i, a in enumerate(TYPES) for b in TYPES[i:]]


def run_pair(lib, g, i):
  dp = ctypes.POINTER(ctypes.c_double)
  keep = [np.ascontiguousarray(g[k][i], dtype=np.float64) for k in ("p1", "r1", "s1", "p2", "r2", "s2")]
  dist = ctypes.c_double(0.0)
  x1, x2 = (ctypes.c_double * 3)(), (ctypes.c_double * 3)()
  n = lib.mjo_ccd(int(g["t1"][i]), int(g["t2"][i]), *[a.ctypes.data_as(dp) for a in keep], ctypes.c_double(float(g["margin"][i])),
                  ctypes.c_double(float(g["cutoff"][i])), ctypes.c_double(1e-6), 35, ctypes.byref(dist), x1, x2)  # fmt: skip
  return n, dist.value, np.array(x1[:]), np.array(x2[:])


@pytest.mark.parametrize("pair", PAIRS, ids=lambda p: f"{NAMES[p[0]]}-{NAMES[p[1]]}")
def test_ccd_matches_reference_execution(pair):
  lib = O.load()
  lib.mjo_ccd.restype = ctypes.c_int
  with np.load(GOLDEN) as z:
    g = {k: z[k] for k in z.files}
  idx = np.nonzero((g["t1"] == pair[0]) & (g["t2"] == pair[1]))[0]
  assert len(idx) >= 30
  loose = npen = 0
  for i in idx:
    n, dist, x1, x2 = run_pair(lib, g, i)
    assert n == int(g["ncollision"][i]), (pair, i)
    if g["dist"][i] > 1e29:
      assert dist > 1e29, (pair, i, dist)
      continue
    npen += g["dist"][i] < 0
    assert abs(dist - g["dist"][i]) < 1e-8, (pair, i, dist, g["dist"][i])
    err = max(np.abs(x1 - g["x1"][i]).max(), np.abs(x2 - g["x2"][i]).max())
    if err > 1e-7:
      loose += 1
      assert err < 5e-2, (pair, i, err)
  assert loose <= 3, (pair, loose)
  assert npen >= 5, (pair, npen)


GOLDEN_MESH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ccd_mesh_pairs.npz")


@pytest.mark.parametrize("t1", TYPES + [7], ids=lambda t: f"{NAMES.get(t, 'mesh')}-mesh")
def test_ccd_mesh_matches_reference_execution(t1):
  """Convex-hull meshes (4-20 vertices, exhaustive support search with the cached start vertex, collision_gjk.py:140-154)
  against every other convex type: 30 pairs each from the reference's `ccd` (make_ccd_golden.py --mesh)."""
  lib = O.load()
  lib.mjo_ccd_mesh.restype = ctypes.c_int
  dp = ctypes.POINTER(ctypes.c_double)
  with np.load(GOLDEN_MESH) as z:
    g = {k: z[k] for k in z.files}
  idx = np.nonzero(g["t1"] == t1)[0]
  assert len(idx) >= 30
  loose = npen = 0
  for i in idx:
    keep = {k: np.ascontiguousarray(g[k][i], dtype=np.float64) for k in ("p1", "r1", "s1", "v1", "p2", "r2", "s2", "v2")}
    P = lambda k: keep[k].ctypes.data_as(dp)  # noqa: E731
    dist = ctypes.c_double(0.0)
    x1, x2 = (ctypes.c_double * 3)(), (ctypes.c_double * 3)()
    n = lib.mjo_ccd_mesh(int(t1), 7, P("p1"), P("r1"), P("s1"), P("v1"), int(g["nv1"][i]), P("p2"), P("r2"), P("s2"), P("v2"), int(g["nv2"][i]),
                         ctypes.c_double(float(g["margin"][i])), ctypes.c_double(float(g["cutoff"][i])), ctypes.c_double(1e-6), 35,
                         ctypes.byref(dist), x1, x2)  # fmt: skip
    assert n == int(g["ncollision"][i]), (t1, i)
    if g["dist"][i] > 1e29:
      assert dist.value > 1e29, (t1, i, dist.value)
      continue
    npen += g["dist"][i] < 0
    assert abs(dist.value - g["dist"][i]) < 1e-8, (t1, i, dist.value, g["dist"][i])
    err = max(np.abs(np.array(x1[:]) - g["x1"][i]).max(), np.abs(np.array(x2[:]) - g["x2"][i]).max())
    if err > 1e-7:
      loose += 1
      assert err < 5e-2, (t1, i, err)
  assert loose <= 3, (t1, loose)
  assert npen >= 5, (t1, npen)


def _build_f32():
  """Host build of the FLOAT32 instantiation of csrc/mjb_ccd_core.h (tests/csrc/ccd_f32.c): the arithmetic of the CUDA kernel."""
  import subprocess
  import tempfile

  here = os.path.dirname(os.path.abspath(__file__))
  out = os.path.join(tempfile.gettempdir(), f"libmjb_ccd_f32_{os.getuid()}.so")
  src = os.path.join(here, "csrc", "ccd_f32.c")
  core = os.path.join(here, "..", "mujoco_warp_b200", "csrc", "mjb_ccd_core.h")
  if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(core)):
    subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-o", out, src, "-lm"], check=True)
  lib = ctypes.CDLL(out)
  lib.ccd32.restype = ctypes.c_int
  return lib


@pytest.mark.parametrize("golden", ["ccd_pairs.npz", "ccd_mesh_pairs.npz"])
def test_float32_build_of_the_device_ccd_core_against_reference_execution(golden):
  """The device runs the same source in float32.  Against the reference's float64 execution: the collision decision is equal
  on every pair; the depth agrees to ~1e-6 (the ccd tolerance) typically, but EPA on smooth shapes in float32 can stop on a
  sliver face of the polytope whose normal is rounding-limited: worst seen 1.1e-4 (sphere-ellipsoid, depth 0.089).  These
  bounds are what tests/test_gpu_reference_golden.py allows the CUDA path on the GJK / EPA fixtures."""
  lib = _build_f32()
  dp = ctypes.POINTER(ctypes.c_double)
  with np.load(os.path.join(os.path.dirname(GOLDEN), golden)) as z:
    g = {k: z[k] for k in z.files}
  mesh = "v1" in g
  errs = []
  for i in range(len(g["dist"])):
    keep = {k: np.ascontiguousarray(g[k][i], dtype=np.float64) for k in ("p1", "r1", "s1", "p2", "r2", "s2")}
    v1 = np.ascontiguousarray(g["v1"][i], dtype=np.float64) if mesh else np.zeros(3)
    v2 = np.ascontiguousarray(g["v2"][i], dtype=np.float64) if mesh else np.zeros(3)
    P = lambda a: a.ctypes.data_as(dp)  # noqa: E731
    dist = ctypes.c_double(0.0)
    x1, x2 = (ctypes.c_double * 3)(), (ctypes.c_double * 3)()
    n = lib.ccd32(int(g["t1"][i]), int(g["t2"][i]), P(keep["p1"]), P(keep["r1"]), P(keep["s1"]), P(v1), int(g["nv1"][i]) if mesh else 0,
                  P(keep["p2"]), P(keep["r2"]), P(keep["s2"]), P(v2), int(g["nv2"][i]) if mesh else 0, ctypes.c_double(float(g["margin"][i])),
                  ctypes.c_double(float(g["cutoff"][i])), ctypes.c_double(1e-6), 35, ctypes.byref(dist), x1, x2)  # fmt: skip
    assert n == int(g["ncollision"][i]), (golden, i)
    if g["dist"][i] > 1e29:
      assert dist.value > 1e29, (golden, i)
      continue
    errs.append(abs(dist.value - g["dist"][i]))
  errs = np.array(errs)
  assert errs.max() < 3e-4, errs.max()
  assert np.percentile(errs, 90) < 5e-6, np.percentile(errs, 90)
  assert np.median(errs) < 2e-7, np.median(errs)
