"""Golden vectors for the convex narrowphase, produced by EXECUTING the reference's own `ccd` (GJK + EPA).

/root/reference/mujoco_warp/_src/collision_gjk.py is imported unmodified under the numpy Warp interpreter
(`warp_interp.py`) and its `ccd` function is called on seeded random pairs of sphere / capsule / ellipsoid / cylinder / box
geoms: separated, touching within the margin, shallow and deep penetrations, with cutoff 0 (contact mode: separated pairs
return FLOAT_MAX) and cutoff 1e32 (distance mode).  Inputs and (ncollision, dist, x1, x2) are stored in `ccd_pairs.npz`;
tests/test_oracle_ccd_golden.py checks the oracle's build of csrc/mjb_ccd_core.h against the file.

  python tests/golden/make_ccd_golden.py          # primitives -> ccd_pairs.npz
  python tests/golden/make_ccd_golden.py --mesh   # convex-hull meshes (exhaustive support search) -> ccd_mesh_pairs.npz
"""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_runner as R  # noqa: E402

OUT = os.path.join(HERE, "ccd_pairs.npz")
TYPES = [2, 3, 4, 5, 6]  # sphere capsule ellipsoid cylinder box
ITER = 35
TOL = 1e-6


def rand_rot(rng):
  q = rng.normal(size=4)
  w, x, y, z = q / np.linalg.norm(q)
  return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                   [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                   [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])  # fmt: skip


def rand_size(rng, t):
  if t == 2:
    return np.array([rng.uniform(0.05, 0.3), 0.0, 0.0])
  if t in (3, 5):
    return np.array([rng.uniform(0.05, 0.2), rng.uniform(0.05, 0.4), 0.0])
  return rng.uniform(0.05, 0.35, 3)


def main(n_per_pair=36, seed=20260202):
  wp, ref = R.load_reference()
  gjk = sys.modules["mujoco_warp._src.collision_gjk"]
  core = sys.modules["mujoco_warp._src.collision_core"]
  rng = np.random.default_rng(seed)
  rec = {k: [] for k in "t1 t2 p1 r1 s1 p2 r2 s2 margin cutoff ncollision dist x1 x2".split()}
  nv, nf = 10 + 2 * ITER, 6 + 5 * ITER
  for i1, t1 in enumerate(TYPES):
    for t2 in TYPES[i1:]:
      for k in range(n_per_pair):
        s1, s2 = rand_size(rng, t1), rand_size(rng, t2)
        r1, r2 = rand_rot(rng), rand_rot(rng)
        reach = np.linalg.norm(s1[:2] if t1 in (3, 5) else s1) + np.linalg.norm(s2[:2] if t2 in (3, 5) else s2)
        sep = rng.uniform(0.0, 1.3) * reach  # centre distance from fully overlapping to clearly separated
        dirn = rng.normal(size=3)
        p1 = rng.uniform(-0.5, 0.5, 3)
        p2 = p1 + sep * dirn / np.linalg.norm(dirn)
        margin = 0.0 if k % 3 else rng.uniform(0.0, 0.05)
        cutoff = 0.0 if k % 2 else 1.0e32
        g = []
        for (p, r, s) in ((p1, r1, s1), (p2, r2, s2)):
          geom = core.Geom()
          geom.pos, geom.rot, geom.size = wp.vec3(*p), wp.mat33(*r.reshape(-1)), wp.vec3(*s)
          geom.margin, geom.index = float(margin), -1
          g.append(geom)
        out = gjk.ccd(TOL, cutoff, ITER, ITER, g[0], g[1], t1, t2, g[0].pos, g[1].pos,
                      wp.zeros(nv, dtype=wp.vec3), wp.zeros(nv, dtype=int), wp.zeros(nf, dtype=int), wp.zeros(nf, dtype=wp.vec3),
                      wp.zeros(nf, dtype=float), wp.zeros(24, dtype=int))  # fmt: skip
        dist, ncol, x1, x2, _ = out
        for key, v in zip(rec, (t1, t2, p1, r1, s1, p2, r2, s2, margin, cutoff, int(ncol), float(dist), np.asarray(x1), np.asarray(x2))):
          rec[key].append(v)
  arrs = {k: np.asarray(v) for k, v in rec.items()}
  np.savez_compressed(OUT, **arrs)
  d = arrs["dist"]
  print("wrote", OUT, len(d), "pairs;", int((d < 0).sum()), "penetrating,", int((d > 1e29).sum()), "early-out,", int(((d >= 0) & (d < 1e29)).sum()), "separated with distance")


OUT_MESH = os.path.join(HERE, "ccd_mesh_pairs.npz")
MAXV = 20


def rand_hull(rng):
  """Vertices of a random convex polytope (points on an ellipsoid are all hull vertices)."""
  n = int(rng.integers(4, MAXV + 1))
  v = rng.normal(size=(n, 3))
  v = v / np.linalg.norm(v, axis=1, keepdims=True) * rng.uniform(0.1, 0.35, 3)
  return v + rng.uniform(-0.05, 0.05, 3)  # the mesh frame is not centred on the hull


def main_mesh(n_per_pair=30, seed=20260303):
  wp, ref = R.load_reference()
  gjk = sys.modules["mujoco_warp._src.collision_gjk"]
  core = sys.modules["mujoco_warp._src.collision_core"]
  rng = np.random.default_rng(seed)
  rec = {k: [] for k in "t1 t2 p1 r1 s1 v1 nv1 p2 r2 s2 v2 nv2 margin cutoff ncollision dist x1 x2".split()}
  nv, nf = 10 + 2 * ITER, 6 + 5 * ITER
  for t1 in TYPES + [7]:
    t2 = 7
    for k in range(n_per_pair):
      shapes = []
      for t in (t1, t2):
        if t == 7:
          v = rand_hull(rng)
          lo, hi = v.min(0), v.max(0)
          shapes.append((0.5 * (hi - lo), v))  # geom_size of a mesh geom: half extents of its bounding box
        else:
          shapes.append((rand_size(rng, t), np.zeros((0, 3))))
      (s1, v1), (s2, v2) = shapes
      r1, r2 = rand_rot(rng), rand_rot(rng)
      reach = sum(np.linalg.norm(v, axis=1).max() if len(v) else np.linalg.norm(s[:2] if t in (3, 5) else s) for (s, v), t in zip(shapes, (t1, t2)))
      sep = rng.uniform(0.0, 1.3) * reach
      dirn = rng.normal(size=3)
      p1 = rng.uniform(-0.5, 0.5, 3)
      p2 = p1 + sep * dirn / np.linalg.norm(dirn)
      margin = 0.0 if k % 3 else rng.uniform(0.0, 0.05)
      cutoff = 0.0 if k % 2 else 1.0e32
      g = []
      for p, r, s, v in ((p1, r1, s1, v1), (p2, r2, s2, v2)):
        geom = core.Geom()
        geom.pos, geom.rot, geom.size = wp.vec3(*p), wp.mat33(*r.reshape(-1)), wp.vec3(*s)
        geom.margin, geom.index = float(margin), -1
        geom.vertadr, geom.vertnum, geom.graphadr = 0, len(v), -1
        geom.vert = wp.array(v.reshape(-1, 3), dtype=wp.vec3)
        geom.graph = wp.zeros(0, dtype=int)
        g.append(geom)
      out = gjk.ccd(TOL, cutoff, ITER, ITER, g[0], g[1], t1, t2, g[0].pos, g[1].pos,
                    wp.zeros(nv, dtype=wp.vec3), wp.zeros(nv, dtype=int), wp.zeros(nf, dtype=int), wp.zeros(nf, dtype=wp.vec3),
                    wp.zeros(nf, dtype=float), wp.zeros(24, dtype=int))  # fmt: skip
      dist, ncol, x1, x2, _ = out
      pad = lambda v: np.vstack([v, np.zeros((MAXV - len(v), 3))])  # noqa: E731
      vals = (t1, t2, p1, r1, s1, pad(v1), len(v1), p2, r2, s2, pad(v2), len(v2), margin, cutoff, int(ncol), float(dist), np.asarray(x1), np.asarray(x2))
      for key, val in zip(rec, vals):
        rec[key].append(val)
  arrs = {k: np.asarray(v) for k, v in rec.items()}
  np.savez_compressed(OUT_MESH, **arrs)
  d = arrs["dist"]
  print("wrote", OUT_MESH, len(d), "pairs;", int((d < 0).sum()), "penetrating,", int((d > 1e29).sum()), "early-out,", int(((d >= 0) & (d < 1e29)).sum()), "separated with distance")


if __name__ == "__main__":
  main_mesh() if "--mesh" in sys.argv else main()
