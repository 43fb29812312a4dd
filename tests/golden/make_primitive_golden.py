"""Golden vectors for the primitive narrowphase, produced by EXECUTING the reference's own functions.

`/root/reference/mujoco_warp/_src/collision_primitive_core.py` imports nothing but `warp`; with the numpy stand-in of
`warp_shim.py` registered as `warp` the file loads unmodified and its `@wp.func`s run as plain Python (float64).  This
script calls them on seeded random geometry and stores inputs and outputs in `primitive_core.npz`; the oracle is
checked against the file in tests/test_oracle_golden.py.  Run here (the reference tree does not exist on the GPU box):

  python tests/golden/make_primitive_golden.py
"""

import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import warp_shim  # noqa: E402

REF = "/root/reference/mujoco_warp/_src/collision_primitive_core.py"
PLANE, SPHERE, CAPSULE, BOX = 0, 2, 3, 6  # mjtGeom


def load_reference():
  wp = warp_shim.install()
  spec = importlib.util.spec_from_file_location("ref_collision_primitive_core", REF)
  mod = importlib.util.module_from_spec(spec)
  spec.loader.exec_module(mod)
  return wp, mod


def rand_rot(rng, aligned=False):
  if aligned:  # signed axis permutations: the degenerate (parallel) configurations
    perm = rng.permutation(3)
    m = np.zeros((3, 3))
    for i in range(3):
      m[i, perm[i]] = rng.choice([-1.0, 1.0])
    if np.linalg.det(m) < 0:
      m[0] *= -1
    return m
  q = rng.normal(size=4)
  w, x, y, z = q / np.linalg.norm(q)
  return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                   [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                   [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])  # fmt: skip


def main(n_per_type=160, seed=20260101):
  wp, ref = load_reference()
  rng = np.random.default_rng(seed)
  rec = {k: [] for k in "t1 t2 p1 r1 s1 p2 r2 s2 margin ncand dist pos normal".split()}

  def add(t1, t2, p1, r1, s1, p2, r2, s2, margin, dist, pos, normal):
    dist = np.atleast_1d(np.asarray(dist, dtype=np.float64))
    n = dist.size
    d8, p8, n8 = np.full(8, 1e10), np.zeros((8, 3)), np.zeros((8, 3))
    d8[:n] = dist
    p8[:n] = np.asarray(pos, dtype=np.float64).reshape(n, 3)
    nn = np.asarray(normal, dtype=np.float64)
    n8[:n] = nn.reshape(n, 3) if nn.size == 3 * n else np.broadcast_to(nn.reshape(-1, 3)[0], (n, 3))
    for k, v in zip(rec, (t1, t2, p1, r1, s1, p2, r2, s2, margin, n, d8, p8, n8)):
      rec[k].append(np.asarray(v, dtype=np.float64))

  V, M = wp.vec3, wp.mat33
  for case in range(n_per_type):
    aligned = case % 8 == 7
    spread = 0.25 if case % 2 else 0.45
    p1, p2 = rng.uniform(-spread, spread, 3), rng.uniform(-spread, spread, 3)
    r1, r2 = rand_rot(rng, aligned), rand_rot(rng, aligned and case % 16 == 15)
    if case % 16 == 7:
      r2 = r1.copy()  # exactly parallel axes
    margin = float(rng.choice([0.0, 0.02, 0.3]))
    rad1, rad2 = rng.uniform(0.05, 0.15, 2)
    hl1, hl2 = rng.uniform(0.1, 0.3, 2)
    b1, b2 = rng.uniform(0.08, 0.3, 3), rng.uniform(0.08, 0.3, 3)
    z1, z2 = r1[:, 2], r2[:, 2]
    sph1, sph2 = np.array([rad1, 0, 0]), np.array([rad2, 0, 0])
    cap1, cap2 = np.array([rad1, hl1, 0]), np.array([rad2, hl2, 0])
    zero = np.zeros(3)

    d, p = ref.plane_sphere(V(z1), V(p1), V(p2), rad2)
    add(PLANE, SPHERE, p1, r1, zero, p2, r2, sph2, margin, d, p, z1)
    d, p, fr = ref.plane_capsule(V(z1), V(p1), V(p2), V(z2), rad2, hl2)
    add(PLANE, CAPSULE, p1, r1, zero, p2, r2, cap2, margin, d, p, np.asarray(fr)[0])
    d, p, nrm = ref.plane_box(V(z1), V(p1), V(p2), M(r2), V(b2))
    add(PLANE, BOX, p1, r1, zero, p2, r2, b2, margin, d, p, nrm)
    d, p, nrm = ref.sphere_sphere(V(p1), rad1, V(p2), rad2)
    add(SPHERE, SPHERE, p1, r1, sph1, p2, r2, sph2, margin, d, p, nrm)
    d, p, nrm = ref.sphere_capsule(V(p1), rad1, V(p2), V(z2), rad2, hl2)
    add(SPHERE, CAPSULE, p1, r1, sph1, p2, r2, cap2, margin, d, p, nrm)
    d, p, nrm = ref.capsule_capsule(V(p1), V(z1), rad1, hl1, V(p2), V(z2), rad2, hl2, margin)
    add(CAPSULE, CAPSULE, p1, r1, cap1, p2, r2, cap2, margin, d, p, nrm)
    d, p, nrm = ref.sphere_box(V(p1), rad1, V(p2), M(r2), V(b2))
    add(SPHERE, BOX, p1, r1, sph1, p2, r2, b2, margin, d, p, nrm)
    d, p, nrm = ref.capsule_box(V(p1), V(z1), rad1, hl1, V(p2), M(r2), V(b2))
    add(CAPSULE, BOX, p1, r1, cap1, p2, r2, b2, margin, d, p, nrm)
    d, p, nrm = ref.box_box(V(p1), M(r1), V(b1), V(p2), M(r2), V(b2), margin)
    add(BOX, BOX, p1, r1, b1, p2, r2, b2, margin, d, p, nrm)

  out = {k: np.stack(v) for k, v in rec.items()}
  for k in ("t1", "t2", "ncand"):
    out[k] = out[k].astype(np.int32)
  path = os.path.join(HERE, "primitive_core.npz")
  np.savez_compressed(path, **out)
  print("wrote", path, {k: v.shape for k, v in out.items()}, os.path.getsize(path), "bytes")


if __name__ == "__main__":
  main()
