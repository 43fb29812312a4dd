"""Validates the CPU oracle through physics invariants (the substitute for a live MuJoCo C
comparison, which is unavailable here) and against the committed golden fixture."""

import ctypes
import os

import numpy as np
import pytest

import mujoco_warp_b200 as mjw
from mujoco_warp_b200 import constants as C
from oracle import oracle as O
from tests import helpers as H

GOLDEN = os.path.join(H.ROOT, "tests", "golden", "humanoid_forward_step.npz")
NW = 6


@pytest.fixture(scope="module")
def fwd(humanoid):
  qpos, qvel, ctrl = H.make_states(humanoid, NW, seed=7)
  om, d = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, d, "forward")
  return om, d


def dense_m(d, w, nv):
  return np.array(d.qM[w, :nv, :nv])


def test_mass_matrix_symmetric_positive_definite(fwd):
  om, d = fwd
  nv = om.sizes["nv"]
  for w in range(NW):
    M = dense_m(d, w, nv)
    np.testing.assert_allclose(M, M.T, atol=1e-12)
    assert np.linalg.eigvalsh(M).min() > 0


def test_cholesky_factor_reproduces_mass_matrix(fwd):
  om, d = fwd
  nv = om.sizes["nv"]
  for w in range(NW):
    L = np.tril(np.array(d.qLD[w]).reshape(nv, nv))
    np.testing.assert_allclose(L @ L.T, dense_m(d, w, nv), atol=1e-10)


def test_smooth_acceleration_solves_m_qacc_eq_qfrc(fwd):
  om, d = fwd
  nv = om.sizes["nv"]
  for w in range(NW):
    np.testing.assert_allclose(dense_m(d, w, nv) @ d.qacc_smooth[w], d.qfrc_smooth[w], atol=1e-9)


def test_mul_m_and_solve_m_helpers_match_dense_algebra(fwd):
  om, d = fwd
  lib = O.load()
  nv, nvp = om.sizes["nv"], om.sizes["nv_pad"]
  rng = np.random.default_rng(0)
  P = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))  # noqa: E731
  for w in range(NW):
    x = rng.normal(size=nv)
    res, sol = np.zeros(nv), np.zeros(nv)
    qM = np.ascontiguousarray(d.qM[w]).reshape(-1)
    lib.mjo_mul_m(nv, nvp, P(qM), P(x), P(res))
    np.testing.assert_allclose(res, dense_m(d, w, nv) @ x, atol=1e-10)
    qLD = np.ascontiguousarray(d.qLD[w]).reshape(-1)
    lib.mjo_solve_m(nv, P(qLD), P(res), P(sol))
    np.testing.assert_allclose(sol, x, atol=1e-8)


def test_contact_frames_are_orthonormal_and_right_handed(fwd):
  _, d = fwd
  n = int(d.nacon[0])
  assert n > 0
  F = np.array(d.contact_frame[:n]).reshape(n, 3, 3)
  for f in F:
    np.testing.assert_allclose(f @ f.T, np.eye(3), atol=1e-12)
    np.testing.assert_allclose(np.cross(f[0], f[1]), f[2], atol=1e-12)


def test_contact_rows_are_consistent_with_contacts(fwd):
  om, d = fwd
  for w in range(NW):
    nefc, ne, nf, nl = int(d.nefc[w]), int(d.ne[w]), int(d.nf[w]), int(d.nl[w])
    ncon_rows = nefc - ne - nf - nl
    ids = d.wcon[w, : int(d.wcon_n[w])]
    expect = 0
    for c in ids:
      dim = int(d.contact_dim[c])
      nrow = 1 if dim == 1 else 2 * (dim - 1)  # pyramidal cone, constraint.py:1669
      active = d.contact_dist[c] < d.contact_includemargin[c]
      adr = d.contact_efc_address[c]
      if active:
        expect += nrow
        assert (adr[:nrow] >= 0).all()
        # contact row position is the contact distance (margin bookkeeping: pos = dist - margin + margin)
        np.testing.assert_allclose(d.efc_pos[w, adr[:nrow]], d.contact_dist[c], atol=1e-12)
      else:
        assert (adr == -1).all()
    assert ncon_rows == expect


def test_solver_satisfies_kkt_conditions(fwd):
  """At the solver's fixed point M qacc - qfrc_smooth = J^T force with force from the row state machine."""
  om, d = fwd
  nv = om.sizes["nv"]
  scale = 1.0 / (om.meaninertia * nv)
  for w in range(NW):
    nefc = int(d.nefc[w])
    J = np.array(d.efc_J[w, :nefc, :nv])
    f = np.array(d.efc_force[w, :nefc])
    res = dense_m(d, w, nv) @ d.qacc[w] - d.qfrc_smooth[w] - J.T @ f
    assert np.linalg.norm(res) * scale < 1e-5
    # complementarity of the inequality rows: force >= 0 and zero where the row is satisfied
    ineq = slice(int(d.ne[w]) + int(d.nf[w]), nefc)
    assert (f[ineq] >= -1e-12).all()
    jar = J @ d.qacc[w] - d.efc_aref[w, :nefc]
    sat = jar[ineq] >= 0
    np.testing.assert_allclose(f[ineq][sat], 0.0, atol=1e-12)
    np.testing.assert_allclose(d.qfrc_constraint[w], J.T @ f, atol=1e-9)


def test_rne_bias_matches_mass_matrix_at_zero_acceleration(humanoid):
  """qfrc_bias(q, v=0) is pure gravity: -J_com^T m g, i.e. its root-translation entries are 0, 0, m g."""
  qpos, qvel, ctrl = H.make_states(humanoid, 2, seed=3)
  qvel[:] = 0
  om, d = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, d, "forward")
  mass = float(np.sum(humanoid.body_mass))
  np.testing.assert_allclose(d.qfrc_bias[:, :3], np.array([[0, 0, 9.81 * mass]] * 2), atol=1e-9)


def test_free_fall_without_contacts_conserves_horizontal_momentum_and_accelerates_at_g(humanoid):
  qpos = np.array(humanoid.key_qpos[0])[None].copy()
  qpos[0, 2] += 5.0  # far above the floor
  om, d = H.oracle_data(humanoid, qpos, np.zeros((1, humanoid.nv)), np.zeros((1, humanoid.nu)), 24, 64)
  O.run(om, d, "forward")
  assert int(d.nacon[0]) == 0
  com0 = np.array(d.subtree_com[0, 1])
  for _ in range(20):
    O.run(om, d, "step")
  O.run(om, d, "forward")
  com1 = np.array(d.subtree_com[0, 1])
  t = 20 * om.opt["timestep"]
  # internal joint forces cannot move the centre of mass (up to the O(dt^2) error of the integrator)
  np.testing.assert_allclose(com1[:2], com0[:2], atol=1e-4)
  # semi-implicit Euler: z(t) = z0 - g dt^2 n(n+1)/2
  np.testing.assert_allclose(com1[2] - com0[2], -9.81 * om.opt["timestep"] ** 2 * 20 * 21 / 2, rtol=5e-3)
  assert abs(float(d.time[0]) - t) < 1e-12


def test_oracle_reproduces_golden_fixture(humanoid):
  g = np.load(GOLDEN)
  om, d = H.oracle_data(humanoid, g["qpos0"], g["qvel0"], g["ctrl"], int(g["nconmax"]), int(g["njmax"]))
  O.run(om, d, "forward")
  for f in ["qM", "qacc_smooth", "efc_J", "efc_D", "efc_aref", "efc_force", "qacc"]:
    np.testing.assert_allclose(getattr(d, f), g["fwd_" + f], rtol=1e-9, atol=1e-9, err_msg=f)
  for f in ["nefc", "ne", "nf", "nl", "wcon_n"]:
    np.testing.assert_array_equal(getattr(d, f), g["fwd_" + f])
  om, d = H.oracle_data(humanoid, g["qpos0"], g["qvel0"], g["ctrl"], int(g["nconmax"]), int(g["njmax"]))
  for k in range(g["traj_qpos"].shape[0]):
    O.run(om, d, "step")
    np.testing.assert_allclose(d.qpos, g["traj_qpos"][k], rtol=1e-8, atol=1e-8)
    np.testing.assert_array_equal(d.nefc, g["traj_nefc"][k])


def test_oracle_threads_and_world_ranges_are_equivalent(humanoid):
  """Running [0, n) in one call, in two halves, or with several OpenMP threads gives identical per-world results."""
  qpos, qvel, ctrl = H.make_states(humanoid, 8, seed=11)
  om, a = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, a, "forward")
  _, b = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, b, "forward", nthread=4)
  for f in ["qacc", "efc_force", "nefc", "qM"]:
    np.testing.assert_array_equal(getattr(a, f), getattr(b, f))
  # contact pool order may differ between thread schedules, per-world content may not
  for w in range(8):
    np.testing.assert_allclose(np.sort(H.world_contacts_orc(a, w, "contact_dist")), np.sort(H.world_contacts_orc(b, w, "contact_dist")))


def test_cg_and_newton_agree(humanoid):
  qpos, qvel, ctrl = H.make_states(humanoid, 4, seed=5)
  om, a = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, a, "forward", opt={"iterations": 100, "tolerance": 1e-10})
  _, b = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, b, "forward", opt={"solver": 1, "iterations": 500, "tolerance": 1e-10})
  np.testing.assert_allclose(a.qacc, b.qacc, rtol=1e-4, atol=1e-4)


# ---- box primitives (collision_primitive_core.py:1103 sphere_box, :1158 capsule_box) against sampled geometry ------
BOX_XML = """
<mujoco>
  <worldbody>
    <body pos="0 0 0"><freejoint/><geom type="box" size=".2 .3 .15" margin="1"/></body>
    <body pos="1 0 0"><freejoint/><geom type="sphere" size=".1" margin="1"/></body>
    <body pos="0 1 0"><freejoint/><geom type="capsule" size=".07 .25" margin="1"/></body>
  </worldbody>
</mujoco>"""


def _box_distance(p, bp, rot, size):
  c = rot.T @ (p - bp)
  if np.all(np.abs(c) < size):
    return -np.min(size - np.abs(c))
  return np.linalg.norm(np.clip(c, -size, size) - c)


def test_sphere_box_and_capsule_box_distances_match_sampled_geometry():
  from tests.test_oracle_kat import run_collision

  mjm = mjw.load_mjcf_string(BOX_XML)
  rng = np.random.default_rng(0)
  nw = 200
  qpos = np.zeros((nw, 21))
  for b in range(3):
    qpos[:, 7 * b : 7 * b + 3] = rng.uniform(-0.45, 0.45, (nw, 3))
    q = rng.normal(size=(nw, 4))
    qpos[:, 7 * b + 3 : 7 * b + 7] = q / np.linalg.norm(q, axis=1, keepdims=True)
  _, d = run_collision(mjm, qpos, nconmax=16)
  size, ts = np.array([0.2, 0.3, 0.15]), np.linspace(-1, 1, 2001)
  nsphere = ncap = nsecond = 0
  for w in range(nw):
    xpos, xmat = d.geom_xpos[w].reshape(-1, 3), d.geom_xmat[w].reshape(-1, 3, 3)
    first = None
    for k in range(int(d.wcon_n[w])):
      cid = int(d.wcon[w, k])
      g, dist = tuple(int(x) for x in d.contact_geom[cid]), float(d.contact_dist[cid])
      frame = d.contact_frame[cid].reshape(3, 3)
      np.testing.assert_allclose(frame @ frame.T, np.eye(3), atol=1e-9)
      if g == (1, 0):  # sphere (type 2) sorts before box (type 6)
        assert abs(_box_distance(xpos[1], xpos[0], xmat[0], size) - 0.1 - dist) < 1e-9
        nsphere += 1
      elif g == (2, 0):
        dd = np.array([_box_distance(xpos[2] + xmat[2][:, 2] * 0.25 * t, xpos[0], xmat[0], size) for t in ts])
        if int(d.contact_geomcollisionid[cid]) == 0:
          first = dist
          if dd.min() > 0:  # segment outside the box: the first contact is the closest approach of the segment
            assert abs(dd.min() - 0.07 - dist) < 5e-4
          ncap += 1
        else:
          nsecond += 1
          assert dist >= dd.min() - 0.07 - 1e-9  # no point of the segment is closer than the closest one
          assert first is not None
  assert nsphere > 100 and ncap > 100 and nsecond > 20


# ---- elliptic friction cones (constraint.py:1940, solver.py:263, :1886, :2592) ---------------------------------------
def test_elliptic_line_cost_derivatives_are_consistent():
  """_eval_elliptic returns (cost, d cost / d alpha, d2 cost / d alpha2): checked by central differences in every zone."""
  import ctypes

  lib = O.load()
  lib.mjo_eval_elliptic.argtypes = [ctypes.c_double] + [ctypes.POINTER(ctypes.c_double)] * 3 + [ctypes.c_double, ctypes.POINTER(ctypes.c_double)]
  lib.mjo_eval_elliptic.restype = None

  def ev(mu, quad, q1, q2, alpha):
    out = (ctypes.c_double * 3)()
    arr = lambda v: (ctypes.c_double * 3)(*v)  # noqa: E731
    lib.mjo_eval_elliptic(mu, arr(quad), arr(q1), arr(q2), alpha, out)
    return np.array(out[:])

  rng = np.random.default_rng(0)
  zones = set()
  for _ in range(300):
    mu, D0, D1, D2 = rng.uniform(0.3, 1.5), *rng.uniform(50, 500, 3)
    fr = np.array([mu, rng.uniform(0.3, 1.5)])
    ja, jv = rng.normal(0, 1, 3), rng.normal(0, 1, 3)
    D = np.array([D0, D1, D2])
    quad = [np.sum(0.5 * ja * ja * D), np.sum(jv * D * ja), np.sum(0.5 * jv * jv * D)]
    u, v = ja[1:] * fr, jv[1:] * fr
    q1, q2 = [ja[0] * mu, jv[0] * mu, u @ u], [u @ v, v @ v, D0 / (mu * mu * (1 + mu * mu))]
    alpha, h = rng.uniform(-0.5, 0.5), 1e-6
    c0, cp, cm = ev(mu, quad, q1, q2, alpha), ev(mu, quad, q1, q2, alpha + h), ev(mu, quad, q1, q2, alpha - h)
    N, T = q1[0] + alpha * q1[1], np.sqrt(q1[2] + alpha * (2 * q2[0] + alpha * q2[1]))
    zone = "top" if N >= mu * T else ("bottom" if mu * N + T <= 0 else "middle")
    if min(abs(N - mu * T), abs(mu * N + T)) < 1e-3:
      continue  # finite differences straddle a zone boundary
    zones.add(zone)
    assert abs((cp[0] - cm[0]) / (2 * h) - c0[1]) < 1e-5 * (1 + abs(c0[1])), zone
    assert abs((cp[1] - cm[1]) / (2 * h) - c0[2]) < 1e-4 * (1 + abs(c0[2])), zone
    if zone == "top":
      assert not c0.any()
  assert zones == {"top", "bottom", "middle"}


def test_elliptic_solution_lies_in_the_friction_cone(humanoid):
  """Converged forces of every elliptic contact satisfy f_n >= 0 and sum_j (f_j / mu_j)^2 <= f_n^2, rows per contact are
  condim (not 2 (condim - 1)), and the stationarity residual M a - qfrc_smooth - J^T f vanishes."""
  qpos, qvel, ctrl = H.make_states(humanoid, 48, seed=1)
  qpos[:, 2] -= 0.12
  om, d = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, d, "forward", opt={"cone": int(C.ConeType.ELLIPTIC)}, nthread=4)
  _, dp = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, dp, "forward", nthread=4)
  states = set()
  for cid in range(int(d.nacon[0])):
    w, adr, dim = int(d.contact_worldid[cid]), d.contact_efc_address[cid], int(d.contact_dim[cid])
    if adr[0] < 0:
      continue
    assert (np.diff(adr[:dim]) == 1).all() and (adr[dim:] == -1).all()
    if dim == 1:
      assert int(d.efc_type[w][adr[0]]) == int(C.ConstraintType.CONTACT_FRICTIONLESS)
      continue
    assert (d.efc_type[w][adr[:dim]] == int(C.ConstraintType.CONTACT_ELLIPTIC)).all()
    f, fri = d.efc_force[w], d.contact_friction[cid]
    fn, ft = f[adr[0]], np.sqrt(sum((f[adr[j]] / fri[j - 1]) ** 2 for j in range(1, dim)))
    states.add(int(d.efc_state[w][adr[0]]))
    assert fn >= -1e-9 and ft <= fn * (1 + 1e-6) + 1e-9
  assert states == {0, 1, 4}  # satisfied, quadratic (sticking), cone (sliding)
  nv = humanoid.nv
  for w in range(48):
    n = int(d.nefc[w])
    J = d.efc_J[w].reshape(64, -1)[:n, :nv]
    res = d.efc_Ma[w] - d.qfrc_smooth[w] - J.T @ d.efc_force[w][:n]
    assert np.abs(res).max() < 1e-3 * (1 + np.abs(d.qfrc_smooth[w]).max())
  # the pyramid is a linearisation of the same cone: accelerations stay close
  assert np.abs(d.qacc - dp.qacc).max() < 0.2 * np.abs(dp.qacc).max()
  assert int(d.solver_niter.max()) < 30


def test_connect_and_weld_equalities_hold_the_bodies_together():
  """Constraint-level property of the connect / weld rows (constraint.py:125, :793): efc_pos is the world-space error of the
  anchors (and of the relative orientation), zero at qpos0 by construction of eq_data (mj_setConst), efc_J is its derivative
  (finite differences through the kinematics), and over 0.6 s of gravity the welded / pinned bodies stay where the equalities hold
  them (a free body would have dropped 1.8 m)."""
  from tests import models as TM

  mjm = mjw.load_mjcf_string(TM.SMALL_MODELS["connect_weld"])
  om = O.OracleModel(mjm)
  d = O.make_data(om, nworld=1, njmax=48)
  O.run(om, d, "forward")
  ne = int(d.ne[0])
  assert ne == 3 + 6 + 3 + 6 + 1  # the inactive weld emits no rows
  assert np.abs(d.efc_pos[0, :ne]).max() < 1e-12
  # derivative check on the hinge / ball / slide coordinates (quaternion coordinates are left alone)
  rng = np.random.default_rng(3)
  q0 = np.array(d.qpos[0])
  hinge = [int(mjm.jnt_qposadr[j]) for j in range(mjm.njnt) if int(mjm.jnt_type[j]) in (2, 3)]
  dofs = [int(mjm.jnt_dofadr[j]) for j in range(mjm.njnt) if int(mjm.jnt_type[j]) in (2, 3)]
  q0[hinge] += rng.uniform(-0.2, 0.2, len(hinge))
  d.qpos[0] = q0
  O.run(om, d, "forward")
  J, pos0 = np.array(d.efc_J[0, :ne, : mjm.nv]), np.array(d.efc_pos[0, :ne])
  eps = 1e-6
  for qa, da in zip(hinge, dofs):
    d.qpos[0] = q0
    d.qpos[0, qa] += eps
    O.run(om, d, "forward")
    fd = (np.array(d.efc_pos[0, :ne]) - pos0) / eps
    rows = slice(0, 18)  # connect / weld rows: their positional rows are exact derivatives; weld rotation rows to first order
    assert np.abs(fd[rows] - J[rows, da]).max() < 2e-2 * (1 + np.abs(J[rows, da]).max()), (qa, np.abs(fd[rows] - J[rows, da]).max())
  d.qpos[0] = q0
  d.qvel[:] = 0
  O.run(om, d, "forward")
  for _ in range(300):
    O.run(om, d, "step")
  O.run(om, d, "forward")
  assert np.abs(d.efc_pos[0, : int(d.ne[0])]).max() < 0.3
  f1, f2, p1 = (mjm.body_names.index(n) for n in ("f1", "f2", "p1"))
  assert abs(d.xpos[0, p1, 2] - 1.0) < 0.2  # pinned to the world
  assert d.xpos[0, f1, 2] < -0.5 and abs(np.linalg.norm(d.xpos[0, f1] - d.xpos[0, f2]) - np.linalg.norm([0.3, 0.0, 0.1])) < 0.02  # fall together
