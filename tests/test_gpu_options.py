"""CUDA path vs oracle under non-default options, on small hand-written models that reach the row types the
humanoid does not (joint equality, dof friction, slide limits, plane-box, sphere-capsule), at the edges
(overflowing njmax / nconmax, nworld 1, worlds not a multiple of the CTA width, no constraints at all),
and size-independent properties at the benchmark's full size (nworld 8192)."""

import ctypes

import numpy as np
import pytest
import torch

import mujoco_warp_b200 as mjw
from mujoco_warp_b200 import constants as C
from tests import helpers as H
from tests import models as TM

pytestmark = pytest.mark.gpu
DB = C.DisableBit


def run_both(mjm, qpos, qvel, ctrl, nconmax, njmax, stage="forward", opt=None, nstep=1):
  from oracle import oracle as O

  om, od = H.oracle_data(mjm, qpos, qvel, ctrl, nconmax, njmax)
  m, d = H.device_data(mjw, mjm, qpos, qvel, ctrl, nconmax, njmax)
  for k, v in (opt or {}).items():
    setattr(m.opt, k, v)
  ostage = "step_implicitfast" if stage == "step" and (opt or {}).get("integrator") == int(C.IntegratorType.IMPLICITFAST) else stage
  for _ in range(nstep):
    O.run(om, od, ostage, opt=opt, nthread=4)
    (mjw.forward if stage == "forward" else mjw.step)(m, d)
  torch.cuda.synchronize()
  return om, od, m, d


def assert_forward_close(od, d, tol_solver=2e-2, tol_smooth=5e-5):
  for f in ["ne", "nf", "nl", "nefc", "wcon_n", "nacon"]:
    np.testing.assert_array_equal(H.dev_np(d, f), getattr(od, f), err_msg=f)
  for f in ["qfrc_bias", "qfrc_passive", "qfrc_actuator", "qfrc_smooth"]:
    assert H.max_err(H.dev_np(d, f), getattr(od, f)) < tol_smooth, f
  if tol_solver is None:
    return
  for f in ["qacc", "qfrc_constraint"]:
    assert H.max_err(H.dev_np(d, f), getattr(od, f)) < tol_solver, f


@pytest.mark.parametrize(
  "flags",
  [DB.CONTACT, DB.LIMIT, DB.CONSTRAINT, DB.GRAVITY, DB.WARMSTART, DB.ACTUATION, DB.CLAMPCTRL, DB.SPRING | DB.DAMPER, DB.EULERDAMP, DB.REFSAFE,
   DB.FILTERPARENT],
)  # fmt: skip
def test_disable_flags(humanoid, flags):
  qpos, qvel, ctrl = H.make_states(humanoid, 16, seed=10)
  ctrl *= 3.0  # exercise the control clamp
  _, od, _, d = run_both(humanoid, qpos, qvel, ctrl, 24, 64, stage="step", opt={"disableflags": int(flags)}, nstep=2)
  # after stepping, the float32 state itself has drifted from the float64 one (qvel ~1e-3)
  assert_forward_close(od, d, tol_solver=5e-2, tol_smooth=2e-3)
  assert H.max_err(H.dev_np(d, "qpos"), od.qpos) < 1e-5
  assert H.max_err(H.dev_np(d, "qvel"), od.qvel) < 5e-3


def test_cg_solver(humanoid):
  qpos, qvel, ctrl = H.make_states(humanoid, 16, seed=11)
  opt = {"solver": int(C.SolverType.CG), "iterations": 100}
  _, od, _, d = run_both(humanoid, qpos, qvel, ctrl, 24, 64, opt=opt)
  assert_forward_close(od, d, tol_solver=5e-2)


@pytest.mark.parametrize("iterations,ls_iterations", [(1, 50), (3, 2), (100, 50)])
def test_iteration_limits(humanoid, iterations, ls_iterations):
  qpos, qvel, ctrl = H.make_states(humanoid, 16, seed=12)
  _, od, _, d = run_both(humanoid, qpos, qvel, ctrl, 24, 64, opt={"iterations": iterations, "ls_iterations": ls_iterations})
  assert int(H.dev_np(d, "solver_niter").max()) <= iterations
  # a truncated solve is not a fixed point: float32 and float64 line searches may stop at different alphas,
  # so the solver outputs are only compared when the iteration budget lets both converge
  assert_forward_close(od, d, tol_solver=5e-2 if iterations >= 100 else None)
  assert not torch.isnan(d.qacc).any()


SMALL_MODELS = TM.SMALL_MODELS


@pytest.mark.parametrize("name", list(SMALL_MODELS))
def test_small_models(name):
  mjm = mjw.load_mjcf_string(SMALL_MODELS[name])
  rng = np.random.default_rng(5)
  nw = 8
  qpos = np.repeat(np.asarray(mjm.qpos0, dtype=np.float64)[None], nw, axis=0)
  for j in range(mjm.njnt):
    jt, qa = int(mjm.jnt_type[j]), int(mjm.jnt_qposadr[j])
    if jt in (2, 3):
      qpos[:, qa] += rng.uniform(-0.3, 0.3, nw)
    elif jt == 0:
      qpos[:, qa + 2] += rng.uniform(-0.02, 0.02, nw)
  qvel = rng.uniform(-0.5, 0.5, (nw, mjm.nv))
  ctrl = rng.uniform(-2, 2, (nw, mjm.nu))
  _, od, _, d = run_both(mjm, qpos, qvel, ctrl, 32, 64, stage="step", nstep=3)
  assert_forward_close(od, d, tol_solver=5e-2, tol_smooth=2e-3)
  assert H.max_err(H.dev_np(d, "qpos"), od.qpos) < 1e-4
  if name == "equality_friction_limits":
    assert int(od.ne.max()) == 1 and int(od.nf.max()) >= 1
  if name == "box_and_spheres_on_plane":
    assert int(od.nacon[0]) > 0


@pytest.mark.parametrize("name", ["humanoid", "equality_friction_limits"])
def test_implicitfast_integrator(humanoid, name):
  """forward.py:495 implicit + derivative.py:321 deriv_smooth_vel (actuator velocity gains, joint damping)."""
  mjm = humanoid if name == "humanoid" else mjw.load_mjcf_string(SMALL_MODELS[name])
  rng = np.random.default_rng(21)
  if name == "humanoid":
    qpos, qvel, ctrl = H.make_states(mjm, 8, seed=21)
  else:
    qpos = np.repeat(np.asarray(mjm.qpos0, dtype=np.float64)[None], 8, axis=0) + rng.uniform(-0.2, 0.2, (8, mjm.nq))
    qvel, ctrl = rng.uniform(-1, 1, (8, mjm.nv)), rng.uniform(-1, 1, (8, mjm.nu))
  opt = {"integrator": int(C.IntegratorType.IMPLICITFAST)}
  _, od, _, d = run_both(mjm, qpos, qvel, ctrl, 24, 64, stage="step", opt=opt, nstep=3)
  assert H.max_err(H.dev_np(d, "qpos"), od.qpos) < 1e-5
  assert H.max_err(H.dev_np(d, "qvel"), od.qvel) < 5e-3
  # and it differs from Euler where actuators have velocity gains
  if name != "humanoid":
    _, oe, _, _ = run_both(mjm, qpos, qvel, ctrl, 24, 64, stage="step", nstep=3)
    assert np.abs(oe.qvel - od.qvel).max() > 1e-4


def test_njmax_overflow_sets_status_and_clamps(humanoid):
  qpos, qvel, ctrl = H.make_states(humanoid, 8, seed=13)
  qpos[:, 2] -= 0.2  # push into the floor: many contacts
  _, od, _m, d = run_both(humanoid, qpos, qvel, ctrl, 24, 8)
  np.testing.assert_array_equal(H.dev_np(d, "nefc"), od.nefc)  # the attempted row count is reported, as in the oracle
  assert int(od.nefc.max()) > 8
  assert int(d.status.max()) & 1  # MJB_STATUS_NEFC_OVERFLOW
  assert mjw.status(_m, d) & 1
  adr_dev = H.dev_np(d, "contact_efc_address")
  assert (adr_dev < 8).all()
  assert not torch.isnan(d.qacc).any()


def test_nconmax_overflow_sets_status(humanoid):
  qpos, qvel, ctrl = H.make_states(humanoid, 8, seed=14)
  qpos[:, 2] -= 0.25
  m, d = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 1, 64)
  mjw.forward(m, d)
  # like the reference's atomic counter (collision_core.py:160-213) nacon counts attempts and may exceed the pool
  assert int(d.nacon[0]) > d.naconmax
  assert int(d.wcon_n.sum()) <= d.naconmax
  assert int(d.status.max()) & 6  # NCON / WCON overflow bits
  assert not torch.isnan(d.qacc).any()


@pytest.mark.parametrize("nworld", [1, 5, 33])
def test_ragged_world_counts(humanoid, nworld):
  qpos, qvel, ctrl = H.make_states(humanoid, nworld, seed=15)
  _, od, _, d = run_both(humanoid, qpos, qvel, ctrl, 24, 64)
  assert_forward_close(od, d)


def test_njmax_zero_means_unconstrained(humanoid):
  qpos, qvel, ctrl = H.make_states(humanoid, 4, seed=16)
  m, d = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 0, 0)
  mjw.forward(m, d)
  assert torch.equal(d.qacc, d.qacc_smooth)
  assert int(d.solver_niter.max()) == 0


def test_rk4_matches_host_composition_of_oracle_forwards(humanoid):
  """forward.py:458 rungekutta4: three extra forward passes; checked against the oracle's forward driven by numpy."""
  from oracle import oracle as O

  lib = O.load()
  qpos, qvel, ctrl = H.make_states(humanoid, 4, seed=17)
  om, od = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  dt = om.opt["timestep"]

  def integrate_pos(q0, v, h):
    out = q0.copy()
    for w in range(q0.shape[0]):
      for j in range(humanoid.njnt):
        jt, qa, da = int(humanoid.jnt_type[j]), int(humanoid.jnt_qposadr[j]), int(humanoid.jnt_dofadr[j])
        if jt == 0:
          out[w, qa : qa + 3] = q0[w, qa : qa + 3] + h * v[w, da : da + 3]
          quat, om3 = q0[w, qa + 3 : qa + 7].copy(), v[w, da + 3 : da + 6].copy()  # copies: the oracle integrates in place
          lib.mjo_quat_integrate(quat.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), om3.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                                 ctypes.c_double(h))  # fmt: skip
          out[w, qa + 3 : qa + 7] = quat
        else:
          out[w, qa] = q0[w, qa] + h * v[w, da]
    return out

  O.run(om, od, "forward")
  A, B = [0.5, 0.5, 1.0], [1 / 6, 1 / 3, 1 / 3, 1 / 6]
  q0, v0 = od.qpos.copy(), od.qvel.copy()
  vrk, ark = B[0] * od.qvel, B[0] * od.qacc
  for i in range(3):
    od.qpos[:] = integrate_pos(q0, od.qvel, A[i] * dt)
    od.qvel[:] = v0 + A[i] * dt * od.qacc
    O.run(om, od, "forward")
    vrk += B[i + 1] * od.qvel
    ark += B[i + 1] * od.qacc
  v1 = v0 + dt * ark
  q1 = integrate_pos(q0, vrk, dt)

  m, d = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  m.opt.integrator = int(C.IntegratorType.RK4)
  mjw.step(m, d)
  eq, ev = H.max_err(H.dev_np(d, "qpos"), q1), H.max_err(H.dev_np(d, "qvel"), v1)
  assert eq < 1e-5 and ev < 5e-3, (eq, ev)
  assert abs(float(d.time[0]) - dt) < 1e-7


def test_reset_then_step_equals_fresh_data(humanoid):
  qpos, qvel, ctrl = H.make_states(humanoid, 8, seed=18)
  m, d = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  for _ in range(3):
    mjw.step(m, d)
  mjw.reset_data(m, d)
  fresh = mjw.make_data(humanoid, nworld=8, nconmax=24, njmax=64)
  mjw.step(m, d)
  mjw.step(m, fresh)
  assert torch.equal(d.qpos, fresh.qpos) and torch.equal(d.qvel, fresh.qvel)


def test_get_data_into_after_forward(humanoid):
  from oracle import oracle as O

  qpos, qvel, ctrl = H.make_states(humanoid, 4, seed=19)
  om, od = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  O.run(om, od, "forward")
  m, d = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  mjw.forward(m, d)
  out = mjw.make_host_data(humanoid)
  mjw.get_data_into(out, humanoid, d, world_id=2)
  assert out.nefc == int(od.nefc[2]) and out.ncon == int(od.wcon_n[2])
  assert out.efc_J.shape == (out.nefc, humanoid.nv)
  assert H.max_err(out.qacc, od.qacc[2]) < 2e-2
  np.testing.assert_array_equal(out.contact.geom, H.world_contacts_orc(od, 2, "contact_geom"))


def test_full_size_properties(humanoid):
  """nworld = 8192 (BASELINE.json configs[1]): size-independent properties + a 32-world oracle sample."""
  from oracle import oracle as O

  nworld = 8192
  qpos, qvel, ctrl = H.make_states(humanoid, nworld, seed=20)
  m, d = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  for _ in range(3):
    mjw.step(m, d)
  mjw.forward(m, d)
  torch.cuda.synchronize()
  assert int(d.status.max()) == 0
  assert not torch.isnan(d.qpos).any() and not torch.isnan(d.qacc).any()
  nefc = d.nefc.long()
  assert int(nefc.max()) <= 64 and int(d.nacon[0]) == int(d.wcon_n.sum()) <= d.naconmax
  assert torch.equal(nefc, (d.ne + d.nf + d.nl).long() + (nefc - (d.ne + d.nf + d.nl).long()))
  # quaternions stay normalised; M qacc_smooth = qfrc_smooth; contact frames orthonormal
  assert float((d.qpos[:, 3:7].norm(dim=1) - 1).abs().max()) < 1e-5
  M = d.qM[:, :27, :27].double()
  res = torch.einsum("wij,wj->wi", M, d.qacc_smooth.double()) - d.qfrc_smooth.double()
  assert float(res.abs().max() / (1 + d.qfrc_smooth.abs().max())) < 1e-3
  n = int(d.nacon[0])
  F = d.contact.frame[:n].double()
  assert float((F @ F.transpose(1, 2) - torch.eye(3, device="cuda", dtype=torch.float64)).abs().max()) < 1e-5
  # KKT residual of every world
  f = torch.where(torch.arange(64, device="cuda")[None] < nefc[:, None], d.efc.force, torch.zeros_like(d.efc.force)).double()
  J = d.efc.J[:, :, :27].double()
  kkt = torch.einsum("wij,wj->wi", M, d.qacc.double()) - d.qfrc_smooth.double() - torch.einsum("wri,wr->wi", J, f)
  scale = 1.0 / (float(m.stat.meaninertia) * 27)
  # float32 noise floor of the gradient: ~1e-4 worst case, ~2e-6 median (tolerance is 1e-6)
  # (a world may also stop on the improvement criterion with a larger gradient, solver.py:3177: allow a tail)
  r = kkt.norm(dim=1) * scale
  assert float(r.median()) < 1e-4 and float(torch.quantile(r, 0.999)) < 5e-3 and float(r.max()) < 5e-2, (
    float(r.median()), float(torch.quantile(r, 0.999)), float(r.max()))
  # a sample of worlds against the oracle, from the device's own state
  idx = np.arange(0, nworld, nworld // 32)[:32]
  om, od = H.oracle_data(humanoid, H.dev_np(d, "qpos")[idx], H.dev_np(d, "qvel")[idx], H.dev_np(d, "ctrl")[idx], 24, 64,
                         warmstart=H.dev_np(d, "qacc_warmstart")[idx])  # fmt: skip
  O.run(om, od, "forward", nthread=4)
  np.testing.assert_array_equal(H.dev_np(d, "nefc")[idx], od.nefc)
  assert H.max_err(H.dev_np(d, "qacc")[idx], od.qacc) < 2e-2


BF = C.BroadphaseFilter


@pytest.mark.parametrize("broadphase", [C.BroadphaseType.SAP_TILE, C.BroadphaseType.SAP_SEGMENTED])
@pytest.mark.parametrize("bfilter", [BF.PLANE | BF.SPHERE | BF.OBB, BF.PLANE | BF.AABB, BF.PLANE | BF.OBB, 0])
def test_sweep_and_prune_broadphase(humanoid, broadphase, bfilter):
  """reference collision_driver.py:554: the sweep result feeds the same filters as NXN.  Counts, contact lists
  and the solved forces must equal the oracle's sweep; with the sphere filter on they also equal NXN's."""
  qpos, qvel, ctrl = H.make_states(humanoid, 33, seed=40)
  opt = {"broadphase": int(broadphase), "broadphase_filter": int(bfilter)}
  _, od, m, d = run_both(humanoid, qpos, qvel, ctrl, 24, 64, opt=opt)
  assert int(d.status.max()) == 0
  np.testing.assert_array_equal(H.dev_np(d, "ncollision"), od.ncollision)
  assert_forward_close(od, d)
  geom_dev, geom_ref = H.dev_np(d, "contact_geom"), od.contact_geom
  for w in range(33):
    n = int(od.wcon_n[w])
    np.testing.assert_array_equal(geom_dev[H.dev_np(d, "wcon")[w, :n]], geom_ref[od.wcon[w, :n]])
  if bfilter & BF.SPHERE:
    ncol, nefc = H.dev_np(d, "ncollision").copy(), H.dev_np(d, "nefc").copy()
    m.opt.broadphase = int(C.BroadphaseType.NXN)
    mjw.forward(m, d)
    np.testing.assert_array_equal(H.dev_np(d, "ncollision"), ncol)
    np.testing.assert_array_equal(H.dev_np(d, "nefc"), nefc)


BOX_PAIRS_XML = TM.BOX_PAIRS_XML


def test_sphere_box_capsule_box_and_box_box_contacts():
  """collision_primitive_core.py:1103 sphere_box, :1158 capsule_box, :648 box_box on 512 random poses (large margins so
  separated pairs report too).  float32 and float64 may pick different closest features at exact ties, so the discrete outcome
  (contact count per world) may differ on a handful of worlds; everywhere else contacts must agree to 1e-4."""
  from oracle import oracle as O

  mjm = mjw.load_mjcf_string(BOX_PAIRS_XML)
  mjm.opt.disableflags |= int(C.DisableBit.NATIVECCD)  # the box_box primitive is the NATIVECCD-disabled route; margins need it too (io.py:386)
  rng = np.random.default_rng(3)
  nw = 512
  qpos = np.zeros((nw, 35))
  for b in range(5):
    qpos[:, 7 * b : 7 * b + 3] = rng.uniform(-0.4, 0.4, (nw, 3))
    q = rng.normal(size=(nw, 4))
    qpos[:, 7 * b + 3 : 7 * b + 7] = q / np.linalg.norm(q, axis=1, keepdims=True)
  qvel, ctrl = np.zeros((nw, mjm.nv)), np.zeros((nw, 0))
  om, od = H.oracle_data(mjm, qpos, qvel, ctrl, 32, 64)
  O.run(om, od, O.STAGE["kinematics"] | O.STAGE["com_pos"] | O.STAGE["collision"], nthread=4)
  m, d = H.device_data(mjw, mjm, qpos, qvel, ctrl, 32, 64)
  mjw.kinematics(m, d)
  mjw.com_pos(m, d)
  mjw.collision(m, d)
  torch.cuda.synchronize()
  assert int(d.status.max()) == 0
  n_dev, n_ref = H.dev_np(d, "wcon_n"), od.wcon_n
  same = n_dev == n_ref
  assert (~same).sum() <= 8, (~same).sum()
  wcon = H.dev_np(d, "wcon")
  fields = {f: H.dev_np(d, f) for f in ("contact_dist", "contact_pos", "contact_frame", "contact_geom", "contact_geomcollisionid")}
  nbox = nboxbox = worst = 0
  for w in np.nonzero(same)[0]:
    a, b = wcon[w, : n_ref[w]], od.wcon[w, : n_ref[w]]
    np.testing.assert_array_equal(fields["contact_geom"][a], od.contact_geom[b])
    np.testing.assert_array_equal(fields["contact_geomcollisionid"][a], od.contact_geomcollisionid[b])
    nbox += int((od.contact_geom[b] == 1).sum())  # geom 1 is the first box
    nboxbox += int(((od.contact_geom[b][:, 0] == 1) & (od.contact_geom[b][:, 1] == 5)).sum())
    err = max(H.max_err(fields[f][a], getattr(od, f)[b]) for f in ("contact_dist", "contact_pos", "contact_frame"))
    worst = max(worst, err)
  assert nbox > 600 and nboxbox > 300, (nbox, nboxbox)
  assert worst < 1e-4, worst


@pytest.mark.parametrize("solver", [C.SolverType.NEWTON, C.SolverType.CG])
def test_elliptic_cones(humanoid, solver):
  """constraint.py:1940 _contact_elliptic, solver.py:263 / :1886 / :2592: condim rows per contact in contact-frame
  axes, cone states and forces, the cone term of the Newton Hessian."""
  qpos, qvel, ctrl = H.make_states(humanoid, 64, seed=31)
  qpos[:, 2] -= 0.1  # feet into the floor: sticking and sliding contacts
  opt = {"cone": int(C.ConeType.ELLIPTIC), "solver": int(solver), "iterations": 100}
  _, od, m, d = run_both(humanoid, qpos, qvel, ctrl, 24, 64, opt=opt)
  assert int(d.status.max()) == 0
  assert_forward_close(od, d, tol_solver=None)
  # sliding contacts leave the acceleration weakly determined along the slip direction, so single elements of the
  # float32 solution move by up to ~0.2 where |qacc| reaches 3e3; the error is therefore measured per world against
  # the world's largest acceleration (the reference's solver tests compare costs, not elements: SURVEY.md section 4)
  for f in ("qacc", "qfrc_constraint"):
    a, b = H.dev_np(d, f), getattr(od, f)
    assert (np.abs(a - b).max(axis=1) / (1.0 + np.abs(b).max(axis=1))).max() < 5e-3, f
  np.testing.assert_array_equal(H.dev_np(d, "efc_type"), od.efc_type)
  for f in H.EFC_FIELDS:
    assert H.max_err(H.dev_np(d, f), getattr(od, f)) < 2e-3, f
  state_dev, state_ref = H.dev_np(d, "efc_state"), od.efc_state
  rows = np.arange(state_ref.shape[1])[None, :] < od.nefc[:, None]
  assert (state_ref[rows] == 4).sum() > 20  # cones in the middle zone are exercised
  assert (state_dev[rows] != state_ref[rows]).mean() < 0.02  # zone boundaries are float-sensitive
  fscale = 1.0 + np.abs(od.efc_force).max(axis=1, keepdims=True)
  assert (np.abs(H.dev_np(d, "efc_force") - od.efc_force) / fscale)[rows].max() < 2e-2
  # and a few steps stay together
  _, od2, _, d2 = run_both(humanoid, qpos, qvel, ctrl, 24, 64, stage="step", opt=opt, nstep=5)
  newton = solver == C.SolverType.NEWTON  # CG stops on its (looser) improvement criterion: float32 / float64 runs part earlier
  assert H.max_err(H.dev_np(d2, "qpos"), od2.qpos) < (1e-4 if newton else 1e-3)
  assert H.max_err(H.dev_np(d2, "qvel"), od2.qvel) < (2e-2 if newton else 1e-1)


def test_rk4_is_native_and_graph_capturable(humanoid):
  """mjb_step with integrator RK4: forward + mjb_rungekutta4 (csrc/mjb_rk4.cu) -- 4 x 5 stage-group launches + 4 RK launches, no
  torch kernels, no host synchronisation, and a CUDA-graph replay equals the eager step bit for bit."""
  from mujoco_warp_b200 import _lib

  lib = _lib.load()
  qpos, qvel, ctrl = H.make_states(humanoid, 16, seed=23)
  m, a = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  _, b = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  m.opt.integrator = int(C.IntegratorType.RK4)
  mjw.prepare(m, a)  # the RK scratch is allocated here, not inside a capture
  mjw.prepare(m, b)
  n0 = lib.mjb_launch_count()
  with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    mjw.step(m, a)
    torch.cuda.synchronize()
  assert lib.mjb_launch_count() - n0 == 4 * 5 + 4  # forward = five stage groups (no integrator group)
  names = [e.key for e in prof.key_averages() if e.device_type == torch.autograd.DeviceType.CUDA]
  foreign = [n for n in names if "mjb_" not in n and "memset" not in n.lower()]
  assert not foreign, foreign
  mjw.step(m, a)
  s = torch.cuda.Stream()
  with torch.cuda.stream(s):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
      mjw.step(m, b)
    g.replay()
    g.replay()
  torch.cuda.synchronize()
  for f in ["qpos", "qvel", "qacc", "qacc_warmstart", "time"]:
    assert torch.equal(a.field(f), b.field(f)), f


def test_rk4_public_function_after_forward_equals_step(humanoid):
  """forward(m, d); rungekutta4(m, d) == step(m, d) with integrator RK4 (forward.py:1013-1016)."""
  qpos, qvel, ctrl = H.make_states(humanoid, 8, seed=29)
  m, a = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  _, b = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  m.opt.integrator = int(C.IntegratorType.RK4)
  mjw.step(m, a)
  mjw.forward(m, b)
  mjw.rungekutta4(m, b)
  for f in ["qpos", "qvel", "qacc", "time"]:
    assert torch.equal(a.field(f), b.field(f)), f


def test_sub_stage_functions_reproduce_fwd_velocity(humanoid):
  """com_vel / passive / rne / energy_pos / energy_vel are public stage functions of the reference too (mujoco_warp/__init__.py): run on
  their own, each writes exactly the outputs it defines with the values `forward` gives them."""
  qpos, qvel, ctrl = H.make_states(humanoid, 8, seed=31)
  m, a = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  _, b = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  m.opt.enableflags = int(m.opt.enableflags) | int(C.EnableBit.ENERGY)
  mjw.forward(m, a)
  mjw.fwd_position(m, b)
  outs = {"com_vel": ["cvel", "cdof_dot"], "passive": ["qfrc_spring", "qfrc_damper", "qfrc_passive"], "rne": ["cacc", "cfrc_int", "qfrc_bias"]}
  for fn in ("rne", "passive", "com_vel"):  # any order: each recomputes what it needs from qpos / qvel and the position stage
    for f in outs[fn]:
      b.field(f).fill_(7.0)
    getattr(mjw, fn)(m, b)
    for f in outs[fn]:
      assert torch.equal(a.field(f), b.field(f)), (fn, f)
  assert torch.all(b.cfrc_ext == 0)
  mjw.energy_pos(m, b)
  mjw.energy_vel(m, b)
  torch.testing.assert_close(a.energy, b.energy, rtol=1e-6, atol=1e-6)
  assert float(b.energy[:, 1].min()) > 0


def test_xfrc_accumulate_matches_the_oracle(humanoid):
  """support.xfrc_accumulate (support.py:229): J^T xfrc_applied.  The oracle folds it into qfrc_smooth (fwd_acceleration); the applied
  part is what is left after the other terms are taken out."""
  from oracle import oracle as O

  rng = np.random.default_rng(5)
  qpos, qvel, ctrl = H.make_states(humanoid, 6, seed=17)
  xfrc = rng.uniform(-5, 5, (6, humanoid.nbody, 6))
  xfrc[:, ::3] = 0.0
  om, od = H.oracle_data(humanoid, qpos, qvel, ctrl, 24, 64)
  od.xfrc_applied[:] = xfrc
  O.run(om, od, "forward", nthread=4)
  want = np.asarray(od.qfrc_smooth) - (np.asarray(od.qfrc_passive) - np.asarray(od.qfrc_bias) + np.asarray(od.qfrc_actuator) + np.asarray(od.qfrc_applied))
  m, d = H.device_data(mjw, humanoid, qpos, qvel, ctrl, 24, 64)
  d.xfrc_applied.copy_(torch.from_numpy(xfrc.astype(np.float32)))
  mjw.kinematics(m, d)
  mjw.com_pos(m, d)
  base = torch.full((6, humanoid.nv), 0.5, device="cuda")
  qfrc = base.clone()
  mjw.xfrc_accumulate(m, d, qfrc)
  torch.cuda.synchronize()
  assert H.max_err((qfrc - base).cpu().numpy(), want) < 2e-5
  with pytest.raises(ValueError):
    mjw.xfrc_accumulate(m, d, torch.zeros((6, humanoid.nv + 1), device="cuda"))


def test_deriv_smooth_vel_is_the_velocity_derivative_of_the_smooth_forces():
  """derivative.deriv_smooth_vel (derivative.py:321): out = qM - dt * d(qfrc_actuator + damping force) / d qvel.  Both forces are
  affine in qvel, so a central difference of the device's own fwd_velocity / fwd_actuation is exact up to rounding; the flags
  that drop a term (DisableBit.ACTUATION, DisableBit.DAMPER) are checked the same way."""
  mjm = mjw.load_mjcf_string(TM.XML["rk4_actuators"])
  nw = 3
  rng = np.random.default_rng(2)
  qpos, qvel, ctrl = H.make_states(mjm, nw, seed=4)
  for flags in (0, int(DB.ACTUATION), int(DB.DAMPER)):
    m, d = H.device_data(mjw, mjm, qpos, qvel, ctrl, 16, 32)
    m.opt.disableflags = flags
    d.act.copy_(torch.from_numpy(rng.uniform(-0.2, 0.2, (nw, mjm.na)).astype(np.float32)))
    mjw.forward(m, d)
    nv, dt = mjm.nv, float(mjm.opt.timestep)
    out = torch.full((nw, nv + 1, nv + 1), 7.0, device="cuda")
    mjw.deriv_smooth_vel(m, d, out)
    qM = d.qM[:, :nv, :nv].double().cpu().numpy()
    got = out.double().cpu().numpy()
    assert np.all(got[:, nv, :] == 7.0) and np.all(got[:, :, nv] == 7.0)  # only the leading nv x nv block is written

    def forces(v):
      d.qvel.copy_(torch.from_numpy(v.astype(np.float32)))
      mjw.fwd_velocity(m, d)
      mjw.fwd_actuation(m, d)
      torch.cuda.synchronize()
      return (d.qfrc_actuator.double() + d.qfrc_damper.double()).cpu().numpy()

    eps = 0.25
    deriv = np.zeros((nw, nv, nv))
    for j in range(nv):
      dv = np.zeros((nw, nv))
      dv[:, j] = eps
      deriv[:, :, j] = (forces(qvel + dv) - forces(qvel - dv)) / (2 * eps)
    want = qM - dt * deriv
    assert H.max_err(got[:, :nv, :nv], want) < 2e-4, flags


@pytest.mark.parametrize("cone", [C.ConeType.PYRAMIDAL, C.ConeType.ELLIPTIC])
def test_contact_rows_for_every_condim(cone):
  """condim 1, 3, 4 and 6 contacts side by side (1, 4, 6, 10 pyramid rows / 1, 3, 4, 6 elliptic rows; constraint.py:1669, :1940), nine
  active contacts per world so that make_constraint's four-contacts-per-pass loop runs full, partial and repeated passes."""
  geoms = []
  for i, cd in enumerate((1, 3, 4, 6, 3, 6, 1, 4, 3)):
    shape = 'type="sphere" size=".1"' if i % 2 == 0 else 'type="box" size=".1 .08 .06"'
    geoms.append(f'<body pos="{0.4 * i} 0 0.09"><freejoint/><geom {shape} condim="{cd}" friction="0.8 0.02 0.003"/></body>')
  xml = f'<mujoco><option timestep="0.002"/><worldbody><geom type="plane" size="5 5 .1" condim="1"/>{"".join(geoms)}</worldbody></mujoco>'
  mjm = mjw.load_mjcf_string(xml)
  nw = 6
  rng = np.random.default_rng(8)
  qpos = np.repeat(np.asarray(mjm.qpos0)[None], nw, axis=0)
  for j in range(mjm.njnt):
    qa = int(mjm.jnt_qposadr[j])
    qpos[:, qa + 2] -= rng.uniform(0.0, 0.03, nw)
    q = qpos[:, qa + 3 : qa + 7] + rng.normal(0, 0.1, (nw, 4))
    qpos[:, qa + 3 : qa + 7] = q / np.linalg.norm(q, axis=1, keepdims=True)
  qvel = rng.uniform(-0.5, 0.5, (nw, mjm.nv))
  ctrl = np.zeros((nw, mjm.nu))
  opt = {"cone": int(cone)}
  _, od, m, d = run_both(mjm, qpos, qvel, ctrl, 48, 160, opt=opt)
  assert int(d.status.max()) == 0
  assert int(od.nacon[0]) >= 9 * nw
  for f in ["ne", "nf", "nl", "nefc", "wcon_n", "nacon"]:
    np.testing.assert_array_equal(H.dev_np(d, f), getattr(od, f), err_msg=f)
  np.testing.assert_array_equal(H.dev_np(d, "efc_type"), od.efc_type)
  for f in H.EFC_FIELDS:
    assert H.max_err(H.dev_np(d, f), getattr(od, f)) < 2e-3, f
