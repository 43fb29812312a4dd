"""Shared helpers of the parity tests: seeded states, oracle / device runners, comparisons."""

import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HUMANOID = os.path.join(ROOT, "benchmarks", "humanoid", "humanoid.xml")

# fields compared after `forward`, grouped by the tolerance class of the reference's tests
# (smooth/forward/constraint 5e-4, solver outputs 0.1 / cost check)
SMOOTH_FIELDS = (
  "xpos xquat xmat xipos ximat xanchor xaxis geom_xpos geom_xmat subtree_com cinert cdof crb qM "
  "actuator_length actuator_moment actuator_velocity cvel cdof_dot qfrc_spring qfrc_damper qfrc_passive "
  "qfrc_bias actuator_force qfrc_actuator qfrc_smooth qacc_smooth qLD cacc cfrc_int"
).split()
EFC_FIELDS = "efc_J efc_pos efc_margin efc_D efc_vel efc_aref efc_frictionloss".split()
CONTACT_FIELDS = "contact_dist contact_pos contact_frame contact_includemargin contact_friction contact_solref contact_solreffriction contact_solimp".split()
CONTACT_INT_FIELDS = "contact_dim contact_geom contact_type contact_geomcollisionid".split()


def make_states(mjm, nworld, seed=0, noise=1.0):
  """Seeded generic states: keyframes (cycled) plus bounded noise on hinge coordinates, root height,
  root orientation, velocities and controls (SURVEY.md 8d randomised-state variant)."""
  rng = np.random.default_rng(seed)
  nkey = max(int(mjm.nkey), 1)
  qpos = np.zeros((nworld, mjm.nq))
  for w in range(nworld):
    base = np.asarray(mjm.key_qpos[w % nkey]) if mjm.nkey else np.asarray(mjm.qpos0)
    q = base.copy()
    for j in range(mjm.njnt):
      jt, qa = int(mjm.jnt_type[j]), int(mjm.jnt_qposadr[j])
      if jt == 0:
        q[qa + 2] += noise * rng.uniform(-0.03, 0.1)
        quat = q[qa + 3 : qa + 7] + noise * rng.normal(0, 0.05, 4)
        q[qa + 3 : qa + 7] = quat / np.linalg.norm(quat)
      elif jt in (2, 3):
        q[qa] += noise * rng.uniform(-0.05, 0.05)
    qpos[w] = q
  qvel = noise * rng.uniform(-0.1, 0.1, (nworld, mjm.nv))
  ctrl = noise * rng.uniform(-1, 1, (nworld, mjm.nu))
  return qpos, qvel, ctrl


def oracle_data(mjm, qpos, qvel, ctrl, nconmax, njmax, warmstart=None):
  from oracle import oracle as O

  om = O.OracleModel(mjm)
  d = O.make_data(om, nworld=qpos.shape[0], nconmax=nconmax, njmax=njmax)
  d.qpos[:], d.qvel[:], d.ctrl[:] = qpos, qvel, ctrl
  if warmstart is not None:
    d.qacc_warmstart[:] = warmstart
  return om, d


def device_data(mjw, mjm, qpos, qvel, ctrl, nconmax, njmax, warmstart=None):
  import torch

  m = mjw.put_model(mjm)
  d = mjw.make_data(mjm, nworld=qpos.shape[0], nconmax=nconmax, njmax=njmax)
  f = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()  # noqa: E731
  d.qpos.copy_(f(qpos))
  d.qvel.copy_(f(qvel))
  d.ctrl.copy_(f(ctrl))
  if warmstart is not None:
    d.qacc_warmstart.copy_(f(warmstart))
  return m, d


def dev_np(d, name):
  return d.field(name).detach().cpu().numpy()


def max_err(a, b):
  """max over elements of |a-b| / (1 + |b|): combined abs/rel error."""
  a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
  if a.size == 0:
    return 0.0
  return float(np.max(np.abs(a - b) / (1.0 + np.abs(b))))


def world_contacts_dev(d, w, name):
  n = int(d.wcon_n[w])
  ids = d.wcon[w, :n].long()
  return d.field(name)[ids].detach().cpu().numpy()


def world_contacts_orc(d, w, name):
  n = int(d.wcon_n[w])
  ids = d.wcon[w, :n]
  return getattr(d, name)[ids]
