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


# ----------------------------------------------------------------------------------------------
# canonical orderings: the reference appends contacts / efc rows through atomic counters, so only the
# SET of rows per world is defined; these helpers sort both sides the same way before comparing
# ----------------------------------------------------------------------------------------------
def _get(obj, name):
  return np.asarray(obj[name] if isinstance(obj, dict) else getattr(obj, name))


def canonical_contacts(obj):
  """Contacts [0, nacon) sorted by (worldid, geom0, geom1, order of appearance); 'rank' maps conid -> sorted position."""
  n = int(_get(obj, "nacon").reshape(-1)[0])
  geom = _get(obj, "contact_geom")[:n].reshape(n, 2)
  wid = _get(obj, "contact_worldid")[:n].reshape(n)
  order = np.lexsort((np.arange(n), geom[:, 1], geom[:, 0], wid))
  out = {"rank": np.empty(n, dtype=np.int64), "order": order}
  out["rank"][order] = np.arange(n)
  for f in "dist pos frame includemargin friction solref solreffriction solimp dim geom worldid".split():
    a = _get(obj, "contact_" + f)[:n]
    out[f] = a.reshape(n, -1)[order] if n else a.reshape(0, int(np.prod(a.shape[1:])) if a.ndim > 1 else 1)
  return out


def canonical_efc(obj, prefix="efc_"):
  """Active rows of every world, sorted per world by (type, canonical id), concatenated over worlds."""
  nefc = _get(obj, "nefc").reshape(-1)
  types, ids = _get(obj, prefix + "type"), _get(obj, prefix + "id")
  rank = canonical_contacts(obj)["rank"]
  fields = {}
  for f in "J pos margin D vel aref frictionloss force state type".split():
    try:
      fields[f] = _get(obj, prefix + f)
    except (KeyError, AttributeError):
      pass
  out = {f: [] for f in fields}
  out["id"] = []
  for w, n in enumerate(nefc):
    n = min(int(n), types.shape[1])
    t = types[w, :n].astype(np.int64)
    i = ids[w, :n].astype(np.int64)
    i = np.where(t >= 5, rank[np.clip(i, 0, max(len(rank) - 1, 0))] if len(rank) else i, i)
    order = np.lexsort((np.arange(n), i, t))
    for f, a in fields.items():
      out[f].append(np.asarray(a[w][:n])[order])
    out["id"].append(i[order])
  return {f: (np.concatenate(v) if v else np.zeros(0)) for f, v in out.items()}


def oracle_rk4_step(om, od, mjm, opt=None):
  """reference forward.py:458 rungekutta4 composed on the host from the oracle's forward (the C oracle has no RK4 driver)."""
  import ctypes

  from oracle import oracle as O

  lib = O.load()
  dt = (opt or {}).get("timestep", om.opt["timestep"])

  def integrate_pos(q0, v, h):
    out = q0.copy()
    for w in range(q0.shape[0]):
      for j in range(mjm.njnt):
        jt, qa, da = int(mjm.jnt_type[j]), int(mjm.jnt_qposadr[j]), int(mjm.jnt_dofadr[j])
        if jt in (0, 1):
          if jt == 0:
            out[w, qa : qa + 3] = q0[w, qa : qa + 3] + h * v[w, da : da + 3]
            qa, da = qa + 3, da + 3
          quat, om3 = q0[w, qa : qa + 4].copy(), v[w, da : da + 3].copy()
          lib.mjo_quat_integrate(quat.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), om3.ctypes.data_as(ctypes.POINTER(ctypes.c_double)),
                                 ctypes.c_double(h))  # fmt: skip
          out[w, qa : qa + 4] = quat
        else:
          out[w, qa] = q0[w, qa] + h * v[w, da]
    return out

  def next_act(act0, act_dot, scale, clamp):  # support.py:38-64 over every actuator with activations
    out = act0.copy()
    for u in range(mjm.nu):
      a0, n = int(mjm.actuator_actadr[u]), int(mjm.actuator_actnum[u])
      if a0 < 0:
        continue
      dyn = int(mjm.actuator_dyntype[u])
      for j in range(a0, a0 + n):
        if dyn == 3:
          tau = max(1e-15, float(mjm.actuator_dynprm[u][0]))
          out[:, j] = act0[:, j] + scale * act_dot[:, j] * tau * (1.0 - np.exp(-dt / tau))
        elif dyn == 5:
          continue
        else:
          out[:, j] = act0[:, j] + scale * act_dot[:, j] * dt
        if clamp and bool(mjm.actuator_actlimited[u]):
          out[:, j] = np.clip(out[:, j], mjm.actuator_actrange[u][0], mjm.actuator_actrange[u][1])
    return out

  O.run(om, od, "forward", opt=opt)
  A, B = [0.5, 0.5, 1.0], [1 / 6, 1 / 3, 1 / 3, 1 / 6]
  q0, v0, t0, a0 = od.qpos.copy(), od.qvel.copy(), od.time.copy(), od.act.copy()
  vrk, ark, drk = B[0] * od.qvel, B[0] * od.qacc, B[0] * od.act_dot
  for i in range(3):
    od.qpos[:] = integrate_pos(q0, od.qvel, A[i] * dt)
    od.qvel[:] = v0 + A[i] * dt * od.qacc
    if mjm.na:
      od.act[:] = next_act(a0, od.act_dot, A[i], False)
    O.run(om, od, "forward", opt=opt)
    vrk += B[i + 1] * od.qvel
    ark += B[i + 1] * od.qacc
    drk += B[i + 1] * od.act_dot
  if mjm.na:
    od.act_dot[:] = drk
    od.act[:] = next_act(a0, drk, 1.0, True)
  od.qvel[:] = v0 + dt * ark
  od.qpos[:] = integrate_pos(q0, vrk, dt)
  od.time[:] = t0 + dt
  od.qacc_warmstart[:] = od.qacc  # forward.py:274 copies d.qacc, which holds the LAST stage's acceleration
  return od


def oracle_steps(om, od, mode, nstep, opt=None, nthread=1):
  """Runs `forward` or `nstep` steps of the oracle under the given option overrides (integrator-aware)."""
  from oracle import oracle as O

  integ = int((opt or {}).get("integrator", om.opt["integrator"]))
  for _ in range(nstep):
    if mode == "forward":
      O.run(om, od, "forward", opt=opt, nthread=nthread)
    elif integ == 1:
      oracle_rk4_step(om, od, om.mjm, opt)
    else:
      O.run(om, od, "step_implicitfast" if integ == 3 else "step", opt=opt, nthread=nthread)
  return od


# ---- support golden (tests/golden/support_golden.npz: outputs of the reference's get_state / set_state / contact_force / reset_data)
SUPPORT_NPZ = os.path.join(ROOT, "tests", "golden", "support_golden.npz")
SUPPORT_CASES = {"humanoid_pyramidal": "humanoid_step5", "humanoid_elliptic": "humanoid_elliptic_newton", "equality": "equality_friction_limits",
                 "box_spheres_pyramidal": "box_and_spheres_on_plane", "box_spheres_elliptic": "box_and_spheres_on_plane"}  # fmt: skip
SUPPORT_STATE_FIELDS = "time qpos qvel act qacc_warmstart ctrl qfrc_applied xfrc_applied eq_active mocap_pos mocap_quat".split()
SUPPORT_RESET_FIELDS = SUPPORT_STATE_FIELDS + "qacc act_dot energy qM solver_niter ne nf nl nefc nacon".split()
SUPPORT_CONTACT_FIELDS = ["contact_" + f for f in "dist pos frame includemargin friction solref solreffriction solimp dim geom efc_address worldid type geomcollisionid".split()]


def load_support_case(name):
  with np.load(SUPPORT_NPZ) as z:
    pre = name + "/"
    return {k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}


def support_fill(rec, prefix, fields, assign):
  """Calls assign(field, array) for every `prefix + field` entry of a support golden record."""
  for f in fields:
    if prefix + f in rec:
      assign(f, rec[prefix + f])


def match_contact_sets(ref, dev, nworld, atol_pos=1e-6, atol_dist=1e-7, atol_nrm=1e-6, strict=True):
  """Compares two contact buffers as per-world, per-pair SETS of (dist, pos, normal): duplicates collapse and order is ignored.
  Used where the reference's selection of a pair's contacts has rounding-level ties (height fields: two prisms of a cell report the
  same contact, and which duplicate is kept is decided in the last bit).  `ref` / `dev`: dicts with nacon, contact_worldid, contact_geom,
  contact_dist, contact_pos, contact_frame.  The touching pairs must be the same and the DEEPEST contact of every pair must match;
  strict: every contact of either side must have a partner on the other.  Returns (matched reference contacts, reference contacts)."""
  def unpack(o):
    n = int(np.asarray(o["nacon"]).reshape(-1)[0])
    return (np.asarray(o["contact_worldid"])[:n].reshape(-1), np.asarray(o["contact_geom"])[:n].reshape(n, 2), np.asarray(o["contact_dist"])[:n].reshape(-1),
            np.asarray(o["contact_pos"])[:n].reshape(n, 3), np.asarray(o["contact_frame"])[:n].reshape(n, 9)[:, :3])  # fmt: skip

  def err_to(b_d, b_p, b_n, d, p, n):
    return np.maximum(np.abs(b_d - d) / atol_dist, np.maximum(np.abs(b_p - p).max(axis=1) / atol_pos, np.abs(b_n - n).max(axis=1) / atol_nrm))

  rw, rg, rd, rp, rn = unpack(ref)
  dw, dg, dd, dp, dn = unpack(dev)
  matched = total = 0
  for w in range(nworld):
    pairs_r = {tuple(x) for x in rg[rw == w].tolist()}
    pairs_d = {tuple(x) for x in dg[dw == w].tolist()}
    assert pairs_r == pairs_d, (w, sorted(pairs_r ^ pairs_d))
    for pair in pairs_r:
      kr = (rw == w) & (rg == pair).all(axis=1)
      kd = (dw == w) & (dg == pair).all(axis=1)
      i0 = int(np.argmin(rd[kr]))  # the deepest contact is the minimum over all prisms: no selection involved
      e0 = err_to(dd[kd], dp[kd], dn[kd], rd[kr][i0], rp[kr][i0], rn[kr][i0]).min()
      assert e0 <= 1.0, (w, pair, "deepest contact", float(rd[kr][i0]), float(dd[kd].min()), float(e0))
      for i in range(int(kr.sum())):
        ok = err_to(dd[kd], dp[kd], dn[kd], rd[kr][i], rp[kr][i], rn[kr][i]).min() <= 1.0
        assert ok or not strict, (w, pair, "ref->dev", float(rd[kr][i]), rp[kr][i].tolist())
        matched += int(ok)
        total += 1
      if strict:
        for i in range(int(kd.sum())):
          assert err_to(rd[kr], rp[kr], rn[kr], dd[kd][i], dp[kd][i], dn[kd][i]).min() <= 1.0, (w, pair, "dev->ref", float(dd[kd][i]), dp[kd][i].tolist())
  return matched, total
