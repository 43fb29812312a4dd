"""Cases of the reference-executed golden vectors (`reference_step.npz`): models, seeded states, option overrides, and
the comparison of a Data-like object against the stored reference outputs.

Shared by the generator (`make_reference_golden.py`, which needs /root/reference) and by the tests (which only read the
committed .npz).  Nothing here imports the Warp interpreter or touches /root/reference.
"""

from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)

import mujoco_warp_b200 as mjw  # noqa: E402
from tests import helpers as H  # noqa: E402
from tests import models as TM  # noqa: E402

NPZ = os.path.join(HERE, "reference_step.npz")

REAL_FIELDS = (
  "xpos xquat xmat xipos ximat xanchor xaxis geom_xpos geom_xmat subtree_com cinert cdof crb qM qLD actuator_length actuator_moment "
  "actuator_velocity cvel cdof_dot qfrc_spring qfrc_damper qfrc_gravcomp qfrc_passive qfrc_bias actuator_force qfrc_actuator qfrc_smooth qacc_smooth "
  "cacc cfrc_int qfrc_constraint qacc qacc_warmstart qpos qvel act act_dot time energy"
).split()
EFC_REAL = "J pos margin D vel aref frictionloss force Ma".split()
EFC_INT = "type id state".split()
CON_REAL = "dist pos frame includemargin friction solref solreffriction solimp".split()
CON_INT = "dim geom efc_address worldid type geomcollisionid".split()
COUNTS = "ne nf nl nefc nacon ncollision solver_niter".split()


def _states(mjm, nworld, seed, sink=0.0, key_first=True):
  qpos, qvel, ctrl = H.make_states(mjm, nworld, seed=seed)
  if key_first and mjm.nkey:
    qpos[0] = np.asarray(mjm.key_qpos[0])
    qvel[0] = 0.0
  if sink:
    qpos[1:, 2] -= sink
  return qpos, qvel, ctrl


def _multi_root_states(mjm, nworld, seed, sink=0.0):
  """make_states on a model without keyframes (several floating trees); every free root is lowered by `sink` in all worlds but the first."""
  qpos, qvel, ctrl = H.make_states(mjm, nworld, seed=seed)
  for j in range(mjm.njnt):
    if int(mjm.jnt_type[j]) == 0:
      qpos[1:, int(mjm.jnt_qposadr[j]) + 2] -= sink
  return qpos, qvel, ctrl


def _small_states(mjm, nworld, seed):
  rng = np.random.default_rng(seed)
  qpos = np.repeat(np.asarray(mjm.qpos0, dtype=np.float64)[None], nworld, axis=0)
  for j in range(mjm.njnt):
    jt, qa = int(mjm.jnt_type[j]), int(mjm.jnt_qposadr[j])
    if jt in (2, 3):
      qpos[:, qa] += rng.uniform(-0.3, 0.3, nworld)
    elif jt == 0:
      qpos[:, qa + 2] += rng.uniform(-0.02, 0.02, nworld)
    elif jt == 1:
      q = qpos[:, qa : qa + 4] + rng.normal(0, 0.2, (nworld, 4))
      qpos[:, qa : qa + 4] = q / np.linalg.norm(q, axis=1, keepdims=True)
  return qpos, rng.uniform(-0.5, 0.5, (nworld, mjm.nv)), rng.uniform(-2, 2, (nworld, mjm.nu))


def _box_pair_states(mjm, nworld, seed):
  rng = np.random.default_rng(seed)
  qpos = np.zeros((nworld, mjm.nq))
  for b in range(mjm.nq // 7):
    qpos[:, 7 * b : 7 * b + 3] = rng.uniform(-0.4, 0.4, (nworld, 3))
    q = rng.normal(size=(nworld, 4))
    qpos[:, 7 * b + 3 : 7 * b + 7] = q / np.linalg.norm(q, axis=1, keepdims=True)
  return qpos, np.zeros((nworld, mjm.nv)), np.zeros((nworld, mjm.nu))


def _box_stack_states(mjm, nworld, seed):
  """Box pairs in face-face, edge-face and corner contact with small penetrations: boxes laid on top of / beside box 0 with yaw
  (flat faces: 4 clipped contacts), tilted about an edge (2 contacts) or standing on a corner (1 contact), plus a random world."""
  rng = np.random.default_rng(seed)
  qpos = np.zeros((nworld, mjm.nq))
  sizes = np.asarray(mjm.geom_size)[1:]  # the plane is geom 0

  def quat(axis, ang):
    axis = np.asarray(axis, dtype=float) / np.linalg.norm(axis)
    return np.concatenate(([np.cos(ang / 2)], np.sin(ang / 2) * axis))

  def qmul(a, b):
    return np.array([a[0]*b[0]-a[1]*b[1]-a[2]*b[2]-a[3]*b[3], a[0]*b[1]+a[1]*b[0]+a[2]*b[3]-a[3]*b[2],
                     a[0]*b[2]-a[1]*b[3]+a[2]*b[0]+a[3]*b[1], a[0]*b[3]+a[1]*b[2]-a[2]*b[1]+a[3]*b[0]])  # fmt: skip

  for w in range(nworld):
    far = np.array([[0, 0, 0], [3, 0, 0], [0, 3, 0], [3, 3, 0], [6, 3, 0]], dtype=float)
    q = np.tile([1.0, 0, 0, 0], (5, 1))
    pos = far.copy()
    q[0] = quat([0.3, -0.2, 1.0], 0.2 * w)  # the base box, slightly rotated
    kind = w % 4
    pen = 0.004 + 0.003 * rng.uniform()
    for b in (1, 2):
      s = sizes[b]
      yaw = quat([0, 0, 1], rng.uniform(0.2, 1.2))
      off = np.array([rng.uniform(-0.1, 0.1), rng.uniform(-0.1, 0.1), 0.0])
      if kind == 0 or (kind == 3 and b == 1):  # flat on top of box 0 (b = 1) / flat below it (b = 2)
        sign = 1.0 if b == 1 else -1.0
        local_q, h = yaw, s[2]
      elif kind == 1:  # tilted about an edge: the lower edge touches the top face
        tilt = quat([1, 0, 0], 0.5)
        local_q, h = qmul(yaw, tilt), abs(np.cos(0.5)) * s[2] + abs(np.sin(0.5)) * s[1]
        sign = 1.0 if b == 1 else -1.0
      else:  # standing on a corner
        tilt = qmul(quat([1, 0, 0], 0.6), quat([0, 1, 0], 0.5))
        local_q = qmul(yaw, tilt)
        Rl = _quat_to_mat(local_q)
        h = float(np.abs(Rl[2] * s).sum())
        sign = 1.0 if b == 1 else -1.0
      if sign < 0:
        local_q = qmul(quat([1, 0, 0], np.pi), local_q)
      local_p = off + np.array([0, 0, sign * (sizes[0][2] + h - pen)])
      R0m = _quat_to_mat(q[0])
      pos[b] = R0m @ local_p
      q[b] = qmul(q[0], local_q)
    if kind == 3:  # boxes 3 and 4 in a random overlapping pose
      pos[3] = [3, 3, 0]
      pos[4] = pos[3] + rng.uniform(-0.12, 0.12, 3)
      for b in (3, 4):
        v = rng.normal(size=4)
        q[b] = v / np.linalg.norm(v)
    for b in range(5):
      qpos[w, 7 * b : 7 * b + 3] = pos[b]
      qpos[w, 7 * b + 3 : 7 * b + 7] = q[b] / np.linalg.norm(q[b])
  return qpos, np.zeros((nworld, mjm.nv)), np.zeros((nworld, mjm.nu))


def _hfield_states(mjm, nworld, seed):
  """Every free body dropped at a random spot of the terrain, a little into its surface (bilinear height at the spot), random
  orientation; the bodies are kept apart from each other in x / y so that only terrain contacts arise."""
  rng = np.random.default_rng(seed)
  nb = mjm.nq // 7
  size = np.asarray(mjm.hfield_size[0])
  nrow, ncol = int(mjm.hfield_nrow[0]), int(mjm.hfield_ncol[0])
  data = np.asarray(mjm.hfield_data).reshape(nrow, ncol)
  gpos, gmat = np.asarray(mjm.geom_pos[0]), _quat_to_mat(np.asarray(mjm.geom_quat[0]))
  qpos = np.zeros((nworld, mjm.nq))
  for w in range(nworld):
    cells = rng.permutation(6)[:nb]  # a 3 x 2 partition of the field, one body per cell
    for b in range(nb):
      cx, cy = cells[b] % 3, cells[b] // 3
      lx = -size[0] + (cx + rng.uniform(0.3, 0.7)) * 2 * size[0] / 3
      ly = -size[1] + (cy + rng.uniform(0.3, 0.7)) * 2 * size[1] / 2
      fx, fy = (lx + size[0]) / (2 * size[0]) * (ncol - 1), (ly + size[1]) / (2 * size[1]) * (nrow - 1)
      c0, r0 = min(int(fx), ncol - 2), min(int(fy), nrow - 2)
      tx, ty = fx - c0, fy - r0
      h = ((1 - tx) * (1 - ty) * data[r0, c0] + tx * (1 - ty) * data[r0, c0 + 1] + (1 - tx) * ty * data[r0 + 1, c0] + tx * ty * data[r0 + 1, c0 + 1]) * size[2]
      p = gmat @ np.array([lx, ly, h + rng.uniform(0.02, 0.09)]) + gpos
      q = rng.normal(size=4)
      qpos[w, 7 * b : 7 * b + 3] = p
      qpos[w, 7 * b + 3 : 7 * b + 7] = q / np.linalg.norm(q)
  return qpos, np.zeros((nworld, mjm.nv)), np.zeros((nworld, mjm.nu))


def _quat_to_mat(q):
  w, x, y, z = q / np.linalg.norm(q)
  return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                   [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                   [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])  # fmt: skip


def _aloha_states(mjm, nworld, seed):
  """Keyframe lift_pot0 / lift_pot4 (arms at the pot; the keyframes hold an all-zero pot quaternion, which MuJoCo's
  mj_kinematics turns into the identity for the reference's put_data; here a slightly tilted one) plus small joint noise."""
  rng = np.random.default_rng(seed)
  keys = [i for i, n in enumerate(mjm.key_names) if n.startswith("lift_pot")]
  qpos = np.stack([np.asarray(mjm.key_qpos[keys[(4 * w) % len(keys)]], dtype=np.float64) for w in range(nworld)])
  ctrl = np.stack([np.asarray(mjm.key_ctrl[keys[(4 * w) % len(keys)]], dtype=np.float64) for w in range(nworld)])
  for j in range(mjm.njnt):
    jt, qa = int(mjm.jnt_type[j]), int(mjm.jnt_qposadr[j])
    if jt == 0:
      quat = np.array([1.0, 0.013, -0.021, 0.005]) + rng.normal(0, 0.004, (nworld, 4))  # slightly tilted: a pot standing exactly flat on
      qpos[:, qa + 3 : qa + 7] = quat / np.linalg.norm(quat, axis=1, keepdims=True)  # the table has no unique witness point (EPA tie)
      qpos[:, qa + 2] -= rng.uniform(0.0, 0.012, nworld)  # the pot a little into the table: box-mesh contacts
    elif jt == 3:
      qpos[:, qa] += rng.uniform(-0.02, 0.02, nworld)
  return qpos, rng.uniform(-0.05, 0.05, (nworld, mjm.nv)), ctrl


def _aloha_replay_states(mjm, nworld, seed):
  """Keyframe lift_pot0 exactly as testspeed --replay starts (qvel 0; the keyframe's all-zero pot quaternion becomes the identity,
  as mj_kinematics does for the reference's put_data) then tilted by a fraction of a degree: a pot standing exactly flat on the
  table has no unique EPA witness point)."""
  rng = np.random.default_rng(seed)
  k0 = [i for i, n in enumerate(mjm.key_names) if n.startswith("lift_pot")][0]
  qpos = np.repeat(np.asarray(mjm.key_qpos[k0], dtype=np.float64)[None], nworld, axis=0)
  ctrl = np.repeat(np.asarray(mjm.key_ctrl[k0], dtype=np.float64)[None], nworld, axis=0)
  for j in range(mjm.njnt):
    if int(mjm.jnt_type[j]) == 0:
      qa = int(mjm.jnt_qposadr[j])
      quat = np.array([1.0, 0.0, 0.0, 0.0]) + rng.normal(0, 0.004, (nworld, 4))
      qpos[:, qa + 3 : qa + 7] = quat / np.linalg.norm(quat, axis=1, keepdims=True)
  # the keyframe holds the lid joint EXACTLY on its lower limit (q = 0, range 0 .. 0.06) with zero load: whether `q - lo < 0` makes
  # a limit row is then decided by rounding noise (1e-24), in the reference as much as here.  Start it a millimetre open instead.
  lid = [j for j in range(mjm.njnt) if int(mjm.jnt_type[j]) in (2, 3) and tuple(np.asarray(mjm.jnt_range[j])) == (0.0, 0.06)]
  for j in lid:
    qpos[:, int(mjm.jnt_qposadr[j])] = 0.001 + 0.0005 * np.arange(nworld)
  return qpos, np.zeros((nworld, mjm.nv)), ctrl


def _panda_states(mjm, nworld, seed):
  """Joint angles a little beyond the joint ranges (active limits), velocities, controls beyond ctrlrange (clamping)."""
  rng = np.random.default_rng(seed)
  lo, hi = np.asarray(mjm.jnt_range)[:, 0], np.asarray(mjm.jnt_range)[:, 1]
  qpos = lo - 0.03 * (hi - lo) + rng.uniform(0, 1, (nworld, mjm.nq)) * 1.06 * (hi - lo)
  clo, chi = np.asarray(mjm.actuator_ctrlrange)[:, 0], np.asarray(mjm.actuator_ctrlrange)[:, 1]
  ctrl = clo - 0.1 * (chi - clo) + rng.uniform(0, 1, (nworld, mjm.nu)) * 1.2 * (chi - clo)
  return qpos, rng.uniform(-0.5, 0.5, (nworld, mjm.nv)), ctrl


# case name -> (model key, state fn, nworld, seed, opt overrides, mode, nstep, nconmax, njmax)
CASES = {
  "humanoid_forward": dict(model="humanoid", nworld=4, seed=3, sink=0.05, mode="forward"),
  "humanoid_step5": dict(model="humanoid", nworld=3, seed=11, sink=0.03, mode="step", nstep=5),
  "humanoid_cg": dict(model="humanoid", nworld=2, seed=5, sink=0.05, mode="forward", opt=dict(solver=1)),
  # CG pinned iteration by iteration: 6 iterations on states where no linesearch bracket test is decided by rounding noise
  # (when a Newton step of the 1-D search lands exactly on the minimiser, the sign of a ~1e-10 derivative picks the branch at
  # solver.py:1294-1316, and two float64 evaluation orders then follow different -- equally valid -- iterates)
  "humanoid_cg_iter6": dict(model="humanoid", nworld=3, seed=21, sink=0.04, mode="forward", opt=dict(solver=1, iterations=6)),
  "humanoid_elliptic_newton": dict(model="humanoid", nworld=3, seed=31, sink=0.1, mode="forward", opt=dict(cone=1)),
  "humanoid_elliptic_cg": dict(model="humanoid", nworld=2, seed=32, sink=0.1, mode="forward", opt=dict(cone=1, solver=1)),
  "humanoid_implicitfast": dict(model="humanoid", nworld=2, seed=7, sink=0.03, mode="step", nstep=3, opt=dict(integrator=3)),
  "humanoid_rk4": dict(model="humanoid", nworld=2, seed=8, sink=0.03, mode="step", nstep=2, opt=dict(integrator=1)),
  # RK4 with stateful actuators (filterexact + actrange clamp, integrator, filter, actearly): next_act in the stage perturbation
  # (unclamped) and the final advance (clamped), forward.py:355-400, :213-274
  "rk4_actuators_step4": dict(model="rk4_actuators", nworld=4, seed=17, mode="step", nstep=4, nconmax=16, njmax=32, opt=dict(integrator=1)),
  "ball_limits_step3": dict(model="ball_limits", nworld=6, seed=23, mode="step", nstep=3, nconmax=8, njmax=16),
  "humanoid_sap_tile": dict(model="humanoid", nworld=2, seed=9, sink=0.05, mode="forward", opt=dict(broadphase=1)),
  "humanoid_sap_segmented": dict(model="humanoid", nworld=2, seed=9, sink=0.05, mode="forward", opt=dict(broadphase=2)),
  "humanoid_no_eulerdamp_refsafe": dict(model="humanoid", nworld=2, seed=12, sink=0.03, mode="step", nstep=2, disable=["EULERDAMP", "REFSAFE"]),
  "humanoid_no_contact_limit": dict(model="humanoid", nworld=2, seed=13, mode="step", nstep=2, disable=["CONTACT", "LIMIT"]),
  "humanoid_no_gravity_actuation": dict(model="humanoid", nworld=2, seed=14, sink=0.03, mode="forward", disable=["GRAVITY", "ACTUATION", "WARMSTART"]),
  "equality_friction_limits": dict(model="equality_friction_limits", nworld=4, seed=5, mode="step", nstep=3, nconmax=32),
  "equality_friction_limits_implicitfast": dict(model="equality_friction_limits", nworld=2, seed=6, mode="step", nstep=3, nconmax=32, opt=dict(integrator=3)),
  "box_and_spheres_on_plane": dict(model="box_and_spheres_on_plane", nworld=4, seed=5, mode="step", nstep=3, nconmax=32),
  "no_constraints_no_actuators": dict(model="no_constraints_no_actuators", nworld=2, seed=5, mode="step", nstep=3, nconmax=32),
  # box-box goes through the primitive only with NATIVECCD disabled (collision_driver.py:736); margins need it too (io.py:389)
  "box_pairs_collision": dict(model="box_pairs", nworld=24, seed=3, mode="forward", nconmax=32, njmax=160, states="box_pairs", disable=["NATIVECCD"]),
  # benchmarks/franka_emika_panda (BASELINE config 3): implicitfast, joint equality, position / general actuators with force
  # and ctrl clamps, active joint limits, plane-box and plane-capsule contacts; mesh geoms are visual-only (contype 0)
  "panda_step3": dict(model="panda", nworld=10, seed=50, mode="step", nstep=3, nconmax=8, njmax=40, states="panda"),
  "panda_deep_forward": dict(model="panda", nworld=10, seed=43, mode="forward", nconmax=8, njmax=80, states="panda"),
  # EnableBit.ENERGY: potential (gravity + joint springs, free / ball / hinge / slide) and kinetic energy (sensor.py:2854, :2922)
  "humanoid_energy_step2": dict(model="humanoid", nworld=3, seed=71, mode="step", nstep=2, enable=["ENERGY"]),
  "springs_energy_forward": dict(model="springs", nworld=4, seed=72, mode="forward", enable=["ENERGY"]),
  "gravcomp_step3": dict(model="gravcomp", nworld=4, seed=81, mode="step", nstep=3),
  "site_equalities_step3": dict(model="site_equalities", nworld=4, seed=91, mode="step", nstep=3, njmax=32),
  # BASELINE config 5 (benchmarks/aloha_pot: 133 convex-hull meshes up to 1159 vertices, 9154 candidate geom pairs, elliptic cones)
  "aloha_pot_forward": dict(model="aloha", nworld=2, seed=95, mode="forward", nconmax=24, njmax=128, states="aloha"),
  # config 5 beyond one forward: three Newton / elliptic steps, the CG solver, and the first 20 steps of the benchmark's own
  # `lift_pot` keyframe replay (testspeed.py --replay; io.py:2603 make_trajectory) with the control trajectory applied per step
  "aloha_pot_step3": dict(model="aloha", nworld=2, seed=96, mode="step", nstep=3, nconmax=24, njmax=128, states="aloha"),
  "aloha_pot_cg_forward": dict(model="aloha", nworld=2, seed=97, mode="forward", nconmax=24, njmax=128, states="aloha", opt=dict(solver=1)),
  "aloha_lift_pot_replay20": dict(model="aloha", nworld=2, seed=98, mode="step", nstep=20, nconmax=24, njmax=128, states="aloha_replay", replay="lift_pot"),
  "connect_weld_forward": dict(model="connect_weld", nworld=5, seed=61, mode="forward", njmax=48),
  "connect_weld_step4": dict(model="connect_weld", nworld=4, seed=62, mode="step", nstep=4, njmax=48),
  "connect_weld_njmax_overflow": dict(model="connect_weld", nworld=2, seed=63, mode="forward", njmax=16),
  # BASELINE config 4 (benchmarks/apptronik_apollo/scene_flat.xml: nv 25, frictionloss on every joint, position actuators with force
  # ranges, capsule / box feet on the plane, eulerdamp disabled, timestep 5 ms).  Its box geoms carry a margin, which the reference
  # only accepts with NATIVECCD disabled (io.py:386-395): box-box then runs the box_box primitive.
  "apollo_forward": dict(model="apollo", nworld=4, seed=31, sink=0.04, mode="forward", nconmax=16, njmax=96, states="humanoid", disable=["NATIVECCD"]),
  "apollo_step3": dict(model="apollo", nworld=3, seed=32, sink=0.03, mode="step", nstep=3, nconmax=16, njmax=64, states="humanoid", disable=["NATIVECCD"]),
  # box-box on the reference's DEFAULT route: GJK / EPA + multi-contact face clipping (flat faces 4 contacts, edge 2, corner 1)
  "box_stack_multicontact": dict(model="box_stack", nworld=12, seed=6, mode="forward", nconmax=32, njmax=160, states="box_stack"),
  # height field against every convex geom type (collision_convex.py:56 _hfield_filter, :158 ccd_hfield_kernel)
  "hfield_collision": dict(model="hfield", nworld=8, seed=14, mode="forward", nconmax=32, njmax=128, states="hfield"),
  # nv = 81: the reference stores qM / qLD / efc.J sparse and factorises the Hessian in 16 x 16 blocks (solver.py:2758, block_cholesky.py);
  # the generator stores qM and efc.J densified (make_reference_golden.py), qLD (sparse L^T D L) is not comparable with a Cholesky factor
  "three_humanoids_forward": dict(model="three_humanoids", nworld=2, seed=41, sink=0.06, mode="forward", nconmax=48, njmax=192, states="multi_root"),
  "three_humanoids_step2": dict(model="three_humanoids", nworld=2, seed=42, sink=0.04, mode="step", nstep=2, nconmax=48, njmax=192, states="multi_root"),
  "convex_pairs_collision": dict(model="convex_pairs", nworld=12, seed=4, mode="forward", nconmax=48, njmax=200, states="box_pairs"),
  "mesh_pairs_collision": dict(model="mesh_pairs", nworld=12, seed=9, mode="forward", nconmax=48, njmax=200, states="box_pairs"),
}


def load_model(key):
  if key == "humanoid":
    return mjw.load_mjcf(H.HUMANOID)
  if key == "three_humanoids":
    return mjw.load_mjcf(os.path.join(ROOT, "benchmarks", "humanoid", "three_humanoids.xml"))
  if key == "aloha":
    return mjw.load_mjcf(os.path.join(ROOT, "benchmarks", "aloha_pot", "scene.xml"))
  if key == "apollo":
    return mjw.load_mjcf(os.path.join(ROOT, "benchmarks", "apptronik_apollo", "scene_flat.xml"))
  if key == "panda":
    return mjw.load_mjcf(os.path.join(ROOT, "benchmarks", "franka_emika_panda", "scene.xml"))
  return mjw.load_mjcf_string(TM.XML[key])


def load_case_model(name):
  """The case's host model with its disable flags set on the model itself: put_model validates against them (io.py:373)."""
  from mujoco_warp_b200 import constants as C

  c = CASES[name]
  mjm = load_model(c["model"])
  for flag in c.get("disable", []):
    mjm.opt.disableflags |= int(getattr(C.DisableBit, flag))
  return mjm


def case_inputs(name):
  c = CASES[name]
  mjm = load_case_model(name)
  kind = c.get("states", "humanoid" if c["model"] == "humanoid" else "small")
  if kind == "humanoid":
    qpos, qvel, ctrl = _states(mjm, c["nworld"], c["seed"], c.get("sink", 0.0))
  elif kind == "multi_root":
    qpos, qvel, ctrl = _multi_root_states(mjm, c["nworld"], c["seed"], c.get("sink", 0.0))
  elif kind == "panda":
    qpos, qvel, ctrl = _panda_states(mjm, c["nworld"], c["seed"])
  elif kind == "aloha":
    qpos, qvel, ctrl = _aloha_states(mjm, c["nworld"], c["seed"])
  elif kind == "aloha_replay":
    qpos, qvel, ctrl = _aloha_replay_states(mjm, c["nworld"], c["seed"])
  elif kind == "box_pairs":
    qpos, qvel, ctrl = _box_pair_states(mjm, c["nworld"], c["seed"])
  elif kind == "hfield":
    qpos, qvel, ctrl = _hfield_states(mjm, c["nworld"], c["seed"])
  elif kind == "box_stack":
    qpos, qvel, ctrl = _box_stack_states(mjm, c["nworld"], c["seed"])
  else:
    qpos, qvel, ctrl = _small_states(mjm, c["nworld"], c["seed"])
  return mjm, qpos, qvel, ctrl


def case_ctrl_traj(name, mjm):
  """(nstep, nu) control trajectory of a keyframe-replay case (reference io.py:2591 find_keys + :2603 make_trajectory), else None."""
  c = CASES[name]
  if "replay" not in c:
    return None
  return mjw.make_trajectory(mjm, mjw.find_keys(mjm, c["replay"]))[: c.get("nstep", 1)]


def _shape_like(r, o):
  r = np.asarray(r, dtype=np.float64)
  o = np.asarray(o, dtype=np.float64)
  if r.shape == o.shape:
    return r, o
  if r.size == o.size:
    return r.reshape(o.shape), o
  if r.ndim == o.ndim:
    sl = tuple(slice(0, min(a, b)) for a, b in zip(r.shape, o.shape))
    return r[sl], o[sl]
  raise ValueError((r.shape, o.shape))


def run_oracle(name):
  from oracle import oracle as O

  c = CASES[name]
  mjm, qpos, qvel, ctrl = case_inputs(name)
  om, od = H.oracle_data(mjm, qpos, qvel, ctrl, c.get("nconmax", 24), c.get("njmax", 64))
  opt = dict(c.get("opt", {}))
  flags = int(om.opt["disableflags"])
  from mujoco_warp_b200 import constants as C

  for flag in c.get("disable", []):
    flags |= int(getattr(C.DisableBit, flag))
  opt["disableflags"] = flags
  eflags = int(om.opt["enableflags"])
  for flag in c.get("enable", []):
    eflags |= int(getattr(C.EnableBit, flag))
  opt["enableflags"] = eflags
  traj = case_ctrl_traj(name, mjm)
  if traj is None:
    H.oracle_steps(om, od, c["mode"], c.get("nstep", 1), opt)
  else:
    for t in range(c.get("nstep", 1)):
      od.ctrl[:] = traj[t]
      H.oracle_steps(om, od, c["mode"], 1, opt)
  return od


def compare(name, ref_out, od, verbose=True):
  """Worst |a-b|/(1+|b|) per field group between the reference output and the oracle's Data."""
  worst = {}
  for f in COUNTS:
    if f == "solver_niter" or not hasattr(od, f):
      continue
    a, b = np.asarray(ref_out[f]).reshape(-1), np.asarray(getattr(od, f)).reshape(-1)
    if not np.array_equal(a, b):
      worst["count:" + f] = float("inf")
      if verbose:
        print("   count mismatch", f, a, b)
  for f in REAL_FIELDS:
    if not hasattr(od, f) or f not in ref_out:  # fields added to the list after a case was recorded
      continue
    r, o = _shape_like(ref_out[f], getattr(od, f))
    if f == "cfrc_int":  # the reference never clears the world body's entry (smooth.py:1199 skips body 0, :1233 adds into it)
      r, o = r[:, 1:], o[:, 1:]
    worst[f] = H.max_err(o, r)
  rows = H.canonical_efc(ref_out, prefix="efc_")
  orows = H.canonical_efc(od, prefix="efc_")
  for f in EFC_REAL + EFC_INT:
    r, o = rows.get(f), orows.get(f)
    if r is None or o is None:
      continue
    worst["efc_" + f] = H.max_err(o, r) if r.shape == o.shape else float("inf")
  rc, oc = H.canonical_contacts(ref_out), H.canonical_contacts(od)
  for f in CON_REAL + ["dim", "geom"]:
    r, o = rc.get(f), oc.get(f)
    if r is None or o is None:
      continue
    worst["contact_" + f] = H.max_err(o, r) if r.shape == o.shape else float("inf")
  return worst




# ----------------------------------------------------------------------------------------------
# used by the tests
# ----------------------------------------------------------------------------------------------
SOLVER_FIELDS = "qfrc_constraint qacc qacc_warmstart efc_force efc_state efc_Ma qpos qvel act act_dot".split()
# cases whose solver does not run to the quadratic-convergence regime: the iterates depend on rounding-level ties in
# the linesearch (see the note at CASES["humanoid_cg_iter6"]), only the converged solution is comparable, to solver tolerance
LOOSE_SOLVER_CASES = ("humanoid_cg", "humanoid_elliptic_cg", "aloha_pot_cg_forward")
# GJK / EPA stop on opt.ccd_tolerance (1e-6): two float64 evaluation orders may differ by one EPA iteration (witness points
# move by < the tolerance; tests/test_oracle_ccd_golden.py quantifies it pair by pair)
CASE_TOL = {"convex_pairs_collision": 5e-6, "mesh_pairs_collision": 5e-6}


def load_golden(name):
  """{field: array} of one case from the committed reference_step.npz."""
  with np.load(NPZ) as z:
    pre = name + "/"
    return {k[len(pre):]: z[k] for k in z.files if k.startswith(pre)}


def case_opt(name):
  """Option overrides of a case as {Option field: value}, disable flags merged into `disableflags`."""
  from mujoco_warp_b200 import constants as C

  c = CASES[name]
  mjm = load_model(c["model"])
  opt = dict(c.get("opt", {}))
  flags = int(mjm.opt.disableflags)
  for flag in c.get("disable", []):
    flags |= int(getattr(C.DisableBit, flag))
  opt["disableflags"] = flags
  eflags = int(mjm.opt.enableflags)
  for flag in c.get("enable", []):
    eflags |= int(getattr(C.EnableBit, flag))
  opt["enableflags"] = eflags
  return opt


class DeviceView:
  """Presents a device Data to `compare` as attributes holding numpy arrays."""

  def __init__(self, d):
    self._d = d

  def __getattr__(self, name):
    try:
      return H.dev_np(self._d, name)
    except Exception as e:  # unknown field
      raise AttributeError(name) from e
