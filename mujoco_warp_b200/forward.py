"""Pipeline driver: step / forward / step1 / step2 and every per-stage function f(m, d).

Same names and call signatures as the reference (reference mujoco_warp/_src/forward.py:973
forward, :1004 step, :1022 step1, :1050 step2, :514 fwd_position, :593 fwd_velocity,
:837 fwd_actuation, :950 fwd_acceleration, :327 euler, :458 rungekutta4; smooth.py:358
kinematics, :602 com_pos, :889 crb, :1104 factor_m, :2848 solve_m, :2605 transmission;
collision_driver.py:755 collision; constraint.py:2210 make_constraint; solver.py:3297 solve;
support.py:133 mul_m).  Each call is ONE kernel launch through the C ABI (mjb_run_stages with
the stage bits of the requested function); nothing is computed on the host and there is no
fallback path.
"""

from __future__ import annotations

import ctypes

import torch

from . import _lib
from . import constants as C
from . import io as _io
from . import types

STAGE_KINEMATICS = 1 << 0
STAGE_COM_POS = 1 << 1
STAGE_CRB = 1 << 2
STAGE_FACTOR_M = 1 << 3
STAGE_COLLISION = 1 << 4
STAGE_MAKE_CONSTRAINT = 1 << 5
STAGE_TRANSMISSION = 1 << 6
STAGE_FWD_VELOCITY = 1 << 7
STAGE_FWD_ACTUATION = 1 << 8
STAGE_FWD_ACCELERATION = 1 << 9
STAGE_FWD_ACC_FACTOR = 1 << 10
STAGE_SOLVE = 1 << 11
STAGE_EULER = 1 << 12
STAGE_IMPLICIT = 1 << 13
STAGE_ZERO_SENSOR = 1 << 14
STAGES_FWD_POSITION = STAGE_KINEMATICS | STAGE_COM_POS | STAGE_CRB | STAGE_COLLISION | STAGE_MAKE_CONSTRAINT | STAGE_TRANSMISSION
STAGES_FORWARD = STAGES_FWD_POSITION | STAGE_ZERO_SENSOR | STAGE_FWD_VELOCITY | STAGE_FWD_ACTUATION | STAGE_FWD_ACC_FACTOR | STAGE_SOLVE


def _ensure_bound(m: types.Model, d: types.Data):
  """Binds (or re-binds, if the user replaced a tensor) the data pointers to the C library."""
  if d._handle is not None and d._model is m and d._bound_epoch == types._EPOCH[0]:
    return  # fast path: no tensor attribute of any Data was assigned since the bind
  if d._handle is None or d._model is not m or d._bound_ptrs != _lib.data_ptr_signature(d):
    _io.bind(m, d)
  else:
    d._bound_epoch = types._EPOCH[0]


def _run(m: types.Model, d: types.Data, stages: int):
  lib = _lib.load()
  _ensure_bound(m, d)
  opt = _lib.make_option(m)
  _lib.check(lib, lib.mjb_run_stages(m._handle, d._handle, ctypes.byref(opt), ctypes.c_uint32(stages), _lib.stream_ptr()))


def kinematics(m, d):
  """Forward kinematics (reference smooth.py:358)."""
  _run(m, d, STAGE_KINEMATICS)


def com_pos(m, d):
  """Subtree CoM, com-frame inertias and motion axes (reference smooth.py:602)."""
  _run(m, d, STAGE_COM_POS)


def crb(m, d):
  """Composite rigid body inertias and the joint-space inertia matrix (reference smooth.py:889)."""
  _run(m, d, STAGE_CRB)


def factor_m(m, d):
  """Cholesky factorisation of qM into qLD (reference smooth.py:1104)."""
  _run(m, d, STAGE_FACTOR_M)


def collision(m, d):
  """Broadphase + narrowphase collision detection (reference collision_driver.py:755)."""
  _run(m, d, STAGE_COLLISION)


def make_constraint(m, d):
  """Constraint rows and dense Jacobian (reference constraint.py:2210)."""
  _run(m, d, STAGE_MAKE_CONSTRAINT)


def transmission(m, d):
  """Actuator lengths and moments (reference smooth.py:2605)."""
  _run(m, d, STAGE_TRANSMISSION)


def fwd_position(m, d, factorize: bool = True):
  """Position-dependent computations (reference forward.py:514)."""
  _run(m, d, STAGES_FWD_POSITION | (STAGE_FACTOR_M if factorize else 0))


def fwd_velocity(m, d):
  """Velocity-dependent computations (reference forward.py:593)."""
  _run(m, d, STAGE_FWD_VELOCITY)


def fwd_actuation(m, d):
  """Actuation-dependent computations (reference forward.py:837)."""
  _run(m, d, STAGE_FWD_ACTUATION)


def fwd_acceleration(m, d, factorize: bool = False):
  """Sums non-constraint forces and computes qacc_smooth (reference forward.py:950)."""
  _run(m, d, STAGE_FWD_ACC_FACTOR if factorize else STAGE_FWD_ACCELERATION)


def solve(m, d):
  """Constraint solver (reference solver.py:3297)."""
  _run(m, d, STAGE_SOLVE)


def euler(m, d):
  """Semi-implicit Euler integrator (reference forward.py:327)."""
  _run(m, d, STAGE_EULER)


def implicit(m, d):
  """implicitfast integrator (reference forward.py:495)."""
  _run(m, d, STAGE_IMPLICIT)


def forward(m, d):
  """Forward dynamics (reference forward.py:973)."""
  _run(m, d, STAGES_FORWARD)


def rungekutta4(m, d):
  """RK4 integrator (reference forward.py:458): three extra `forward` calls on perturbed states."""
  A = [0.5, 0.5, 1.0]
  B = [1.0 / 6.0, 1.0 / 3.0, 1.0 / 3.0, 1.0 / 6.0]
  dt = float(m.opt.timestep.reshape(-1)[0])
  qpos_t0, qvel_t0 = d.qpos.clone(), d.qvel.clone()
  qvel_rk, qacc_rk = B[0] * d.qvel, B[0] * d.qacc
  has_act = m.sizes["na"] > 0
  if has_act:
    act_t0, act_dot_rk = d.act.clone(), B[0] * d.act_dot
  for i in range(3):
    a, b = A[i], B[i + 1]
    _rk_integrate_pos(m, d, qpos_t0, d.qvel, a * dt)
    d.qvel.copy_(qvel_t0 + a * dt * d.qacc)
    if has_act:
      d.act.copy_(act_t0 + a * dt * d.act_dot)
    forward(m, d)
    qvel_rk += b * d.qvel
    qacc_rk += b * d.qacc
    if has_act:
      act_dot_rk += b * d.act_dot
  d.qpos.copy_(qpos_t0)
  d.qvel.copy_(qvel_t0)
  if has_act:
    d.act.copy_(act_t0)
    d.act_dot.copy_(act_dot_rk)
    d.act.add_(dt * d.act_dot)
  d.qvel.add_(dt * qacc_rk)
  _rk_integrate_pos(m, d, d.qpos.clone(), qvel_rk, dt)
  d.time.add_(dt)
  d.qacc_warmstart.copy_(d.qacc)


def _rk_integrate_pos(m, d, qpos0, qvel, h):
  """qpos <- integrate(qpos0, qvel, h) with quaternion joints (reference forward.py:52-112), torch ops."""
  hm = m._host
  out = qpos0.clone()
  for j in range(hm.sizes["njnt"]):
    jt = int(hm.ints["jnt_type"][j])
    qa, da = int(hm.ints["jnt_qposadr"][j]), int(hm.ints["jnt_dofadr"][j])
    if jt in (C.JointType.FREE, C.JointType.BALL):
      if jt == C.JointType.FREE:
        out[:, qa : qa + 3] = qpos0[:, qa : qa + 3] + h * qvel[:, da : da + 3]
        qa, da = qa + 3, da + 3
      q = torch.nn.functional.normalize(qpos0[:, qa : qa + 4], dim=1)
      v = qvel[:, da : da + 3]
      n = v.norm(dim=1, keepdim=True)
      ax = torch.where(n > 0, v / n.clamp_min(1e-30), torch.zeros_like(v))
      ang = 0.5 * h * n
      qr = torch.cat([torch.cos(ang), ax * torch.sin(ang)], dim=1)
      w0, x0, y0, z0 = q.unbind(1)
      w1, x1, y1, z1 = qr.unbind(1)
      res = torch.stack(
        [w0 * w1 - x0 * x1 - y0 * y1 - z0 * z1, w0 * x1 + x0 * w1 + y0 * z1 - z0 * y1, w0 * y1 - x0 * z1 + y0 * w1 + z0 * x1, w0 * z1 + x0 * y1 - y0 * x1 + z0 * w1],
        dim=1,
      )
      out[:, qa : qa + 4] = torch.nn.functional.normalize(res, dim=1)
    else:
      out[:, qa] = qpos0[:, qa] + h * qvel[:, da]
  d.qpos.copy_(out)


def step(m, d):
  """Advances the simulation by one timestep (reference forward.py:1004)."""
  integrator = int(m.opt.integrator)
  if integrator in (C.IntegratorType.EULER, C.IntegratorType.IMPLICITFAST):
    lib = _lib.load()
    _ensure_bound(m, d)
    opt = _lib.make_option(m)
    _lib.check(lib, lib.mjb_step(m._handle, d._handle, ctypes.byref(opt), _lib.stream_ptr()))
  elif integrator == C.IntegratorType.RK4:
    forward(m, d)
    rungekutta4(m, d)
  else:
    raise NotImplementedError(f"integrator {integrator} not implemented.")


def step1(m, d):
  """First half of a two-phase step (reference forward.py:1022)."""
  lib = _lib.load()
  _ensure_bound(m, d)
  opt = _lib.make_option(m)
  _lib.check(lib, lib.mjb_step1(m._handle, d._handle, ctypes.byref(opt), _lib.stream_ptr()))


def step2(m, d):
  """Second half of a two-phase step (reference forward.py:1050)."""
  lib = _lib.load()
  _ensure_bound(m, d)
  opt = _lib.make_option(m)
  _lib.check(lib, lib.mjb_step2(m._handle, d._handle, ctypes.byref(opt), _lib.stream_ptr()))


def solve_m(m, d, x: torch.Tensor, y: torch.Tensor):
  """x = M^-1 y using the factor in d.qLD (reference smooth.py:2848)."""
  lib = _lib.load()
  _ensure_bound(m, d)
  assert x.is_contiguous() and y.is_contiguous() and x.dtype == torch.float32 and y.dtype == torch.float32
  _lib.check(lib, lib.mjb_solve_m(m._handle, d._handle, ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(y.data_ptr()), _lib.stream_ptr()))


def mul_m(m, d, res: torch.Tensor, vec: torch.Tensor):
  """res = qM @ vec (reference support.py:133)."""
  lib = _lib.load()
  _ensure_bound(m, d)
  assert res.is_contiguous() and vec.is_contiguous()
  _lib.check(lib, lib.mjb_mul_m(m._handle, d._handle, ctypes.c_void_p(res.data_ptr()), ctypes.c_void_p(vec.data_ptr()), _lib.stream_ptr()))


def ctrl_noise(m, d, step_index: int, ctrl_center=None, ctrlnoisestd: float = 0.01, ctrlnoiserate: float = 0.1):
  """Ornstein-Uhlenbeck control noise of the benchmark loop (reference benchmark.py:42)."""
  lib = _lib.load()
  _ensure_bound(m, d)
  opt = _lib.make_option(m)
  cptr = ctypes.c_void_p(ctrl_center.data_ptr()) if ctrl_center is not None else ctypes.c_void_p(0)
  _lib.check(lib, lib.mjb_ctrl_noise(m._handle, d._handle, ctypes.byref(opt), cptr, int(step_index), float(ctrlnoisestd), float(ctrlnoiserate), _lib.stream_ptr()))


def status(m, d) -> int:
  """ORed per-world status bits (MJB_STATUS_*: nefc / contact-pool overflow, NaN in qacc).

  Replaces the reference's device printf diagnostics (forward.py:194-210).  Synchronises the stream."""
  lib = _lib.load()
  _ensure_bound(m, d)
  flags = ctypes.c_uint32(0)
  _lib.check(lib, lib.mjb_status(d._handle, ctypes.byref(flags), _lib.stream_ptr()))
  return int(flags.value)
