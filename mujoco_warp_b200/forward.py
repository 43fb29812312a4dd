"""Pipeline driver: step / forward / step1 / step2 and every per-stage function f(m, d).

Same names and call signatures as the reference (reference mujoco_warp/_src/forward.py:973
forward, :1004 step, :1022 step1, :1050 step2, :514 fwd_position, :593 fwd_velocity,
:837 fwd_actuation, :950 fwd_acceleration, :327 euler, :458 rungekutta4; smooth.py:358
kinematics, :602 com_pos, :889 crb, :1104 factor_m, :2848 solve_m, :2605 transmission;
collision_driver.py:755 collision; constraint.py:2210 make_constraint; solver.py:3297 solve;
support.py:133 mul_m).  Each call goes through the C ABI (mjb_run_stages with the stage bits of the
requested function: one kernel launch per stage group touched, six for a whole step); nothing is
computed on the host and there is no fallback path.
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
STAGE_COM_VEL = 1 << 15
STAGE_PASSIVE = 1 << 16
STAGE_RNE = 1 << 17
STAGE_ENERGY_POS = 1 << 18
STAGE_ENERGY_VEL = 1 << 19
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


def prepare(m: types.Model, d: types.Data):
  """Resolves launch configurations and library-owned scratch for the current `m.opt` ahead of time (mjb_prepare), so that the next
  call can be captured into a CUDA graph.  `_ensure_bound` calls it on every (re)bind; call it again after changing
  `m.opt.cone / broadphase / integrator` if the next call is to be captured."""
  lib = _lib.load()
  _ensure_bound(m, d)
  opt = _lib.make_option(m)
  _lib.check(lib, lib.mjb_prepare(m._handle, d._handle, ctypes.byref(opt)))


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


def com_vel(m, d):
  """Body velocities in the com frame and the time derivatives of the motion axes (reference smooth.py:2015)."""
  _run(m, d, STAGE_COM_VEL)


def passive(m, d):
  """Passive forces: joint springs, dampers, gravity compensation (reference passive.py:729)."""
  _run(m, d, STAGE_PASSIVE)


def rne(m, d):
  """Recursive Newton-Euler bias forces: cacc, cfrc_int, qfrc_bias (reference smooth.py:1259)."""
  _run(m, d, STAGE_RNE)


def energy_pos(m, d):
  """Potential energy into d.energy[:, 0] (reference sensor.py:2854)."""
  _run(m, d, STAGE_ENERGY_POS)


def energy_vel(m, d):
  """Kinetic energy into d.energy[:, 1] (reference sensor.py:2922)."""
  _run(m, d, STAGE_ENERGY_VEL)


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
  """RK4 integrator (reference forward.py:458), called after `forward` like the reference's.

  Native: four small save / accumulate / perturb / advance launches around three `forward` passes (csrc/mjb_rk4.cu); activations go
  through the device `next_act` (filterexact dynamics, actrange clamp on the final advance)."""
  lib = _lib.load()
  _ensure_bound(m, d)
  opt = _lib.make_option(m)
  _lib.check(lib, lib.mjb_rungekutta4(m._handle, d._handle, ctypes.byref(opt), _lib.stream_ptr()))


def step(m, d):
  """Advances the simulation by one timestep (reference forward.py:1004)."""
  integrator = int(m.opt.integrator)
  if integrator not in (C.IntegratorType.EULER, C.IntegratorType.RK4, C.IntegratorType.IMPLICITFAST):
    raise NotImplementedError(f"integrator {integrator} not implemented.")
  lib = _lib.load()
  _ensure_bound(m, d)
  opt = _lib.make_option(m)
  _lib.check(lib, lib.mjb_step(m._handle, d._handle, ctypes.byref(opt), _lib.stream_ptr()))


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


def xfrc_accumulate(m, d, qfrc: torch.Tensor):
  """qfrc += J^T d.xfrc_applied (reference support.py:229)."""
  lib = _lib.load()
  _ensure_bound(m, d)
  if tuple(qfrc.shape) != (d.nworld, m.nv) or qfrc.dtype != torch.float32 or not qfrc.is_contiguous() or not qfrc.is_cuda:
    raise ValueError(f"qfrc must be a contiguous float32 CUDA tensor of shape {(d.nworld, m.nv)}")
  _lib.check(lib, lib.mjb_xfrc_accumulate(m._handle, d._handle, ctypes.c_void_p(qfrc.data_ptr()), _lib.stream_ptr()))


def deriv_smooth_vel(m, d, out: torch.Tensor):
  """out = qM - dt * qDeriv, the matrix implicitfast factorises (reference derivative.py:321); out: (nworld, n, n) with n >= nv."""
  lib = _lib.load()
  _ensure_bound(m, d)
  if out.dim() != 3 or out.shape[0] != d.nworld or out.shape[1] != out.shape[2] or out.shape[1] < m.nv or out.dtype != torch.float32 \
      or not out.is_contiguous() or not out.is_cuda:
    raise ValueError(f"out must be a contiguous float32 CUDA tensor of shape (nworld, n, n) with n >= nv = {m.nv}")
  opt = _lib.make_option(m)
  _lib.check(lib, lib.mjb_deriv_smooth_vel(m._handle, d._handle, ctypes.byref(opt), ctypes.c_void_p(out.data_ptr()), int(out.shape[1]), _lib.stream_ptr()))


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
