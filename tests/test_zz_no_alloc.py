"""The allocation-free RK4 entry point (`mjb_rungekutta4_static`: scratch in a file-scope `static __device__` array, no cudaMalloc)
against the allocating one (`mjb_rungekutta4`), through the C ABI, on two calls in a row.

Collected last on purpose (a failure here must not hide the other GPU tests).  Passes alone in a fresh
process and after the rest of the suite: it touches no state shared with the other tests except the library's static scratch,
which every call overwrites before reading.
"""

import ctypes

import numpy as np
import pytest
import torch

import mujoco_warp_b200 as mjw
from mujoco_warp_b200 import _lib
from mujoco_warp_b200 import constants as C
from tests import helpers as H

pytestmark = pytest.mark.gpu

FIELDS = ["qpos", "qvel", "act", "act_dot", "qacc", "qacc_warmstart", "time"]


def _forward_then(lib, entry, m, d):
  """mjb_forward + `entry` (an RK4 integrator entry point) on the current stream, straight through the C ABI."""
  opt = _lib.make_option(m)
  s = _lib.stream_ptr()
  _lib.check(lib, lib.mjb_forward(m._handle, d._handle, ctypes.byref(opt), s))
  _lib.check(lib, entry(m._handle, d._handle, ctypes.byref(opt), s))


@pytest.mark.parametrize("model", ["humanoid", "rk4_actuators"])
def test_static_scratch_rk4_matches_allocating_rk4_on_two_calls_in_a_row(model):
  from tests import models as TM

  lib = _lib.load()
  if model == "humanoid":
    mjm = mjw.load_mjcf(H.HUMANOID)
    qpos, qvel, ctrl = H.make_states(mjm, 64, seed=41)
  else:  # stateful actuators: the act / act_dot part of the scratch
    from tests.golden import reference_cases as RC

    mjm = mjw.load_mjcf_string(TM.XML[model])
    qpos, qvel, ctrl = RC._small_states(mjm, 64, 43)
  m, a = H.device_data(mjw, mjm, qpos, qvel, ctrl, 24, 64)
  _, b = H.device_data(mjw, mjm, qpos, qvel, ctrl, 24, 64)
  m.opt.integrator = int(C.IntegratorType.RK4)
  from mujoco_warp_b200.forward import _ensure_bound

  _ensure_bound(m, a)
  _ensure_bound(m, b)
  for call in range(2):
    _forward_then(lib, lib.mjb_rungekutta4, m, a)
    _forward_then(lib, lib.mjb_rungekutta4_static, m, b)
    torch.cuda.synchronize()
    for f in FIELDS:
      # same kernels, same launch geometry, same arithmetic order: bit-identical; stated as the tolerance the RK4 tests use (exact)
      assert torch.equal(a.field(f), b.field(f)), (call, f, H.max_err(H.dev_np(a, f), H.dev_np(b, f)))
  assert not torch.isnan(b.qpos).any() and float(b.time[0]) == pytest.approx(2 * float(m.opt.timestep[0]), rel=1e-6)
  assert not np.array_equal(H.dev_np(b, "qpos"), qpos.astype(np.float32))  # the state did advance


def test_static_scratch_rk4_rejects_sizes_beyond_its_maximum():
  lib = _lib.load()
  mjm = mjw.load_mjcf(H.HUMANOID)
  m = mjw.put_model(mjm)
  stride = (mjm.nq + 3 * mjm.nv + 2 * mjm.na + 3) // 4 * 4
  nworld = int(lib.mjb_rk_static_max_floats()) // stride + 1
  d = mjw.make_data(mjm, nworld=nworld, nconmax=1, njmax=8)
  from mujoco_warp_b200.forward import _ensure_bound

  _ensure_bound(m, d)
  opt = _lib.make_option(m)
  rc = lib.mjb_rungekutta4_static(m._handle, d._handle, ctypes.byref(opt), _lib.stream_ptr())
  assert rc == -6
  msg = lib.mjb_last_error().decode()
  assert "exceeds the static scratch" in msg and "MJB_RK_STATIC_MAX_FLOATS" in msg
  # one world fewer fits
  d2 = mjw.make_data(mjm, nworld=nworld - 1, nconmax=1, njmax=8)
  _ensure_bound(m, d2)
  _forward_then(lib, lib.mjb_rungekutta4_static, m, d2)
  torch.cuda.synchronize()
  assert not torch.isnan(d2.qpos).any()
