"""State packing and contact forces: the callers either side of `step` in an RL loop.

Same names / argument meaning as the reference (mujoco_warp/_src/support.py:572 get_state, :711 set_state, :354
contact_force).  Each call is one kernel launch through the C ABI (mjb_get_state / mjb_set_state / mjb_contact_force):
asynchronous on the current stream, CUDA-graph capturable, no CPU path.

The packed layout is MuJoCo's mj_getState order (time, qpos, qvel, act, qacc_warmstart, ctrl, qfrc_applied, xfrc_applied,
eq_active, mocap_pos, mocap_quat), which is what the reference's tests compare against (support_test.py:121-180).  Two
slips of the reference's kernels are NOT reproduced: `adr += j` after eq_active (support.py:663, :799: the cursor advances by
neq - 1) and the swapped x / y of mocap_pos in set_state (support.py:803-806) -- with either, set_state(get_state(d)) would
not round-trip.
"""

from __future__ import annotations

import ctypes

import torch

from . import _lib
from . import constants as C
from . import types
from .forward import _ensure_bound


def state_size(m: types.Model, d: types.Data, sig: int) -> int:
  """Floats per world of the state components selected by `sig` (MuJoCo mj_stateSize)."""
  if int(sig) >= (1 << int(C.State.NSTATE)) or int(sig) < 0:
    raise ValueError(f"invalid state signature {sig} >= 2^mjNSTATE")
  s = m.sizes
  n = {C.State.TIME: 1, C.State.QPOS: s["nq"], C.State.QVEL: s["nv"], C.State.ACT: s["na"], C.State.WARMSTART: s["nv"], C.State.CTRL: s["nu"],
       C.State.QFRC_APPLIED: s["nv"], C.State.XFRC_APPLIED: 6 * s["nbody"], C.State.EQ_ACTIVE: s["neq"], C.State.MOCAP_POS: 3 * s["nmocap"],
       C.State.MOCAP_QUAT: 4 * s["nmocap"]}  # fmt: skip
  return sum(v for k, v in n.items() if int(sig) & int(k))


def _mask_ptr(d, active, keep):
  if active is None:
    return ctypes.c_void_p(0)
  a = torch.as_tensor(active, device=d.device)
  if a.shape != (d.nworld,):
    raise ValueError(f"active must have shape ({d.nworld},), got {tuple(a.shape)}")
  a = a.to(torch.uint8) if a.dtype != torch.bool else a.view(torch.uint8)
  a = a.contiguous()
  keep.append(a)
  return ctypes.c_void_p(a.data_ptr())


def _state_call(fn_name, m, d, state, sig, active):
  if int(sig) >= (1 << int(C.State.NSTATE)) or int(sig) < 0:
    raise ValueError(f"invalid state signature {sig} >= 2^mjNSTATE")
  n = state_size(m, d, sig)
  if not (torch.is_tensor(state) and state.dtype == torch.float32 and state.dim() == 2 and state.stride(1) == 1):
    raise ValueError("state must be a float32 tensor of shape (nworld, >= state size) with contiguous rows")
  # a pinned host tensor is device-accessible (mapped, unified addressing): the kernel then reads / writes host memory directly
  # over PCIe, which saves the staging buffer and the DMA copy of an observation / action exchange with the host
  on_device = state.device == d.qpos.device or (state.device.type == "cpu" and state.is_pinned())
  if state.shape[0] != d.nworld or state.shape[1] < n or not on_device:
    raise ValueError(f"state must have shape ({d.nworld}, >= {n}) on {d.qpos.device} (or in pinned host memory), got {tuple(state.shape)} on {state.device}")
  lib = _lib.load()
  _ensure_bound(m, d)
  keep = []
  _lib.check(lib, getattr(lib, fn_name)(m._handle, d._handle, ctypes.c_void_p(state.data_ptr()), int(state.stride(0)), int(sig),
                                        _mask_ptr(d, active, keep), _lib.stream_ptr()))  # fmt: skip


def get_state(m: types.Model, d: types.Data, state: torch.Tensor, sig: int, active=None):
  """Copies the concatenated state components selected by `sig` (bits of `State`) from Data into `state` (nworld, nstate).

  active: optional per-world bool mask; rows of inactive worlds are left untouched (reference support.py:572)."""
  _state_call("mjb_get_state", m, d, state, sig, active)


def set_state(m: types.Model, d: types.Data, state: torch.Tensor, sig: int, active=None):
  """Copies the concatenated state components selected by `sig` from `state` into Data (reference support.py:711)."""
  _state_call("mjb_set_state", m, d, state, sig, active)


def contact_force(m: types.Model, d: types.Data, contact_ids: torch.Tensor, to_world_frame: bool, force: torch.Tensor):
  """6D force:torque of the contacts `contact_ids` (int32, device) into `force` (n, 6), in the contact frame or rotated into
  the world frame (reference support.py:354).  Rows of ids >= nacon are not written."""
  n = int(contact_ids.numel())
  if contact_ids.dtype != torch.int32 or not contact_ids.is_contiguous():
    raise ValueError("contact_ids must be a contiguous int32 tensor")
  if force.dtype != torch.float32 or not force.is_contiguous() or force.numel() < 6 * n:
    raise ValueError("force must be a contiguous float32 tensor of shape (n, 6)")
  lib = _lib.load()
  _ensure_bound(m, d)
  _lib.check(lib, lib.mjb_contact_force(m._handle, d._handle, int(m.opt.cone), ctypes.c_void_p(contact_ids.data_ptr()), n,
                                        int(bool(to_world_frame)), ctypes.c_void_p(force.data_ptr()), _lib.stream_ptr()))  # fmt: skip
