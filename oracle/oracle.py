"""Python wrapper of the CPU oracle (oracle/mjoracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  It may import the product's host tables (mujoco_warp_b200.io.host_model,
types) to share field names and shapes; the product never imports it.

Parity status: PINNED to outputs of the reference itself (its unmodified sources executed here in a
float64 numpy interpreter of Warp, golden vectors under tests/golden/); see the header of mjoracle.c
and DESIGN.md section 4.
"""

from __future__ import annotations

import ctypes
import os
import subprocess
from types import SimpleNamespace

import numpy as np

from mujoco_warp_b200 import io as _io
from mujoco_warp_b200 import types as _types

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(_HERE, "mjoracle.c")
LIB = os.path.join(_HERE, "_build", "libmjoracle.so")

STAGE = {
  "kinematics": 1 << 0, "com_pos": 1 << 1, "crb": 1 << 2, "factor_m": 1 << 3, "collision": 1 << 4,
  "make_constraint": 1 << 5, "transmission": 1 << 6, "fwd_velocity": 1 << 7, "fwd_actuation": 1 << 8,
  "fwd_acceleration": 1 << 9, "fwd_acc_factor": 1 << 10, "solve": 1 << 11, "euler": 1 << 12,
  "implicit": 1 << 13, "zero_sensor": 1 << 14,
}  # fmt: skip
STAGE["fwd_position"] = STAGE["kinematics"] | STAGE["com_pos"] | STAGE["crb"] | STAGE["collision"] | STAGE["make_constraint"] | STAGE["transmission"]
STAGE["forward"] = STAGE["fwd_position"] | STAGE["zero_sensor"] | STAGE["fwd_velocity"] | STAGE["fwd_actuation"] | STAGE["fwd_acc_factor"] | STAGE["solve"]
STAGE["step"] = STAGE["forward"] | STAGE["euler"]
STAGE["step_implicitfast"] = STAGE["forward"] | STAGE["implicit"]

_lib = None


def build(force: bool = False) -> str:
  os.makedirs(os.path.dirname(LIB), exist_ok=True)
  root = os.path.dirname(_HERE)
  deps = [SRC, os.path.join(root, "include", "mjb_fields.h"), os.path.join(root, "mujoco_warp_b200", "csrc", "mjb_ccd_core.h")]
  if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= max(os.path.getmtime(p) for p in deps):
    return LIB
  cmd = ["gcc", "-O3", "-march=x86-64-v3", "-fPIC", "-shared", "-fopenmp", "-o", LIB, SRC, "-lm"]
  res = subprocess.run(cmd, capture_output=True, text=True)
  if res.returncode != 0:
    raise RuntimeError("gcc failed:\n" + res.stderr)
  return LIB


def load():
  global _lib
  if _lib is None:
    build()
    _lib = ctypes.CDLL(LIB)
    _lib.mjo_field_name.restype = ctypes.c_char_p
    _lib.mjo_field_name.argtypes = [ctypes.c_int, ctypes.c_int]
    _lib.mjo_run_stages.restype = ctypes.c_int
  return _lib


class OracleModel:
  def __init__(self, mjm):
    self.hm = _io.host_model(mjm)
    self.sizes = dict(self.hm.sizes)
    self.sizes.update(self.hm.extra)
    self.opt = dict(self.hm.opt)
    self.meaninertia = self.hm.meaninertia
    self.mjm = mjm


def make_data(om: OracleModel, nworld=1, nconmax=None, njmax=None) -> SimpleNamespace:
  """float64 / int32 numpy arrays with the same names and logical shapes as the device Data."""
  sizes, _, _ = _io._data_sizes(om.mjm, om.hm, nworld, nconmax, None, njmax, None, None, None)
  d = SimpleNamespace(sizes=sizes, nworld=nworld)
  for name, shape in _types.DATA_INT_FIELDS.items():
    setattr(d, name, np.zeros(_types.resolve_shape(shape, sizes), dtype=np.int32))
  for name, shape in _types.DATA_REAL_FIELDS.items():
    setattr(d, name, np.zeros(_types.resolve_shape(shape, sizes), dtype=np.float64))
  d.contact_efc_address[:] = -1
  d.qpos[:] = om.hm.reals["qpos0"]
  k, gx, gm = _io._static_frames(om.mjm, om.hm)
  d.xquat[:], d.xmat[:], d.ximat[:], d.xpos[:], d.xipos[:] = k.xquat, k.xmat, k.ximat, k.xpos, k.xipos
  d.geom_xpos[:], d.geom_xmat[:] = gx, gm
  if sizes["neq"]:
    d.eq_active[:] = np.asarray(om.mjm.eq_active0).astype(np.int32)
  return d


def run(om: OracleModel, d, stages, w0=0, w1=None, nthread=1, opt=None):
  """Runs the given stage mask (int or stage name) over worlds [w0, w1)."""
  lib = load()
  if isinstance(stages, str):
    stages = STAGE[stages]
  w1 = d.nworld if w1 is None else w1
  o = dict(om.opt)
  if opt:
    o.update(opt)
  g = np.asarray(o["gravity"], dtype=np.float64).reshape(-1)
  flat = dict(o, gravity_x=g[0], gravity_y=g[1], gravity_z=g[2], meaninertia=om.meaninertia)
  msizes = (ctypes.c_int * len(_types.MODEL_SIZES))(*[int(om.sizes[n]) for n in _types.MODEL_SIZES])
  mint = (ctypes.c_void_p * len(_types.MODEL_INT_FIELDS))(*[om.hm.ints[n].ctypes.data for n in _types.MODEL_INT_FIELDS])
  mreal = (ctypes.c_void_p * len(_types.MODEL_REAL_FIELDS))(*[om.hm.reals[n].ctypes.data for n in _types.MODEL_REAL_FIELDS])
  oint = (ctypes.c_int * len(_types.OPT_INT_FIELDS))(*[int(flat[n]) for n in _types.OPT_INT_FIELDS])
  oreal = (ctypes.c_double * len(_types.OPT_REAL_FIELDS))(*[float(flat[n]) for n in _types.OPT_REAL_FIELDS])
  dsizes = (ctypes.c_int * len(_types.DATA_SIZES))(*[int(d.sizes[n]) for n in _types.DATA_SIZES])
  for n in list(_types.DATA_INT_FIELDS) + list(_types.DATA_REAL_FIELDS):
    a = getattr(d, n)
    assert a.flags["C_CONTIGUOUS"], n
  dint = (ctypes.c_void_p * len(_types.DATA_INT_FIELDS))(*[getattr(d, n).ctypes.data for n in _types.DATA_INT_FIELDS])
  dreal = (ctypes.c_void_p * len(_types.DATA_REAL_FIELDS))(*[getattr(d, n).ctypes.data for n in _types.DATA_REAL_FIELDS])
  rc = lib.mjo_run_stages(msizes, mint, mreal, oint, oreal, dsizes, dint, dreal, ctypes.c_uint32(stages), int(w0), int(w1), int(nthread))
  if rc != 0:
    raise RuntimeError(f"oracle error {rc}")
  return d


def check_tables():
  lib = load()
  want = [
    _types.MODEL_SIZES, list(_types.MODEL_INT_FIELDS), list(_types.MODEL_REAL_FIELDS), _types.OPT_INT_FIELDS,
    _types.OPT_REAL_FIELDS, _types.DATA_SIZES, list(_types.DATA_INT_FIELDS), list(_types.DATA_REAL_FIELDS),
  ]  # fmt: skip
  for kind, names in enumerate(want):
    got, i = [], 0
    while True:
      s = lib.mjo_field_name(kind, i)
      if not s:
        break
      got.append(s.decode())
      i += 1
    assert got == list(names), (kind, set(got) ^ set(names))


# ---- callers either side of `step` (test checker for mjb_get_state / mjb_set_state / mjb_reset_data / mjb_contact_force) ----
_STATE_FIELDS = ("time", "qpos", "qvel", "act", "qacc_warmstart", "ctrl", "qfrc_applied", "xfrc_applied", "eq_active", "mocap_pos", "mocap_quat")


def _state_segments(d, sig):
  """[(field, column offset, width)] of the packed state row: MuJoCo mj_getState order = the bit order of the reference's
  `State` enum (types.py:598-637), the layout support.py:572-708 writes (see mujoco_warp_b200/support.py on its two slips)."""
  segs, adr = [], 0
  for bit, name in enumerate(_STATE_FIELDS):
    if sig & (1 << bit):
      width = int(np.prod(getattr(d, name).shape[1:], dtype=np.int64))
      segs.append((name, adr, width))
      adr += width
  return segs, adr


def get_state(om, d, sig, active=None, state=None):
  """support.py:572 get_state -> (nworld, nstate) float64; rows of inactive worlds keep the values of `state` (zeros if None)."""
  segs, n = _state_segments(d, int(sig))
  out = np.zeros((d.nworld, n)) if state is None else state
  rows = np.ones(d.nworld, dtype=bool) if active is None else np.asarray(active, dtype=bool)
  for name, adr, width in segs:
    out[rows, adr : adr + width] = getattr(d, name).reshape(d.nworld, -1)[rows].astype(np.float64)
  return out


def set_state(om, d, state, sig, active=None):
  """support.py:711 set_state."""
  segs, _ = _state_segments(d, int(sig))
  rows = np.ones(d.nworld, dtype=bool) if active is None else np.asarray(active, dtype=bool)
  for name, adr, width in segs:
    f = getattr(d, name)
    v = np.asarray(state)[:, adr : adr + width]
    f.reshape(d.nworld, -1)[rows] = (v != 0).astype(f.dtype)[rows] if name == "eq_active" else v[rows]


def reset_data(om, d, reset=None):
  """io.py:1458-1640 reset_data: contacts of the reset worlds cleared in place, then the per-world defaults."""
  rows = np.ones(d.nworld, dtype=bool) if reset is None else np.asarray(reset, dtype=bool)
  hm, nacon = om.hm, int(d.nacon.reshape(-1)[0])
  wid = d.contact_worldid[:nacon]
  clear = np.nonzero((wid < 0) | rows[np.clip(wid, 0, d.nworld - 1)])[0]  # :1622-1626
  for name in ("dist", "pos", "frame", "includemargin", "friction", "solref", "solreffriction", "solimp", "dim", "geom", "worldid", "type", "geomcollisionid"):
    getattr(d, "contact_" + name)[clear] = 0
  d.contact_efc_address[clear] = -1
  for name in ("xfrc_applied", "qM", "solver_niter", "ne", "nf", "nl", "nefc", "time", "energy", "qvel", "qacc_warmstart", "qfrc_applied", "qacc", "ctrl"):
    getattr(d, name)[rows] = 0
  if rows[0]:
    d.nacon.reshape(-1)[0] = 0  # :1525 only a reset of world 0 zeroes the pooled counter
  d.qpos[rows] = hm.reals["qpos0"].reshape(-1)
  k = min(om.sizes["nu"], om.sizes["na"])  # :1541-1545
  d.act[rows, :k] = 0
  d.act_dot[rows, :k] = 0
  if om.sizes["neq"]:
    d.eq_active[rows] = hm.ints["eq_active0"]
  for b, mid in enumerate(hm.ints["body_mocapid"]):
    if mid >= 0:
      d.mocap_pos.reshape(d.nworld, -1, 3)[rows, mid] = hm.reals["body_pos"].reshape(-1, 3)[b]
      d.mocap_quat.reshape(d.nworld, -1, 4)[rows, mid] = hm.reals["body_quat"].reshape(-1, 4)[b]


def contact_force(om, d, contact_ids, to_world_frame, cone=None, force=None):
  """support.py:354 contact_force (kernel :317, contact_force_fn :263, _decode_pyramid :239) -> (n, 6)."""
  cone = int(om.opt["cone"]) if cone is None else int(cone)
  ids = np.asarray(contact_ids).reshape(-1)
  out = np.zeros((len(ids), 6)) if force is None else force
  nacon, njmax = int(d.nacon.reshape(-1)[0]), int(d.sizes["njmax"])
  adr_all = d.contact_efc_address.reshape(len(d.contact_dim), -1)
  for t, c in enumerate(ids):
    if c >= nacon:
      continue  # :331 row left untouched
    f = np.zeros(6)
    if c >= 0:
      w, condim, adr0 = int(d.contact_worldid[c]), int(d.contact_dim[c]), int(adr_all[c, 0])
      frc = d.efc_force[w]
      if adr0 >= 0:
        if cone == 0:
          if condim == 1:
            f[0] = frc[adr0]
          else:
            for i in range(condim - 1):
              a = adr0 + 2 * i
              d1 = frc[a] if a < njmax else 0.0
              d2 = frc[a + 1] if a + 1 < njmax else 0.0
              f[0] += d1 + d2
              f[i + 1] = (d1 - d2) * d.contact_friction[c, i]
        else:
          for i in range(min(condim, adr_all.shape[1])):
            a = int(adr_all[c, i])
            if 0 <= a < njmax:
              f[i] = frc[a]
      if to_world_frame:
        fr = d.contact_frame[c].reshape(3, 3)
        f = np.concatenate([f[:3] @ fr, f[3:] @ fr])
    out[t] = f
  return out
