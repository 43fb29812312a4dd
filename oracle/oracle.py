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
  hdr = os.path.join(os.path.dirname(_HERE), "include", "mjb_fields.h")
  if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= max(os.path.getmtime(SRC), os.path.getmtime(hdr)):
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
