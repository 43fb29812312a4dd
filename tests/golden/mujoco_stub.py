"""A stand-in for the `mujoco` Python module, just large enough for the reference's `types.py`, `io.put_model` and
`io.make_data` to import and run against this repo's `MjModelLite` (mujoco itself is not installable here).

It provides (a) the numeric values of the public MuJoCo enums / constants the reference reads at import time
(mjmodel.h / mjdata.h; values that the step path never compares against anything outside the reference itself are
numbered in declaration order), (b) `mj_isSparse`, `mj_name2id`, `mj_id2name`, and (c) `MjData` + `mj_kinematics`
for the static-geom poses `make_data` precomputes (reference io.py:958-959) — computed with the standard forward
kinematics recursion on the host model.

Only the golden-vector generators import this file.
"""

from __future__ import annotations

import sys
import types as _types

import numpy as np

_ENUMS = {
  "mjtDisableBit": dict(
    mjDSBL_CONSTRAINT=1 << 0, mjDSBL_EQUALITY=1 << 1, mjDSBL_FRICTIONLOSS=1 << 2, mjDSBL_LIMIT=1 << 3, mjDSBL_CONTACT=1 << 4,
    mjDSBL_SPRING=1 << 5, mjDSBL_DAMPER=1 << 6, mjDSBL_GRAVITY=1 << 7, mjDSBL_CLAMPCTRL=1 << 8, mjDSBL_WARMSTART=1 << 9,
    mjDSBL_FILTERPARENT=1 << 10, mjDSBL_ACTUATION=1 << 11, mjDSBL_REFSAFE=1 << 12, mjDSBL_SENSOR=1 << 13, mjDSBL_MIDPHASE=1 << 14,
    mjDSBL_EULERDAMP=1 << 15, mjDSBL_AUTORESET=1 << 16, mjDSBL_NATIVECCD=1 << 17, mjDSBL_ISLAND=1 << 18,
  ),
  "mjtEnableBit": dict(mjENBL_OVERRIDE=1, mjENBL_ENERGY=2, mjENBL_FWDINV=4, mjENBL_INVDISCRETE=8, mjENBL_MULTICCD=16, mjENBL_SLEEP=32),
  "mjtJoint": dict(mjJNT_FREE=0, mjJNT_BALL=1, mjJNT_SLIDE=2, mjJNT_HINGE=3),
  "mjtGeom": dict(mjGEOM_PLANE=0, mjGEOM_HFIELD=1, mjGEOM_SPHERE=2, mjGEOM_CAPSULE=3, mjGEOM_ELLIPSOID=4, mjGEOM_CYLINDER=5,
                  mjGEOM_BOX=6, mjGEOM_MESH=7, mjGEOM_SDF=8, mjGEOM_FLEX=105),
  "mjtCamLight": dict(mjCAMLIGHT_FIXED=0, mjCAMLIGHT_TRACK=1, mjCAMLIGHT_TRACKCOM=2, mjCAMLIGHT_TARGETBODY=3, mjCAMLIGHT_TARGETBODYCOM=4),
  "mjtProjection": dict(mjPROJ_PERSPECTIVE=0, mjPROJ_ORTHOGRAPHIC=1),
  "mjtIntegrator": dict(mjINT_EULER=0, mjINT_RK4=1, mjINT_IMPLICIT=2, mjINT_IMPLICITFAST=3),
  "mjtCone": dict(mjCONE_PYRAMIDAL=0, mjCONE_ELLIPTIC=1),
  "mjtJacobian": dict(mjJAC_DENSE=0, mjJAC_SPARSE=1, mjJAC_AUTO=2),
  "mjtSolver": dict(mjSOL_PGS=0, mjSOL_CG=1, mjSOL_NEWTON=2),
  "mjtEq": dict(mjEQ_CONNECT=0, mjEQ_WELD=1, mjEQ_JOINT=2, mjEQ_TENDON=3, mjEQ_FLEX=4, mjEQ_DISTANCE=5),
  "mjtWrap": dict(mjWRAP_NONE=0, mjWRAP_JOINT=1, mjWRAP_PULLEY=2, mjWRAP_SITE=3, mjWRAP_SPHERE=4, mjWRAP_CYLINDER=5),
  "mjtTrn": dict(mjTRN_JOINT=0, mjTRN_JOINTINPARENT=1, mjTRN_SLIDERCRANK=2, mjTRN_TENDON=3, mjTRN_SITE=4, mjTRN_BODY=5, mjTRN_UNDEFINED=1000),
  "mjtDyn": dict(mjDYN_NONE=0, mjDYN_INTEGRATOR=1, mjDYN_FILTER=2, mjDYN_FILTEREXACT=3, mjDYN_MUSCLE=4, mjDYN_USER=5),
  "mjtGain": dict(mjGAIN_FIXED=0, mjGAIN_AFFINE=1, mjGAIN_MUSCLE=2, mjGAIN_USER=3),
  "mjtBias": dict(mjBIAS_NONE=0, mjBIAS_AFFINE=1, mjBIAS_MUSCLE=2, mjBIAS_USER=3),
  "mjtObj": dict(mjOBJ_UNKNOWN=0, mjOBJ_BODY=1, mjOBJ_XBODY=2, mjOBJ_JOINT=3, mjOBJ_DOF=4, mjOBJ_GEOM=5, mjOBJ_SITE=6, mjOBJ_CAMERA=7,
                 mjOBJ_LIGHT=8, mjOBJ_FLEX=9, mjOBJ_MESH=10, mjOBJ_SKIN=11, mjOBJ_HFIELD=12, mjOBJ_TEXTURE=13, mjOBJ_MATERIAL=14,
                 mjOBJ_PAIR=15, mjOBJ_EXCLUDE=16, mjOBJ_EQUALITY=17, mjOBJ_TENDON=18, mjOBJ_ACTUATOR=19, mjOBJ_SENSOR=20,
                 mjOBJ_NUMERIC=21, mjOBJ_TEXT=22, mjOBJ_TUPLE=23, mjOBJ_KEY=24, mjOBJ_PLUGIN=25),
  "mjtConstraint": dict(mjCNSTR_EQUALITY=0, mjCNSTR_FRICTION_DOF=1, mjCNSTR_FRICTION_TENDON=2, mjCNSTR_LIMIT_JOINT=3,
                        mjCNSTR_LIMIT_TENDON=4, mjCNSTR_CONTACT_FRICTIONLESS=5, mjCNSTR_CONTACT_PYRAMIDAL=6, mjCNSTR_CONTACT_ELLIPTIC=7),
  "mjtConstraintState": dict(mjCNSTRSTATE_SATISFIED=0, mjCNSTRSTATE_QUADRATIC=1, mjCNSTRSTATE_LINEARNEG=2, mjCNSTRSTATE_LINEARPOS=3,
                             mjCNSTRSTATE_CONE=4),
  "mjtStage": dict(mjSTAGE_NONE=0, mjSTAGE_POS=1, mjSTAGE_VEL=2, mjSTAGE_ACC=3),
  "mjtDataType": dict(mjDATATYPE_REAL=0, mjDATATYPE_POSITIVE=1, mjDATATYPE_AXIS=2, mjDATATYPE_QUATERNION=3),
  "mjtState": dict(
    mjSTATE_TIME=1 << 0, mjSTATE_QPOS=1 << 1, mjSTATE_QVEL=1 << 2, mjSTATE_ACT=1 << 3, mjSTATE_WARMSTART=1 << 4, mjSTATE_CTRL=1 << 5,
    mjSTATE_QFRC_APPLIED=1 << 6, mjSTATE_XFRC_APPLIED=1 << 7, mjSTATE_EQ_ACTIVE=1 << 8, mjSTATE_MOCAP_POS=1 << 9, mjSTATE_MOCAP_QUAT=1 << 10,
    mjSTATE_USERDATA=1 << 11, mjSTATE_PLUGIN=1 << 12, mjNSTATE=13, mjSTATE_PHYSICS=(1 << 1) | (1 << 2) | (1 << 3),
    mjSTATE_FULLPHYSICS=(1 << 0) | (1 << 1) | (1 << 2) | (1 << 3) | (1 << 12),
    mjSTATE_USER=(1 << 5) | (1 << 6) | (1 << 7) | (1 << 8) | (1 << 9) | (1 << 10) | (1 << 11), mjSTATE_INTEGRATION=(1 << 13) - 1,
  ),
}
_AUTO_BITS = set()


class _Enum:
  """Known names resolve to the tabulated value; unknown names of declaration-ordered enums are numbered on first use."""

  def __init__(self, name):
    self._name = name
    self._tab = dict(_ENUMS.get(name, {}))
    self._auto = name not in _ENUMS

  def __getattr__(self, k):
    if k.startswith("_"):
      raise AttributeError(k)
    if k not in self._tab:
      if not self._auto:
        raise AttributeError(f"mujoco stub: {self._name}.{k} is not tabulated")
      n = len(self._tab)
      self._tab[k] = (1 << n) if self._name in _AUTO_BITS else n
    return self._tab[k]


class MjData:
  """Holds only what reference io.make_data reads after mj_kinematics (io.py:958-995)."""

  def __init__(self, mjm):
    nb, ng = mjm.nbody, mjm.ngeom
    self.xpos, self.xquat = np.zeros((nb, 3)), np.zeros((nb, 4))
    self.xmat, self.ximat = np.zeros((nb, 9)), np.zeros((nb, 9))
    self.xipos = np.zeros((nb, 3))
    self.geom_xpos, self.geom_xmat = np.zeros((ng, 3)), np.zeros((ng, 9))
    self.qpos = np.array(mjm.qpos0, dtype=np.float64)
    self.mocap_pos = np.zeros((mjm.nmocap, 3))
    self.mocap_quat = np.zeros((mjm.nmocap, 4))


def _qmul(a, b):
  return np.array([
    a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
    a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
    a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
    a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0],
  ])


def _q2m(q):
  w, x, y, z = q
  return np.array([
    [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
    [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
    [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)],
  ])


def mj_kinematics(mjm, mjd):
  """Forward kinematics at mjd.qpos (textbook recursion; free / ball / slide / hinge joints)."""
  q = mjd.qpos
  for b in range(mjm.nbody):
    if b == 0:
      pos, quat = np.zeros(3), np.array([1.0, 0, 0, 0])
    else:
      p = int(mjm.body_parentid[b])
      pos = mjd.xpos[p] + _q2m(mjd.xquat[p]) @ mjm.body_pos[b]
      quat = _qmul(mjd.xquat[p], mjm.body_quat[b])
      for j in range(int(mjm.body_jntadr[b]), int(mjm.body_jntadr[b]) + int(mjm.body_jntnum[b])):
        t, a = int(mjm.jnt_type[j]), int(mjm.jnt_qposadr[j])
        if t == 0:
          pos, quat = q[a : a + 3].copy(), q[a + 3 : a + 7] / np.linalg.norm(q[a + 3 : a + 7])
        elif t == 1:
          anchor = pos + _q2m(quat) @ mjm.jnt_pos[j]
          quat = _qmul(quat, q[a : a + 4] / np.linalg.norm(q[a : a + 4]))
          pos = anchor - _q2m(quat) @ mjm.jnt_pos[j]
        elif t == 2:
          pos = pos + _q2m(quat) @ mjm.jnt_axis[j] * (q[a] - mjm.qpos0[a])
        else:
          anchor = pos + _q2m(quat) @ mjm.jnt_pos[j]
          ang = q[a] - mjm.qpos0[a]
          ax = mjm.jnt_axis[j]
          quat = _qmul(quat, np.concatenate(([np.cos(ang / 2)], np.sin(ang / 2) * ax)))
          pos = anchor - _q2m(quat) @ mjm.jnt_pos[j]
      quat = quat / np.linalg.norm(quat)
    mjd.xpos[b], mjd.xquat[b] = pos, quat
    mjd.xmat[b] = _q2m(quat).reshape(-1)
    mjd.xipos[b] = pos + _q2m(quat) @ mjm.body_ipos[b]
    mjd.ximat[b] = _q2m(_qmul(quat, mjm.body_iquat[b])).reshape(-1)
  for g in range(mjm.ngeom):
    b = int(mjm.geom_bodyid[g])
    mjd.geom_xpos[g] = mjd.xpos[b] + _q2m(mjd.xquat[b]) @ mjm.geom_pos[g]
    mjd.geom_xmat[g] = _q2m(_qmul(mjd.xquat[b], mjm.geom_quat[g])).reshape(-1)


def build():
  mj = _types.ModuleType("mujoco")
  mj.mjMINVAL, mj.mjMAXVAL, mj.mjMINIMP, mj.mjMAXIMP, mj.mjMAXCONPAIR, mj.mjMINMU = 1e-15, 1e10, 1e-4, 0.9999, 50, 1e-5
  mj.mjNEQDATA, mj.mjNDYN, mj.mjNGAIN, mj.mjNBIAS, mj.mjNREF, mj.mjNIMP, mj.mjNSOLVER = 11, 10, 10, 10, 2, 5, 200
  mj.mjNFLUID = 12
  _cache = {}

  def _getattr(name):
    if name.startswith("mjt"):
      if name not in _cache:
        _cache[name] = _Enum(name)
      return _cache[name]
    raise AttributeError(f"mujoco stub has no attribute {name}")

  mj.__getattr__ = _getattr
  mj.MjModel = type("MjModel", (), {})
  mj.MjData = MjData
  mj.MjSpec = type("MjSpec", (), {})
  mj.mj_kinematics = mj_kinematics
  mj.mj_isSparse = lambda m: int(m.opt.jacobian == 1 or (m.opt.jacobian == 2 and m.nv >= 60))
  mj.mj_name2id = lambda m, objtype, name: -1
  mj.mj_id2name = lambda m, objtype, i: None
  mj._is_stub = True
  return mj


def install():
  if "mujoco" in sys.modules and not getattr(sys.modules["mujoco"], "_is_stub", False):
    raise RuntimeError("a real mujoco is importable: use it instead of the stub")
  mj = build()
  sys.modules["mujoco"] = mj
  return mj
