"""Enums and constants of the MuJoCo data contract, hard-coded.

The reference imports these values from the `mujoco` package (reference
mujoco_warp/_src/types.py:23-28,115-637); `mujoco` is not available here so the numeric values of
MuJoCo's public headers (mjmodel.h) are restated.  The CUDA side mirrors them in csrc/mjb_types.h.
"""

import enum

MJ_MINVAL = 1e-15
MJ_MAXVAL = 1e10
MJ_MINIMP = 1e-4
MJ_MAXIMP = 0.9999
MJ_MINMU = 1e-5
MJ_MAXCONPAIR = 50


class DisableBit(enum.IntFlag):
  CONSTRAINT = 1 << 0
  EQUALITY = 1 << 1
  FRICTIONLOSS = 1 << 2
  LIMIT = 1 << 3
  CONTACT = 1 << 4
  SPRING = 1 << 5
  DAMPER = 1 << 6
  GRAVITY = 1 << 7
  CLAMPCTRL = 1 << 8
  WARMSTART = 1 << 9
  FILTERPARENT = 1 << 10
  ACTUATION = 1 << 11
  REFSAFE = 1 << 12
  SENSOR = 1 << 13
  MIDPHASE = 1 << 14
  EULERDAMP = 1 << 15
  AUTORESET = 1 << 16
  NATIVECCD = 1 << 17
  ISLAND = 1 << 18


class EnableBit(enum.IntFlag):
  OVERRIDE = 1 << 0
  ENERGY = 1 << 1
  FWDINV = 1 << 2
  INVDISCRETE = 1 << 3
  MULTICCD = 1 << 4
  SLEEP = 1 << 5


class JointType(enum.IntEnum):
  FREE = 0
  BALL = 1
  SLIDE = 2
  HINGE = 3


class GeomType(enum.IntEnum):
  PLANE = 0
  HFIELD = 1
  SPHERE = 2
  CAPSULE = 3
  ELLIPSOID = 4
  CYLINDER = 5
  BOX = 6
  MESH = 7
  SDF = 8


class IntegratorType(enum.IntEnum):
  EULER = 0
  RK4 = 1
  IMPLICIT = 2
  IMPLICITFAST = 3


class ConeType(enum.IntEnum):
  PYRAMIDAL = 0
  ELLIPTIC = 1


class SolverType(enum.IntEnum):
  PGS = 0
  CG = 1
  NEWTON = 2


class JacobianType(enum.IntEnum):
  DENSE = 0
  SPARSE = 1
  AUTO = 2


class ConstraintType(enum.IntEnum):
  EQUALITY = 0
  FRICTION_DOF = 1
  FRICTION_TENDON = 2
  LIMIT_JOINT = 3
  LIMIT_TENDON = 4
  CONTACT_FRICTIONLESS = 5
  CONTACT_PYRAMIDAL = 6
  CONTACT_ELLIPTIC = 7


class ConstraintState(enum.IntEnum):
  SATISFIED = 0
  QUADRATIC = 1
  LINEARNEG = 2
  LINEARPOS = 3
  CONE = 4


class TrnType(enum.IntEnum):
  JOINT = 0
  JOINTINPARENT = 1
  SLIDERCRANK = 2
  TENDON = 3
  SITE = 4
  BODY = 5


class DynType(enum.IntEnum):
  NONE = 0
  INTEGRATOR = 1
  FILTER = 2
  FILTEREXACT = 3
  MUSCLE = 4
  USER = 5


class GainType(enum.IntEnum):
  FIXED = 0
  AFFINE = 1
  MUSCLE = 2
  USER = 3


class BiasType(enum.IntEnum):
  NONE = 0
  AFFINE = 1
  MUSCLE = 2
  USER = 3


class EqType(enum.IntEnum):
  CONNECT = 0
  WELD = 1
  JOINT = 2
  TENDON = 3
  FLEX = 4
  DISTANCE = 5


class BroadphaseType(enum.IntEnum):
  NXN = 0
  SAP_TILE = 1
  SAP_SEGMENTED = 2


class BroadphaseFilter(enum.IntFlag):
  PLANE = 1
  SPHERE = 2
  AABB = 4
  OBB = 8


class State(enum.IntEnum):
  """State component bitflags (reference types.py:598-637 = MuJoCo mjtState)."""

  TIME = 1 << 0
  QPOS = 1 << 1
  QVEL = 1 << 2
  ACT = 1 << 3
  WARMSTART = 1 << 4
  CTRL = 1 << 5
  QFRC_APPLIED = 1 << 6
  XFRC_APPLIED = 1 << 7
  EQ_ACTIVE = 1 << 8
  MOCAP_POS = 1 << 9
  MOCAP_QUAT = 1 << 10
  NSTATE = 13
  PHYSICS = (1 << 1) | (1 << 2) | (1 << 3)
  FULLPHYSICS = (1 << 0) | (1 << 1) | (1 << 2) | (1 << 3) | (1 << 12)
  USER = (1 << 5) | (1 << 6) | (1 << 7) | (1 << 8) | (1 << 9) | (1 << 10) | (1 << 11)
  INTEGRATION = (1 << 13) - 1
  # unsupported (carry no data): USERDATA = 1 << 11, PLUGIN = 1 << 12


class ContactType(enum.IntFlag):
  CONSTRAINT = 1
  SENSOR = 2
