"""Data contract of the batched step: Option / Model / Data / Contact / Constraint.

Mirrors the reference dataclasses (reference mujoco_warp/_src/types.py:706 Option, :834 Model,
:1618 Contact, :1659 Constraint, :1703 Data): same field names, same logical shapes and dtypes.
Array fields are `torch.Tensor`s (float32 / int32) instead of Warp arrays; Warp vector element
types become trailing dimensions (vec3 -> 3, quat -> 4 (w,x,y,z), mat33 -> (3,3),
spatial_vector -> 6, vec5 -> 5, vec10 -> 10, vec2 -> 2).

The tables below give the logical shape of every field that crosses the C ABI.  Their names and
order come from include/mjb_fields.h through `mjb_field_name` (checked in tests/test_abi.py).
"""

from __future__ import annotations

import dataclasses
from typing import Any, Callable, Optional

from .constants import (  # noqa: F401  (re-exported like the reference's types module)
  MJ_MAXCONPAIR,
  MJ_MAXIMP,
  MJ_MAXVAL,
  MJ_MINIMP,
  MJ_MINMU,
  MJ_MINVAL,
  BiasType,
  BroadphaseFilter,
  BroadphaseType,
  ConeType,
  ConstraintState,
  ConstraintType,
  ContactType,
  DisableBit,
  DynType,
  EnableBit,
  EqType,
  GainType,
  GeomType,
  IntegratorType,
  JacobianType,
  JointType,
  SolverType,
  TrnType,
)

TILE_SIZE_JTDAJ_SPARSE = 16  # reference types.py:34
TILE_SIZE_JTDAJ_DENSE = 16  # reference types.py:35

# ----------------------------------------------------------------------------------------------
# field tables (logical shapes; a str names a size, an int is literal)
# ----------------------------------------------------------------------------------------------
MODEL_SIZES = (
  "nq nv nu na nbody njnt ngeom nsite nmocap neq npair nxn nlevel nv_pad nlimit nlimitball "
  "nfrictiondof neqjnt nJmom nmaxcondim nmaxpyramid ngravcomp is_sparse nM ntree ngeompair nmesh nmeshvert nmeshgraph neqconnect neqweld blob_hot_bytes nhfield nhfielddata ngroup ngrouppair"
).split()

MODEL_INT_FIELDS = {
  "body_parentid": ("nbody",),
  "body_rootid": ("nbody",),
  "body_weldid": ("nbody",),
  "body_mocapid": ("nbody",),
  "eq_active0": ("neq",),
  "body_jntnum": ("nbody",),
  "body_jntadr": ("nbody",),
  "body_dofnum": ("nbody",),
  "body_dofadr": ("nbody",),
  "body_subtreeend": ("nbody",),
  "body_tree": ("nbody",),
  "body_tree_adr": ("nlevel+1",),
  "jnt_type": ("njnt",),
  "jnt_qposadr": ("njnt",),
  "jnt_dofadr": ("njnt",),
  "jnt_bodyid": ("njnt",),
  "jnt_limited_adr": ("nlimit",),
  "jnt_limited_ball_adr": ("nlimitball",),
  "jnt_actfrclimited": ("njnt",),
  "jnt_actgravcomp": ("njnt",),
  "dof_bodyid": ("nv",),
  "dof_jntid": ("nv",),
  "dof_parentid": ("nv",),
  "dof_friction_adr": ("nfrictiondof",),
  "dof_tree0": ("nv",),
  "geom_type": ("ngeom",),
  "geom_bodyid": ("ngeom",),
  "geom_condim": ("ngeom",),
  "geom_priority": ("ngeom",),
  "geom_static": ("ngeom",),
  "site_bodyid": ("nsite",),
  "nxn_pair": ("nxn", 2),
  "nxn_pairid": ("nxn", 2),
  "nxn_lookup": ("ngeompair",),
  "pair_dim": ("npair",),
  "actuator_trntype": ("nu",),
  "actuator_trnid": ("nu", 2),
  "actuator_dyntype": ("nu",),
  "actuator_gaintype": ("nu",),
  "actuator_biastype": ("nu",),
  "actuator_actadr": ("nu",),
  "actuator_actnum": ("nu",),
  "actuator_ctrllimited": ("nu",),
  "actuator_forcelimited": ("nu",),
  "actuator_actlimited": ("nu",),
  "actuator_actearly": ("nu",),
  "eq_type": ("neq",),
  "eq_obj1id": ("neq",),
  "eq_obj2id": ("neq",),
  "eq_objtype": ("neq",),
  "eq_jnt_adr": ("neqjnt",),
  "eq_connect_adr": ("neqconnect",),
  "eq_wld_adr": ("neqweld",),
  "geom_dataid": ("ngeom",),
  "mesh_vertadr": ("nmesh",),
  "mesh_vertnum": ("nmesh",),
  "mesh_graphadr": ("nmesh",),
  "mesh_graph": ("nmeshgraph",),
  "hfield_nrow": ("nhfield",),
  "hfield_ncol": ("nhfield",),
  "hfield_adr": ("nhfield",),
  "group_bodyid": ("ngroup",),
  "grouppair": ("ngrouppair", 2),
  "grouppair_adr": ("ngrouppair+1",),
  "grouppair_always": ("ngrouppair",),
}

# reference batches these over a leading `*` dimension (1 or nworld); here always 1
MODEL_REAL_FIELDS = {
  "qpos0": ("nq",),
  "qpos_spring": ("nq",),
  "body_pos": ("nbody", 3),
  "body_quat": ("nbody", 4),
  "body_ipos": ("nbody", 3),
  "body_iquat": ("nbody", 4),
  "body_mass": ("nbody",),
  "body_subtreemass": ("nbody",),
  "body_inertia": ("nbody", 3),
  "body_invweight0": ("nbody", 2),
  "body_gravcomp": ("nbody",),
  "jnt_pos": ("njnt", 3),
  "jnt_axis": ("njnt", 3),
  "jnt_stiffness": ("njnt",),
  "jnt_range": ("njnt", 2),
  "jnt_margin": ("njnt",),
  "jnt_solref": ("njnt", 2),
  "jnt_solimp": ("njnt", 5),
  "jnt_actfrcrange": ("njnt", 2),
  "dof_armature": ("nv",),
  "dof_damping": ("nv",),
  "dof_invweight0": ("nv",),
  "dof_frictionloss": ("nv",),
  "dof_solref": ("nv", 2),
  "dof_solimp": ("nv", 5),
  "geom_size": ("ngeom", 3),
  "geom_pos": ("ngeom", 3),
  "geom_quat": ("ngeom", 4),
  "geom_friction": ("ngeom", 3),
  "geom_solmix": ("ngeom",),
  "geom_solref": ("ngeom", 2),
  "geom_solimp": ("ngeom", 5),
  "geom_margin": ("ngeom",),
  "geom_gap": ("ngeom",),
  "geom_rbound": ("ngeom",),
  "geom_aabb": ("ngeom", 2, 3),
  "site_pos": ("nsite", 3),
  "site_quat": ("nsite", 4),
  "pair_solref": ("npair", 2),
  "pair_solreffriction": ("npair", 2),
  "pair_solimp": ("npair", 5),
  "pair_margin": ("npair",),
  "pair_gap": ("npair",),
  "pair_friction": ("npair", 5),
  "actuator_dynprm": ("nu", 10),
  "actuator_gainprm": ("nu", 10),
  "actuator_biasprm": ("nu", 10),
  "actuator_ctrlrange": ("nu", 2),
  "actuator_forcerange": ("nu", 2),
  "actuator_actrange": ("nu", 2),
  "actuator_gear": ("nu", 6),
  "eq_solref": ("neq", 2),
  "eq_solimp": ("neq", 5),
  "eq_data": ("neq", 11),
  "mesh_vert": ("nmeshvert", 3),
  "hfield_size": ("nhfield", 4),
  "hfield_data": ("nhfielddata",),
  "group_sphere": ("ngroup", 4),
}

OPT_INT_FIELDS = (
  "integrator cone solver iterations ls_iterations disableflags enableflags broadphase broadphase_filter run_collision_detection ccd_iterations"
).split()
OPT_REAL_FIELDS = "timestep tolerance ls_tolerance gravity_x gravity_y gravity_z impratio_invsqrt meaninertia ccd_tolerance".split()

DATA_SIZES = "nworld naconmax njmax njmax_pad nwconmax".split()

# name -> shape; names with a contact_/efc_ prefix live in Data.contact / Data.efc
DATA_INT_FIELDS = {
  "solver_niter": ("nworld",),
  "ne": ("nworld",),
  "nf": ("nworld",),
  "nl": ("nworld",),
  "nefc": ("nworld",),
  "nacon": (1,),
  "ncollision": (1,),
  "eq_active": ("nworld", "neq"),
  "moment_rownnz": ("nworld", "nu"),
  "moment_rowadr": ("nworld", "nu"),
  "moment_colind": ("nworld", "nJmom"),
  "contact_dim": ("naconmax",),
  "contact_geom": ("naconmax", 2),
  "contact_efc_address": ("naconmax", "nmaxpyramid"),
  "contact_worldid": ("naconmax",),
  "contact_type": ("naconmax",),
  "contact_geomcollisionid": ("naconmax",),
  "efc_type": ("nworld", "njmax"),
  "efc_id": ("nworld", "njmax"),
  "efc_state": ("nworld", "njmax_pad"),
  "wcon_n": ("nworld",),
  "wcon": ("nworld", "nwconmax"),
  "status": ("nworld",),
}

DATA_REAL_FIELDS = {
  "time": ("nworld",),
  "energy": ("nworld", 2),
  "qpos": ("nworld", "nq"),
  "qvel": ("nworld", "nv"),
  "act": ("nworld", "na"),
  "qacc_warmstart": ("nworld", "nv"),
  "ctrl": ("nworld", "nu"),
  "qfrc_applied": ("nworld", "nv"),
  "xfrc_applied": ("nworld", "nbody", 6),
  "mocap_pos": ("nworld", "nmocap", 3),
  "mocap_quat": ("nworld", "nmocap", 4),
  "qacc": ("nworld", "nv"),
  "act_dot": ("nworld", "na"),
  "xpos": ("nworld", "nbody", 3),
  "xquat": ("nworld", "nbody", 4),
  "xmat": ("nworld", "nbody", 3, 3),
  "xipos": ("nworld", "nbody", 3),
  "ximat": ("nworld", "nbody", 3, 3),
  "xanchor": ("nworld", "njnt", 3),
  "xaxis": ("nworld", "njnt", 3),
  "geom_xpos": ("nworld", "ngeom", 3),
  "geom_xmat": ("nworld", "ngeom", 3, 3),
  "site_xpos": ("nworld", "nsite", 3),
  "site_xmat": ("nworld", "nsite", 3, 3),
  "subtree_com": ("nworld", "nbody", 3),
  "cdof": ("nworld", "nv", 6),
  "cinert": ("nworld", "nbody", 10),
  "actuator_length": ("nworld", "nu"),
  "actuator_moment": ("nworld", "nJmom"),
  "crb": ("nworld", "nbody", 10),
  "qM": ("nworld", "nv_pad", "nv_pad"),
  "qLD": ("nworld", "nv", "nv"),
  "qLDiagInv": ("nworld", "nv"),
  "actuator_velocity": ("nworld", "nu"),
  "cvel": ("nworld", "nbody", 6),
  "cdof_dot": ("nworld", "nv", 6),
  "qfrc_bias": ("nworld", "nv"),
  "qfrc_spring": ("nworld", "nv"),
  "qfrc_damper": ("nworld", "nv"),
  "qfrc_gravcomp": ("nworld", "nv"),
  "qfrc_fluid": ("nworld", "nv"),
  "qfrc_passive": ("nworld", "nv"),
  "actuator_force": ("nworld", "nu"),
  "qfrc_actuator": ("nworld", "nv"),
  "qfrc_smooth": ("nworld", "nv"),
  "qacc_smooth": ("nworld", "nv"),
  "qfrc_constraint": ("nworld", "nv"),
  "cacc": ("nworld", "nbody", 6),
  "cfrc_int": ("nworld", "nbody", 6),
  "cfrc_ext": ("nworld", "nbody", 6),
  "contact_dist": ("naconmax",),
  "contact_pos": ("naconmax", 3),
  "contact_frame": ("naconmax", 3, 3),
  "contact_includemargin": ("naconmax",),
  "contact_friction": ("naconmax", 5),
  "contact_solref": ("naconmax", 2),
  "contact_solreffriction": ("naconmax", 2),
  "contact_solimp": ("naconmax", 5),
  "efc_J": ("nworld", "njmax_pad", "nv_pad"),
  "efc_pos": ("nworld", "njmax"),
  "efc_margin": ("nworld", "njmax"),
  "efc_D": ("nworld", "njmax_pad"),
  "efc_vel": ("nworld", "njmax"),
  "efc_aref": ("nworld", "njmax"),
  "efc_frictionloss": ("nworld", "njmax"),
  "efc_force": ("nworld", "njmax"),
  "efc_Ma": ("nworld", "nv"),
}

# reference Data fields that exist for API completeness but that no stage on the supported
# path computes (sensors, tendons, flex, cameras, islands, inverse dynamics): zero-filled.
DATA_EXTRA_FIELDS = {
  "sensordata": ("f", ("nworld", "nsensordata")),
  "subtree_linvel": ("f", ("nworld", "nbody", 3)),
  "subtree_angmom": ("f", ("nworld", "nbody", 3)),
  "qfrc_inverse": ("f", ("nworld", "nv")),
  "cam_xpos": ("f", ("nworld", "ncam", 3)),
  "cam_xmat": ("f", ("nworld", "ncam", 3, 3)),
  "light_xpos": ("f", ("nworld", "nlight", 3)),
  "light_xdir": ("f", ("nworld", "nlight", 3)),
  "ten_length": ("f", ("nworld", "ntendon")),
  "ten_velocity": ("f", ("nworld", "ntendon")),
  "ten_J": ("f", ("nworld", "nJten")),
  "nisland": ("i", ("nworld",)),
  "tree_island": ("i", ("nworld", "ntree")),
}


def resolve_shape(shape, sizes) -> tuple:
  out = []
  for s in shape:
    if isinstance(s, int):
      out.append(s)
    elif s.endswith("+1"):
      out.append(int(sizes[s[:-2]]) + 1)
    else:
      out.append(int(sizes[s]))
  return tuple(out)


def data_field_numel(name: str, sizes) -> Optional[int]:
  """Element count of an ABI data field for the given model + data sizes, or None if a symbol of its shape is unknown."""
  shape = DATA_INT_FIELDS.get(name) or DATA_REAL_FIELDS.get(name)
  if shape is None:
    return None
  try:
    dims = resolve_shape(shape, sizes)
  except KeyError:
    return None
  n = 1
  for x in dims:
    n *= x
  return n


# ----------------------------------------------------------------------------------------------
# dataclasses
# ----------------------------------------------------------------------------------------------
@dataclasses.dataclass
class Option:
  """Physics options (reference types.py:706-772).

  Array-typed options of the reference (`timestep`, `tolerance`, `ls_tolerance`, `gravity`,
  `impratio_invsqrt`, ...) are host tensors of leading dimension 1: they are passed by value
  to every kernel launch, so in-place edits take effect on the next call without a device
  sync.  Per-world option batching (leading dimension nworld) is not supported.
  """

  timestep: Any
  tolerance: Any
  ls_tolerance: Any
  ccd_tolerance: Any
  gravity: Any
  wind: Any
  magnetic: Any
  density: Any
  viscosity: Any
  integrator: int
  cone: int
  solver: int
  iterations: int
  ls_iterations: int
  ccd_iterations: int
  disableflags: int
  enableflags: int
  sdf_initpoints: int
  sdf_iterations: int
  impratio_invsqrt: Any
  ls_parallel: bool
  ls_parallel_min_step: float
  broadphase: int
  broadphase_filter: int
  graph_conditional: bool
  run_collision_detection: bool
  contact_sensor_maxmatch: int


@dataclasses.dataclass
class Statistic:
  """Model statistics (reference types.py:775-783)."""

  meaninertia: Any


@dataclasses.dataclass
class Callback:
  """User callbacks between stages (reference types.py:811-831)."""

  passive: Optional[Callable] = None
  control: Optional[Callable] = None
  act_dyn: Optional[Callable] = None
  act_gain: Optional[Callable] = None
  act_bias: Optional[Callable] = None
  sensor: Optional[Callable] = None
  contactfilter: Optional[Callable] = None


class Model:
  """Device model (reference types.py:834-1603).

  Every array in MODEL_INT_FIELDS / MODEL_REAL_FIELDS is an attribute that is a *view* into one
  contiguous device blob (`_blob`), which is what the kernels stage into shared memory; editing
  a field in place (e.g. `m.geom_friction[0, 3, 0] = 0.5`) is therefore seen by the next launch.
  Real fields carry the reference's leading `*` dimension (always 1 here).
  """

  def __init__(self):
    self.opt: Option = None
    self.stat: Statistic = None
    self.callback = Callback()
    self.sizes: dict = {}
    self._blob = None
    self._handle = None
    self._keepalive = []

  def __getattr__(self, name):
    # sizes are exposed as attributes (m.nv, m.nbody, ...)
    sizes = self.__dict__.get("sizes", {})
    if name in sizes:
      return sizes[name]
    raise AttributeError(name)

  def __setattr__(self, name, value):
    # The kernels read the blob, not this object's attributes: re-assigning a blob field after put_model (the reference's
    # domain-randomisation idiom `m.geom_friction = <(nworld, ngeom, 3) array>`, indexed there as [worldid % shape[0]]) would be
    # silently ignored.  Edit in place (`m.geom_friction[0, g] = ...`, `m.dof_damping.copy_(x)`) instead.
    if self.__dict__.get("_frozen") and (name in MODEL_INT_FIELDS or name in MODEL_REAL_FIELDS):
      raise NotImplementedError(
        f"Model.{name} is a view into the packed device model: assign in place (m.{name}[...] = value / m.{name}.copy_(value)). "
        "Per-world (nworld-batched) model fields are not implemented."
      )
    object.__setattr__(self, name, value)


# Bumped whenever a tensor attribute of any Data / Contact / Constraint object is (re)assigned; lets the
# step path detect a replaced tensor (`d.ctrl = new`) without walking all ~90 fields on every call.
_EPOCH = [0]


class _TracksAssignment:
  def __setattr__(self, name, value):
    if not name.startswith("_"):  # private bookkeeping (handles, bind epoch) is not device data
      _EPOCH[0] += 1
    object.__setattr__(self, name, value)


@dataclasses.dataclass
class Contact(_TracksAssignment):
  """Pooled contact arrays (reference types.py:1618-1655); row i belongs to world worldid[i]."""

  dist: Any = None
  pos: Any = None
  frame: Any = None
  includemargin: Any = None
  friction: Any = None
  solref: Any = None
  solreffriction: Any = None
  solimp: Any = None
  dim: Any = None
  geom: Any = None
  flex: Any = None
  vert: Any = None
  efc_address: Any = None
  worldid: Any = None
  type: Any = None
  geomcollisionid: Any = None


@dataclasses.dataclass
class Constraint(_TracksAssignment):
  """Per-world constraint rows (reference types.py:1659-1700), dense Jacobian."""

  type: Any = None
  id: Any = None
  J_rownnz: Any = None
  J_rowadr: Any = None
  J_colind: Any = None
  J: Any = None
  pos: Any = None
  margin: Any = None
  D: Any = None
  vel: Any = None
  aref: Any = None
  frictionloss: Any = None
  force: Any = None
  state: Any = None
  Ma: Any = None


class Data(_TracksAssignment):
  """nworld-batched simulation state and outputs (reference types.py:1703-1896)."""

  def __init__(self):
    self.contact = Contact()
    self.efc = Constraint()
    self.nworld = 0
    self.naconmax = 0
    self.naccdmax = 0
    self.njmax = 0
    self.njmax_pad = 0
    self.njmax_nnz = 0
    self._handle = None
    self._model = None
    self._sizes = {}

  def field(self, name: str):
    """Returns the tensor of an ABI field name (`efc_J` -> self.efc.J)."""
    if name.startswith("contact_"):
      return getattr(self.contact, name[len("contact_") :])
    if name.startswith("efc_"):
      return getattr(self.efc, name[len("efc_") :])
    return getattr(self, name)

  def set_field(self, name: str, value):
    if name.startswith("contact_"):
      setattr(self.contact, name[len("contact_") :], value)
    elif name.startswith("efc_"):
      setattr(self.efc, name[len("efc_") :], value)
    else:
      setattr(self, name, value)
