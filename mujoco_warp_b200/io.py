"""Host <-> device boundary: put_model / make_data / put_data / get_data_into / reset_data.

Same names, arguments and error behaviour as the reference (reference
mujoco_warp/_src/io.py:77 put_model, :859 make_data, :1016 put_data, :1243 get_data_into,
:1458 reset_data).  `mjm` is any object exposing MuJoCo's `MjModel` attribute names: a real
`mujoco.MjModel` if the package is importable, or `mjcf.MjModelLite` from this package.
"""

from __future__ import annotations

import ctypes
import dataclasses
import warnings
from types import SimpleNamespace
from typing import Optional

import numpy as np
import torch

from . import _lib
from . import constants as C
from . import types

_NV_MAX_DENSE = 60  # reference io.py:142
_NV_MAX = 128  # this library: dense storage for every model, the per-world Hessian and its factor live in shared memory


# ----------------------------------------------------------------------------------------------
# host model tables (numpy, float64 / int32) -- shared by put_model and by the test oracle wrapper
# ----------------------------------------------------------------------------------------------
def is_sparse(mjm) -> bool:
  """reference io.py:67-74."""
  jac = int(mjm.opt.jacobian)
  if jac == C.JacobianType.AUTO:
    return mjm.nv > 32
  return jac == C.JacobianType.SPARSE


def _get_padded_sizes(nv: int, njmax: int, sparse: bool, tile_size: int):
  """reference io.py:650-661."""

  def round_up(x, multiple):
    return ((x + multiple - 1) // multiple) * multiple

  njmax_padded = round_up(njmax, tile_size)
  nv_padded = round_up(nv, tile_size) if (sparse or nv > 32) else round_up(nv, 4)
  return njmax_padded, nv_padded


def _default_nconmax(mjm, mjd=None) -> int:
  """reference io.py:664-674."""
  valid_sizes = (2 + (np.arange(19) % 2)) * (2 ** (np.arange(19) // 2 + 3))
  nhfield = getattr(mjm, "nhfield", 0)
  nconmax = max(mjm.nv * 0.35 * (nhfield > 0) * 10 + 45, 256 * (getattr(mjm, "nflex", 0) > 0), getattr(mjd, "ncon", 0) if mjd else 0)
  return int(valid_sizes[np.searchsorted(valid_sizes, nconmax)])


def _default_njmax(mjm, mjd=None) -> int:
  """reference io.py:677-687."""
  valid_sizes = (2 + (np.arange(19) % 2)) * (2 ** (np.arange(19) // 2 + 3))
  nhfield = getattr(mjm, "nhfield", 0)
  njmax = max(mjm.nv * 2.26 * (nhfield > 0) * 18 + 53, 512 * (getattr(mjm, "nflex", 0) > 0), getattr(mjd, "nefc", 0) if mjd else 0)
  return int(valid_sizes[np.searchsorted(valid_sizes, njmax)])


_GROUP_BROADPHASE_MIN_PAIRS = 1024  # filtered geom pairs from which the two-level (group sphere, then geom pair) broadphase is used
_BLOB_SMEM_LIMIT = 96 * 1024  # bytes of model tables the stage-group kernels keep in shared memory (mjb_api.cu)


def _check_supported(mjm):
  """Feature gating with the reference's exception types (reference io.py:90-144)."""
  for field, enum_cls in (
    (mjm.actuator_trntype, C.TrnType),
    (mjm.actuator_dyntype, C.DynType),
    (mjm.actuator_gaintype, C.GainType),
    (mjm.actuator_biastype, C.BiasType),
    (mjm.eq_type, C.EqType),
    (mjm.geom_type, C.GeomType),
  ):
    valid = [int(v) for v in enum_cls]
    missing = ~np.isin(np.asarray(field), valid)
    if missing.any():
      raise NotImplementedError(f"{np.asarray(field)[missing].tolist()} not supported.")
  for value, enum_cls in ((mjm.opt.integrator, C.IntegratorType), (mjm.opt.cone, C.ConeType), (mjm.opt.solver, C.SolverType)):
    if int(value) not in [int(v) for v in enum_cls]:
      raise NotImplementedError(f"{value} is unsupported.")
  if int(mjm.opt.solver) == C.SolverType.PGS:
    raise NotImplementedError("PGS is unsupported.")
  if int(mjm.opt.integrator) == C.IntegratorType.IMPLICIT:
    raise NotImplementedError("IMPLICIT is unsupported.")
  if getattr(mjm.opt, "noslip_iterations", 0) > 0:
    raise NotImplementedError("noslip solver not implemented.")
  if mjm.nv > _NV_MAX_DENSE and int(mjm.opt.jacobian) == C.JacobianType.DENSE:
    raise ValueError(f"Dense is unsupported for nv > {_NV_MAX_DENSE} (nv = {mjm.nv}).")
  # features of the reference that this library does not implement yet (scope)
  # Models the reference stores sparse (nv > 32 under jacobian=auto, or jacobian=sparse) are accepted and stored DENSE here: qM, qLD
  # and efc.J keep their dense shapes (Model.is_sparse is False), every other field is identical.  One warp per world with the
  # Hessian in shared memory bounds nv.
  if mjm.nv > _NV_MAX:
    raise NotImplementedError(f"nv > {_NV_MAX} is not implemented (nv = {mjm.nv}).")
  for name in ("ntendon", "nflex"):
    if getattr(mjm, name, 0):
      raise NotImplementedError(f"{name} > 0 is not implemented yet.")
  for t in np.asarray(mjm.actuator_trntype):
    if int(t) not in (C.TrnType.JOINT, C.TrnType.JOINTINPARENT):
      raise NotImplementedError("only joint / jointinparent actuator transmissions are implemented.")
  for i, t in enumerate(np.asarray(mjm.eq_type)):
    if int(t) not in (C.EqType.JOINT, C.EqType.CONNECT, C.EqType.WELD):
      raise NotImplementedError("only connect / weld / joint equality constraints are implemented.")
    if int(t) != C.EqType.JOINT and int(np.asarray(getattr(mjm, "eq_objtype", np.ones(mjm.neq)))[i]) not in (1, 6):
      raise NotImplementedError("connect / weld equality constraints are implemented between bodies or sites only.")
  for arr, bad in ((mjm.actuator_dyntype, (C.DynType.MUSCLE, C.DynType.USER)), (mjm.actuator_gaintype, (C.GainType.MUSCLE, C.GainType.USER)), (mjm.actuator_biastype, (C.BiasType.MUSCLE, C.BiasType.USER))):
    if np.isin(np.asarray(arr), [int(b) for b in bad]).any():
      raise NotImplementedError("muscle / user actuators are not implemented.")
  if float(mjm.opt.density) > 0 or float(mjm.opt.viscosity) > 0 or np.asarray(mjm.opt.wind).any():
    raise NotImplementedError("fluid forces are not implemented yet.")


def host_model(mjm) -> SimpleNamespace:
  """Builds every model table the kernels need (numpy float64 / int32).

  Follows reference io.py:212-647 for the derived fields (nv_pad, body_tree, limited joint
  lists, the statically filtered NXN geom pairs, ...).
  """
  _check_supported(mjm)
  i32, f64 = np.int32, np.float64
  nbody, njnt, ngeom, nv, nu = int(mjm.nbody), int(mjm.njnt), int(mjm.ngeom), int(mjm.nv), int(mjm.nu)
  ints, reals = {}, {}

  def A(name, dtype):
    return np.ascontiguousarray(np.asarray(getattr(mjm, name)), dtype=dtype)

  for name in (
    "body_parentid body_rootid body_weldid body_mocapid body_jntnum body_jntadr body_dofnum body_dofadr "
    "jnt_type jnt_qposadr jnt_dofadr jnt_bodyid jnt_actfrclimited jnt_actgravcomp dof_bodyid dof_jntid dof_parentid "
    "geom_type geom_bodyid geom_condim geom_priority site_bodyid pair_dim actuator_trntype actuator_trnid "
    "actuator_dyntype actuator_gaintype actuator_biastype actuator_actadr actuator_actnum actuator_ctrllimited "
    "actuator_forcelimited actuator_actlimited actuator_actearly eq_type eq_obj1id eq_obj2id"
  ).split():
    ints[name] = A(name, i32)
  ints["eq_active0"] = np.asarray(getattr(mjm, "eq_active0", np.ones(int(getattr(mjm, "neq", 0)), dtype=bool))).astype(i32).reshape(-1)
  # mesh tables (absent on models without mesh assets)
  nmesh = int(getattr(mjm, "nmesh", 0))
  ints["geom_dataid"] = A("geom_dataid", i32) if hasattr(mjm, "geom_dataid") else np.full(ngeom, -1, dtype=i32)
  ints["mesh_vertadr"] = A("mesh_vertadr", i32) if nmesh else np.zeros(0, dtype=i32)
  ints["mesh_vertnum"] = A("mesh_vertnum", i32) if nmesh else np.zeros(0, dtype=i32)
  ints["mesh_graphadr"] = A("mesh_graphadr", i32) if nmesh and hasattr(mjm, "mesh_graphadr") else np.full(nmesh, -1, dtype=i32)
  ints["mesh_graph"] = A("mesh_graph", i32).reshape(-1) if nmesh and hasattr(mjm, "mesh_graph") else np.zeros(0, dtype=i32)
  for name in ("hfield_nrow", "hfield_ncol", "hfield_adr"):
    ints[name] = A(name, i32) if int(getattr(mjm, "nhfield", 0)) else np.zeros(0, dtype=i32)
  for name in types.MODEL_REAL_FIELDS:
    if name == "mesh_vert":
      reals[name] = A(name, f64).reshape(-1, 3) if nmesh else np.zeros((0, 3))
    elif name == "actuator_gear":
      reals[name] = A(name, f64).reshape(nu, 6)
    elif name == "group_sphere":
      continue  # derived below (two-level broadphase)
    elif name in ("hfield_size", "hfield_data"):
      nhf = int(getattr(mjm, "nhfield", 0))
      reals[name] = (A(name, f64).reshape(-1, 4) if name == "hfield_size" else A(name, f64).reshape(-1)) if nhf else np.zeros((0, 4) if name == "hfield_size" else 0)
    else:
      reals[name] = A(name, f64)

  parent = ints["body_parentid"]
  # subtree ranges: MuJoCo numbers bodies depth-first, so a subtree is the id range [b, end)
  end = np.arange(1, nbody + 1, dtype=i32)
  for b in range(nbody - 1, 0, -1):
    end[parent[b]] = max(end[parent[b]], end[b])
  for b in range(1, nbody):
    if not (parent[b] < b and end[parent[b]] >= end[b]):
      raise ValueError("bodies are not in depth-first order")
  # every body in [b, end[b]) must descend from b
  for b in range(1, nbody):
    for c in range(b + 1, end[b]):
      p = c
      while p > b:
        p = parent[p]
      if p != b:
        raise ValueError("bodies are not in depth-first (pre-order) order")
  ints["body_subtreeend"] = end

  # body ids grouped by tree depth (reference io.py:235-239)
  depth = np.zeros(nbody, dtype=int) - 1
  levels = {}
  for b in range(nbody):
    depth[b] = depth[parent[b]] + 1 if b > 0 else 0
    levels.setdefault(int(depth[b]), []).append(b)
  tree, tree_adr = [], [0]
  for lv in sorted(levels):
    tree += levels[lv]
    tree_adr.append(len(tree))
  ints["body_tree"] = np.array(tree, dtype=i32)
  ints["body_tree_adr"] = np.array(tree_adr, dtype=i32)
  nlevel = len(levels)

  jnt_type = ints["jnt_type"]
  limited = np.asarray(mjm.jnt_limited).astype(bool)
  ints["jnt_limited_adr"] = np.nonzero(limited & np.isin(jnt_type, (C.JointType.SLIDE, C.JointType.HINGE)))[0].astype(i32)
  ints["jnt_limited_ball_adr"] = np.nonzero(limited & (jnt_type == C.JointType.BALL))[0].astype(i32)
  ints["dof_friction_adr"] = np.nonzero(reals["dof_frictionloss"] > 0)[0].astype(i32)
  tree0 = np.arange(nv, dtype=i32)  # first dof of the dof's kinematic tree (parents precede children)
  for i in range(nv):
    p = int(ints["dof_parentid"][i])
    if p >= 0:
      tree0[i] = tree0[p]
  ints["dof_tree0"] = tree0
  ints["eq_objtype"] = np.asarray(getattr(mjm, "eq_objtype", np.ones(int(mjm.neq)))).astype(i32).reshape(-1)  # mjtObj: 1 body, 3 joint, 6 site
  ints["eq_jnt_adr"] = np.nonzero(ints["eq_type"] == C.EqType.JOINT)[0].astype(i32)
  ints["eq_connect_adr"] = np.nonzero(ints["eq_type"] == C.EqType.CONNECT)[0].astype(i32)  # reference io.py:445-446
  ints["eq_wld_adr"] = np.nonzero(ints["eq_type"] == C.EqType.WELD)[0].astype(i32)

  # static geoms: attached to the world and not descended from a mocap body (smooth.py:195)
  bodyid = ints["geom_bodyid"]
  ints["geom_static"] = ((ints["body_weldid"][bodyid] == 0) & (ints["body_mocapid"][ints["body_rootid"][bodyid]] == -1)).astype(i32)

  # statically filtered geom pairs (reference io.py:270-358)
  filterparent = not (int(mjm.opt.disableflags) & C.DisableBit.FILTERPARENT)
  geom1, geom2 = np.triu_indices(ngeom, k=1)
  bodyid1, bodyid2 = bodyid[geom1], bodyid[geom2]
  contype, conaff = A("geom_contype", i32), A("geom_conaffinity", i32)
  weld = ints["body_weldid"]
  weldid1, weldid2 = weld[bodyid1], weld[bodyid2]
  weld_parentid1, weld_parentid2 = weld[parent[weldid1]], weld[parent[weldid2]]
  self_collision = weldid1 == weldid2
  parent_child = filterparent & (weldid1 != 0) & (weldid2 != 0) & ((weldid1 == weld_parentid2) | (weldid2 == weld_parentid1))
  mask = ((contype[geom1] & conaff[geom2]) | (contype[geom2] & conaff[geom1])).astype(bool)
  exclude = np.isin((bodyid1.astype(np.int64) << 16) + bodyid2, np.asarray(getattr(mjm, "exclude_signature", []), dtype=np.int64))
  pairid_contact = -1 * np.ones(len(geom1), dtype=int)
  pairid_contact[~(mask & ~self_collision & ~parent_child & ~exclude)] = -2

  def upper_tri_index(n, i, j):
    i, j = (j, i) if j < i else (i, j)
    return (i * (2 * n - i - 3)) // 2 + j - 1

  npair = int(getattr(mjm, "npair", 0))
  for i in range(npair):
    pairid_contact[upper_tri_index(ngeom, int(mjm.pair_geom1[i]), int(mjm.pair_geom2[i]))] = i
  pairid_collision = -1 * np.ones(len(geom1), dtype=int)  # collision sensors: not supported (nsensor == 0)
  include = pairid_contact > -2
  ints["nxn_pair"] = np.stack((geom1, geom2), axis=1)[include].astype(i32).reshape(-1, 2)
  ints["nxn_pairid"] = np.stack((pairid_contact, pairid_collision), axis=1)[include].astype(i32).reshape(-1, 2)
  nxn = int(include.sum())
  lookup = -1 * np.ones(len(geom1), dtype=i32)  # pair table position -> filtered pair index (sweep-and-prune path)
  lookup[include] = np.arange(nxn, dtype=i32)
  ints["nxn_lookup"] = lookup

  # box-box runs GJK / EPA + multi-contact unless DisableBit.NATIVECCD; multi-contact does not support margins (reference io.py:372-403)
  if not (int(mjm.opt.disableflags) & C.DisableBit.NATIVECCD):
    gt_, gm_ = ints["geom_type"], reals["geom_margin"]
    for (ga, gb), (pid, _) in zip(ints["nxn_pair"], ints["nxn_pairid"]):
      if int(gt_[ga]) == C.GeomType.BOX and int(gt_[gb]) == C.GeomType.BOX:
        margin = float(reals["pair_margin"][pid]) if pid >= 0 else max(float(gm_[ga]), float(gm_[gb]))
        if margin:
          raise NotImplementedError(f"geom pair ({ga}, {gb}) has non-zero margin with NATIVECCD enabled. Set margin to 0 or disable NATIVECCD.")

  # Two-level broadphase for large pair lists (aloha_pot: 9154 filtered pairs, of which a handful are ever close): geoms are grouped
  # (all geoms of a moving body; every world geom alone), each group gets a bounding sphere in its body's frame that contains the
  # members' bounding spheres with their margins, and the pair list is sorted by group pair.  The collision kernel tests the
  # group spheres first and walks the geom pairs of the surviving group pairs only.  The test is conservative with respect to
  # the reference's per-pair sphere filter (collision_driver.py:91-111), so the set of pairs that pass -- and ncollision -- is
  # unchanged; only the order of the pair list (hence of a world's contacts, arbitrary in the reference) differs.
  ngroup = ngrouppair = 0
  grp_body, grp_sphere = np.zeros(0, dtype=i32), np.zeros((0, 4))
  grp_pair, grp_adr, grp_always = np.zeros((0, 2), dtype=i32), np.zeros(1, dtype=i32), np.zeros(0, dtype=i32)
  if nxn >= _GROUP_BROADPHASE_MIN_PAIRS:
    geom_group = np.zeros(ngeom, dtype=np.int64)
    members = []
    by_body = {}
    for g in range(ngeom):
      b = int(bodyid[g])
      if b == 0:
        geom_group[g] = len(members)
        members.append([g])
      else:
        if b not in by_body:
          by_body[b] = len(members)
          members.append([])
        geom_group[g] = by_body[b]
        members[by_body[b]].append(g)
    ngroup = len(members)
    grp_body = np.array([int(bodyid[mm[0]]) for mm in members], dtype=i32)
    grp_sphere = np.zeros((ngroup, 4))
    has_plane = np.zeros(ngroup, dtype=bool)
    gpos, grb, gmargin = reals["geom_pos"], reals["geom_rbound"], reals["geom_margin"]
    for k, mm in enumerate(members):
      c = gpos[mm].mean(axis=0)
      r = max(float(np.linalg.norm(gpos[g] - c)) + float(grb[g]) + float(gmargin[g]) for g in mm)
      grp_sphere[k] = [c[0], c[1], c[2], r * (1.0 + 1e-4) + 1e-6]  # slack: the device evaluates both tests in float32
      has_plane[k] = any(float(grb[g]) == 0.0 for g in mm)
    pr = ints["nxn_pair"]
    ga, gb = geom_group[pr[:, 0]], geom_group[pr[:, 1]]
    lo, hi = np.minimum(ga, gb), np.maximum(ga, gb)
    order = np.lexsort((np.arange(nxn), hi, lo))
    ints["nxn_pair"] = np.ascontiguousarray(pr[order])
    ints["nxn_pairid"] = np.ascontiguousarray(ints["nxn_pairid"][order])
    inv = np.empty(nxn, dtype=i32)
    inv[order] = np.arange(nxn, dtype=i32)
    lookup[include] = inv
    ints["nxn_lookup"] = lookup
    key = (lo[order] * ngroup + hi[order]).astype(np.int64)
    starts = np.concatenate(([0], np.nonzero(np.diff(key))[0] + 1)).astype(i32)
    ngrouppair = int(starts.size)
    grp_adr = np.concatenate((starts, [nxn])).astype(i32)
    grp_pair = np.stack((lo[order][starts], hi[order][starts]), axis=1).astype(i32)
    sensor = ints["nxn_pairid"][:, 1] >= 0
    grp_always = np.array(
      [int(has_plane[a] or has_plane[b] or sensor[grp_adr[k] : grp_adr[k + 1]].any()) for k, (a, b) in enumerate(grp_pair)], dtype=i32)
  ints["group_bodyid"], ints["grouppair"], ints["grouppair_adr"], ints["grouppair_always"] = grp_body, grp_pair, grp_adr, grp_always
  reals["group_sphere"] = grp_sphere

  # supported narrowphase pair types: analytic primitives + GJK / EPA for ellipsoid / cylinder / mesh pairs (no hfields yet)
  supported = {
    (C.GeomType.PLANE, C.GeomType.SPHERE),
    (C.GeomType.PLANE, C.GeomType.CAPSULE),
    (C.GeomType.PLANE, C.GeomType.BOX),
    (C.GeomType.SPHERE, C.GeomType.SPHERE),
    (C.GeomType.SPHERE, C.GeomType.CAPSULE),
    (C.GeomType.CAPSULE, C.GeomType.CAPSULE),
    (C.GeomType.SPHERE, C.GeomType.BOX),
    (C.GeomType.CAPSULE, C.GeomType.BOX),
    (C.GeomType.BOX, C.GeomType.BOX),
    (C.GeomType.PLANE, C.GeomType.ELLIPSOID),
    (C.GeomType.PLANE, C.GeomType.CYLINDER),
    (C.GeomType.SPHERE, C.GeomType.CYLINDER),
    # convex narrowphase (GJK / EPA, reference collision_driver.py:43 MJ_COLLISION_TABLE entries of type CONVEX)
    (C.GeomType.SPHERE, C.GeomType.ELLIPSOID),
    (C.GeomType.CAPSULE, C.GeomType.ELLIPSOID),
    (C.GeomType.CAPSULE, C.GeomType.CYLINDER),
    (C.GeomType.ELLIPSOID, C.GeomType.ELLIPSOID),
    (C.GeomType.ELLIPSOID, C.GeomType.CYLINDER),
    (C.GeomType.ELLIPSOID, C.GeomType.BOX),
    (C.GeomType.CYLINDER, C.GeomType.CYLINDER),
    (C.GeomType.CYLINDER, C.GeomType.BOX),
    # convex-hull meshes: plane_convex (exhaustive vertex search) and GJK / EPA with the exhaustive support function
    (C.GeomType.PLANE, C.GeomType.MESH),
    (C.GeomType.SPHERE, C.GeomType.MESH),
    (C.GeomType.CAPSULE, C.GeomType.MESH),
    (C.GeomType.ELLIPSOID, C.GeomType.MESH),
    (C.GeomType.CYLINDER, C.GeomType.MESH),
    (C.GeomType.BOX, C.GeomType.MESH),
    (C.GeomType.MESH, C.GeomType.MESH),
    # height fields: one GJK / EPA run per prism of the sub-grid under the geom (collision_convex.py:158)
    (C.GeomType.HFIELD, C.GeomType.SPHERE),
    (C.GeomType.HFIELD, C.GeomType.CAPSULE),
    (C.GeomType.HFIELD, C.GeomType.ELLIPSOID),
    (C.GeomType.HFIELD, C.GeomType.CYLINDER),
    (C.GeomType.HFIELD, C.GeomType.BOX),
    (C.GeomType.HFIELD, C.GeomType.MESH),
  }
  gtype = ints["geom_type"]
  for ga, gb in ints["nxn_pair"]:
    t = tuple(sorted((int(gtype[ga]), int(gtype[gb]))))
    if t not in supported:
      raise NotImplementedError(f"collision between geom types {C.GeomType(t[0]).name} and {C.GeomType(t[1]).name} is not implemented yet.")

  # actuator moment layout is static for joint transmissions (reference assigns it per step by
  # atomics, smooth.py:2094/2132/2144)
  nJmom = 0
  for a in range(nu):
    jt = int(jnt_type[ints["actuator_trnid"][a, 0]])
    nJmom += 6 if jt == C.JointType.FREE else (3 if jt == C.JointType.BALL else 1)

  sparse = False  # storage is dense for every model (see _check_supported); padded sizes do not depend on it (both tile sizes are 16)
  nv_pad = _get_padded_sizes(nv, 0, is_sparse(mjm), types.TILE_SIZE_JTDAJ_SPARSE if sparse else types.TILE_SIZE_JTDAJ_DENSE)[1]
  nmaxcondim = int(np.concatenate(([0], ints["geom_condim"], ints["pair_dim"])).max())
  sizes = dict(
    nq=int(mjm.nq), nv=nv, nu=nu, na=int(mjm.na), nbody=nbody, njnt=njnt, ngeom=ngeom, nsite=int(mjm.nsite),
    nmocap=int(mjm.nmocap), neq=int(mjm.neq), npair=npair, nxn=nxn, nlevel=nlevel, nv_pad=nv_pad,
    nlimit=int(ints["jnt_limited_adr"].size), nlimitball=int(ints["jnt_limited_ball_adr"].size),
    nfrictiondof=int(ints["dof_friction_adr"].size), neqjnt=int(ints["eq_jnt_adr"].size), neqconnect=int(ints["eq_connect_adr"].size),
    neqweld=int(ints["eq_wld_adr"].size), nJmom=nJmom,
    nmaxcondim=nmaxcondim, nmaxpyramid=int(max(1, 2 * (nmaxcondim - 1))),
    ngravcomp=int((reals["body_gravcomp"] != 0).sum()), is_sparse=int(sparse), nM=int(getattr(mjm, "nM", 0)),
    ntree=int(getattr(mjm, "ntree", 0)), ngeompair=int(len(geom1)), nmesh=nmesh, nmeshvert=int(reals["mesh_vert"].shape[0]), nmeshgraph=int(ints["mesh_graph"].size), blob_hot_bytes=0,
    ngroup=ngroup, ngrouppair=ngrouppair, nhfield=int(getattr(mjm, "nhfield", 0)), nhfielddata=int(reals["hfield_data"].size),
  )  # fmt: skip
  extra = dict(
    nsensordata=int(getattr(mjm, "nsensordata", 0)), ncam=int(getattr(mjm, "ncam", 0)), nlight=int(getattr(mjm, "nlight", 0)),
    ntendon=0, nJten=0,
  )  # fmt: skip

  # shape check against the tables
  for name, shape in types.MODEL_INT_FIELDS.items():
    want = types.resolve_shape(shape, sizes)
    ints[name] = np.ascontiguousarray(ints[name].reshape(want), dtype=i32)
  for name, shape in types.MODEL_REAL_FIELDS.items():
    want = types.resolve_shape(shape, sizes)
    reals[name] = np.ascontiguousarray(reals[name].reshape(want), dtype=f64)

  o = mjm.opt
  opt = dict(
    timestep=float(o.timestep), tolerance=max(float(o.tolerance), 1e-6),  # reference io.py:185
    ls_tolerance=float(o.ls_tolerance), ccd_tolerance=float(getattr(o, "ccd_tolerance", 1e-6)),
    gravity=np.asarray(o.gravity, dtype=f64).copy(), wind=np.asarray(o.wind, dtype=f64).copy(),
    magnetic=np.asarray(o.magnetic, dtype=f64).copy(), density=float(o.density), viscosity=float(o.viscosity),
    integrator=int(o.integrator), cone=int(o.cone), solver=int(o.solver), iterations=int(o.iterations),
    ls_iterations=int(o.ls_iterations), ccd_iterations=int(getattr(o, "ccd_iterations", 35)),
    disableflags=int(o.disableflags), enableflags=int(o.enableflags),
    sdf_initpoints=int(getattr(o, "sdf_initpoints", 40)), sdf_iterations=int(getattr(o, "sdf_iterations", 10)),
    impratio_invsqrt=1.0 / np.sqrt(max(float(o.impratio), C.MJ_MINVAL)),  # reference io.py:179
    ls_parallel=False, ls_parallel_min_step=1.0e-6, broadphase=int(C.BroadphaseType.NXN),
    broadphase_filter=int(C.BroadphaseFilter.PLANE | C.BroadphaseFilter.SPHERE | C.BroadphaseFilter.OBB),  # io.py:192
    graph_conditional=True, run_collision_detection=True, contact_sensor_maxmatch=64,
  )  # fmt: skip
  return SimpleNamespace(sizes=sizes, extra=extra, ints=ints, reals=reals, opt=opt, meaninertia=float(mjm.stat.meaninertia))


def _static_frames(mjm, hm):
  """World-body and static-geom frames at qpos0 (reference io.py:958-983 uses mj_kinematics)."""
  from . import mjcf

  k = mjcf.host_kinematics(mjm, np.asarray(mjm.qpos0, dtype=np.float64))
  ngeom = hm.sizes["ngeom"]
  gx, gm = np.zeros((ngeom, 3)), np.zeros((ngeom, 3, 3))
  for g in range(ngeom):
    b = int(hm.ints["geom_bodyid"][g])
    gx[g] = k.xpos[b] + k.xmat[b] @ hm.reals["geom_pos"][g]
    gm[g] = mjcf.quat_to_mat(mjcf.quat_mul(k.xquat[b], hm.reals["geom_quat"][g]))
  return k, gx, gm


# ----------------------------------------------------------------------------------------------
# put_model
# ----------------------------------------------------------------------------------------------
def _device(device):
  if device is None:
    if not torch.cuda.is_available():
      raise RuntimeError("mujoco_warp_b200 needs a CUDA device (there is no CPU fallback); pass device='cpu' only to inspect host-side structures.")
    return torch.device("cuda", torch.cuda.current_device())
  return torch.device(device)


def put_model(mjm, device=None) -> types.Model:
  """Creates a model on device (reference io.py:77).

  All model arrays are packed back to back (16-byte aligned) into one device blob that the
  kernels stage into shared memory with a single TMA bulk copy; `Model` attributes are views
  into that blob.
  """
  dev = _device(device)
  hm = host_model(mjm)
  m = types.Model()
  m.sizes = dict(hm.sizes)
  m.sizes.update(hm.extra)
  m._host = hm

  o = hm.opt
  t = lambda v: torch.tensor(np.atleast_1d(np.asarray(v, dtype=np.float32)))  # noqa: E731
  m.opt = types.Option(
    timestep=t(o["timestep"]), tolerance=t(o["tolerance"]), ls_tolerance=t(o["ls_tolerance"]),
    ccd_tolerance=t(o["ccd_tolerance"]), gravity=t(o["gravity"]).reshape(1, 3), wind=t(o["wind"]).reshape(1, 3),
    magnetic=t(o["magnetic"]).reshape(1, 3), density=t(o["density"]), viscosity=t(o["viscosity"]),
    integrator=o["integrator"], cone=o["cone"], solver=o["solver"], iterations=o["iterations"],
    ls_iterations=o["ls_iterations"], ccd_iterations=o["ccd_iterations"], disableflags=o["disableflags"],
    enableflags=o["enableflags"], sdf_initpoints=o["sdf_initpoints"], sdf_iterations=o["sdf_iterations"],
    impratio_invsqrt=t(o["impratio_invsqrt"]), ls_parallel=o["ls_parallel"], ls_parallel_min_step=o["ls_parallel_min_step"],
    broadphase=o["broadphase"], broadphase_filter=o["broadphase_filter"], graph_conditional=o["graph_conditional"],
    run_collision_detection=o["run_collision_detection"], contact_sensor_maxmatch=o["contact_sensor_maxmatch"],
  )  # fmt: skip
  m.stat = types.Statistic(meaninertia=t(hm.meaninertia))

  # pack the blob
  int_names, real_names = list(types.MODEL_INT_FIELDS), list(types.MODEL_REAL_FIELDS)
  # hot tables first, then the large / streamed ones (pair lists, mesh vertices and graphs): when the whole blob does not fit the
  # kernels' shared-memory budget they stage the hot prefix only and read the cold tables from global memory (mjb_types.cuh MIC / MRC)
  cold = ("nxn_pair", "nxn_pairid", "nxn_lookup", "mesh_graph", "mesh_vert", "hfield_data")
  offsets, off, hot_bytes = {}, 0, 0
  for name in [n for n in int_names + real_names if n not in cold] + list(cold):
    arr = hm.ints[name] if name in hm.ints else hm.reals[name]
    if name == cold[0]:
      hot_bytes = off
    offsets[name] = off
    off += (arr.size * 4 + 15) // 16 * 16
  nbytes = max(off, 16)
  m.sizes["blob_hot_bytes"] = hot_bytes if nbytes > _BLOB_SMEM_LIMIT else 0  # 0: stage everything
  host = np.zeros(nbytes, dtype=np.uint8)
  for name in int_names:
    a = hm.ints[name].astype(np.int32).ravel()
    host[offsets[name] : offsets[name] + a.size * 4] = a.view(np.uint8)
  for name in real_names:
    a = hm.reals[name].astype(np.float32).ravel()
    host[offsets[name] : offsets[name] + a.size * 4] = a.view(np.uint8)
  m._blob = torch.from_numpy(host).to(dev)
  m._offsets = offsets
  for name in int_names:
    shape = hm.ints[name].shape
    n = int(np.prod(shape))
    setattr(m, name, m._blob[offsets[name] : offsets[name] + 4 * n].view(torch.int32).view(shape))
  for name in real_names:
    shape = hm.reals[name].shape
    n = int(np.prod(shape))
    setattr(m, name, m._blob[offsets[name] : offsets[name] + 4 * n].view(torch.float32).view((1,) + tuple(shape)))
  # fields of the reference Model that are host-side scalars / flags here
  m.is_sparse = bool(hm.sizes["is_sparse"])
  m.has_fluid = False
  m.has_sdf_geom = False
  m.device = dev

  if dev.type == "cuda":
    lib = _lib.load()
    m._handle = _lib.model_create(lib, m, int_names, real_names)
  m._frozen = True  # blob fields can no longer be re-assigned (types.Model.__setattr__)
  return m


# ----------------------------------------------------------------------------------------------
# make_data / put_data
# ----------------------------------------------------------------------------------------------
def _resolve_batch_size(na, n, nworld, default):
  if na is not None:
    return na
  if n is not None:
    return n * nworld
  return default


def _data_sizes(mjm, hm, nworld, nconmax, nccdmax, njmax, njmax_nnz, naconmax, naccdmax):
  """Argument validation and size resolution of reference io.py:887-933."""
  if nconmax is None:
    nconmax = _default_nconmax(mjm)
  if njmax is None:
    njmax = _default_njmax(mjm)
  if nconmax < 0:
    raise ValueError("nconmax must be >= 0")
  if njmax < 0:
    raise ValueError("njmax must be >= 0")
  if nworld < 1:
    raise ValueError("nworld must be >= 1")
  naconmax = _resolve_batch_size(naconmax, nconmax, nworld, 0)
  if naconmax < 0:
    raise ValueError("naconmax must be >= 0")
  naccdmax = _resolve_batch_size(naccdmax, nccdmax, nworld, naconmax)
  if naccdmax < 0:
    raise ValueError("naccdmax must be >= 0")
  elif naccdmax > naconmax:
    raise ValueError(f"naccdmax ({naccdmax}) must be <= naconmax ({naconmax})")
  if nccdmax is not None:
    if nccdmax < 0:
      raise ValueError("nccdmax must be >= 0")
    elif nccdmax > nconmax:
      raise ValueError(f"nccdmax ({nccdmax}) must be <= nconmax ({nconmax})")
  sparse = bool(hm.sizes["is_sparse"])
  tile = types.TILE_SIZE_JTDAJ_SPARSE if sparse else types.TILE_SIZE_JTDAJ_DENSE
  njmax_pad, _ = _get_padded_sizes(hm.sizes["nv"], njmax, sparse, tile)
  if njmax_nnz is None:
    njmax_nnz = njmax * hm.sizes["nv"]
  # per-world contact index list: a world may hold more than nconmax contacts (reference
  # io.py:874) as long as the pool has room
  nwconmax = int(min(max(naconmax, 1), max(4 * max(nconmax, 1), 64)))
  sizes = dict(hm.sizes)
  sizes.update(hm.extra)
  sizes.update(nworld=nworld, naconmax=naconmax, njmax=njmax, njmax_pad=njmax_pad, nwconmax=nwconmax)
  return sizes, naccdmax, njmax_nnz


def make_data(
  mjm,
  nworld: int = 1,
  nconmax: Optional[int] = None,
  nccdmax: Optional[int] = None,
  njmax: Optional[int] = None,
  njmax_nnz: Optional[int] = None,
  naconmax: Optional[int] = None,
  naccdmax: Optional[int] = None,
  device=None,
  m: Optional[types.Model] = None,
) -> types.Data:
  """Creates a data object on device (reference io.py:859).

  `m` (optional, not in the reference signature) binds the data to an existing device model;
  otherwise the binding happens lazily on the first stage call `f(m, d)`.
  """
  dev = _device(device) if m is None else m.device
  hm = m._host if m is not None else host_model(mjm)
  sizes, naccdmax, njmax_nnz = _data_sizes(mjm, hm, nworld, nconmax, nccdmax, njmax, njmax_nnz, naconmax, naccdmax)

  d = types.Data()
  d._sizes = sizes
  d.nworld, d.naconmax, d.naccdmax = nworld, sizes["naconmax"], naccdmax
  d.njmax, d.njmax_pad, d.njmax_nnz = sizes["njmax"], sizes["njmax_pad"], njmax_nnz
  d.device = dev
  for name, shape in types.DATA_INT_FIELDS.items():
    d.set_field(name, torch.zeros(types.resolve_shape(shape, sizes), dtype=torch.int32, device=dev))
  for name, shape in types.DATA_REAL_FIELDS.items():
    d.set_field(name, torch.zeros(types.resolve_shape(shape, sizes), dtype=torch.float32, device=dev))
  for name, (kind, shape) in types.DATA_EXTRA_FIELDS.items():
    dt = torch.float32 if kind == "f" else torch.int32
    setattr(d, name, torch.zeros(types.resolve_shape(shape, sizes), dtype=dt, device=dev))
  d.contact.efc_address.fill_(-1)
  d.contact.flex = torch.zeros((sizes["naconmax"], 2), dtype=torch.int32, device=dev)
  d.contact.vert = torch.zeros((sizes["naconmax"], 2), dtype=torch.int32, device=dev)
  d.efc.J_rownnz = torch.zeros((nworld, 0), dtype=torch.int32, device=dev)
  d.efc.J_rowadr = torch.zeros((nworld, 0), dtype=torch.int32, device=dev)
  d.efc.J_colind = torch.zeros((nworld, 0, 0), dtype=torch.int32, device=dev)

  f32 = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dev)  # noqa: E731
  d.qpos.copy_(f32(hm.reals["qpos0"]).unsqueeze(0).expand(nworld, -1))
  k, gx, gm = _static_frames(mjm, hm)
  d.xquat.copy_(f32(k.xquat).unsqueeze(0).expand(nworld, -1, -1))
  d.xmat.copy_(f32(k.xmat).unsqueeze(0).expand(nworld, -1, -1, -1))
  d.ximat.copy_(f32(k.ximat).unsqueeze(0).expand(nworld, -1, -1, -1))
  d.xpos.copy_(f32(k.xpos).unsqueeze(0).expand(nworld, -1, -1))
  d.xipos.copy_(f32(k.xipos).unsqueeze(0).expand(nworld, -1, -1))
  d.geom_xpos.copy_(f32(gx).unsqueeze(0).expand(nworld, -1, -1))
  d.geom_xmat.copy_(f32(gm).unsqueeze(0).expand(nworld, -1, -1, -1))
  nmocap = hm.sizes["nmocap"]
  if nmocap:
    mocap_body = np.nonzero(hm.ints["body_mocapid"] >= 0)[0]
    order = mocap_body[np.argsort(hm.ints["body_mocapid"][mocap_body])]
    d.mocap_pos.copy_(f32(hm.reals["body_pos"][order]).unsqueeze(0).expand(nworld, -1, -1))
    d.mocap_quat.copy_(f32(hm.reals["body_quat"][order]).unsqueeze(0).expand(nworld, -1, -1))
  if hm.sizes["neq"]:
    act0 = torch.from_numpy(np.asarray(mjm.eq_active0).astype(np.int32)).to(dev)
    d.eq_active.copy_(act0.unsqueeze(0).expand(nworld, -1))
  # static actuator moment layout
  if hm.sizes["nu"]:
    rownnz, rowadr, colind, adr = [], [], [], 0
    for a in range(hm.sizes["nu"]):
      j = int(hm.ints["actuator_trnid"][a, 0])
      jt = int(hm.ints["jnt_type"][j])
      n = 6 if jt == C.JointType.FREE else (3 if jt == C.JointType.BALL else 1)
      rownnz.append(n)
      rowadr.append(adr)
      colind += [int(hm.ints["jnt_dofadr"][j]) + i for i in range(n)]
      adr += n
    i32t = lambda a: torch.tensor(a, dtype=torch.int32, device=dev)  # noqa: E731
    d.moment_rownnz.copy_(i32t(rownnz).unsqueeze(0).expand(nworld, -1))
    d.moment_rowadr.copy_(i32t(rowadr).unsqueeze(0).expand(nworld, -1))
    d.moment_colind.copy_(i32t(colind).unsqueeze(0).expand(nworld, -1))
  if m is not None:
    bind(m, d)
  return d


def bind(m: types.Model, d: types.Data):
  """(Re)binds the device pointers of `d` to the C library (mjb_data_bind)."""
  if m.device.type != "cuda":
    raise RuntimeError("mujoco_warp_b200 kernels need a CUDA device; there is no CPU fallback.")
  lib = _lib.load()
  if d._handle is not None:
    _lib.data_destroy(lib, d._handle)
  d._handle = _lib.data_bind(lib, m, d)
  opt = _lib.make_option(m)
  _lib.check(lib, lib.mjb_prepare(m._handle, d._handle, ctypes.byref(opt)))  # first-call work done here: the next call is graph-capturable
  d._model = m
  d._bound_ptrs = _lib.data_ptr_signature(d)
  d._bound_epoch = types._EPOCH[0]  # set last: the assignments above bump the epoch


class MjDataLite(SimpleNamespace):
  """Host state container with `mujoco.MjData` attribute names (used when mujoco is absent)."""


def make_host_data(mjm) -> MjDataLite:
  """mujoco.MjData(mjm) equivalent for MjModelLite: state at qpos0."""
  return MjDataLite(
    time=0.0,
    qpos=np.asarray(mjm.qpos0, dtype=np.float64).copy(),
    qvel=np.zeros(mjm.nv),
    act=np.zeros(mjm.na),
    ctrl=np.zeros(mjm.nu),
    qacc_warmstart=np.zeros(mjm.nv),
    qfrc_applied=np.zeros(mjm.nv),
    xfrc_applied=np.zeros((mjm.nbody, 6)),
    mocap_pos=np.zeros((mjm.nmocap, 3)),
    mocap_quat=np.tile([1.0, 0, 0, 0], (mjm.nmocap, 1)),
    eq_active=np.asarray(getattr(mjm, "eq_active0", np.zeros(0)), dtype=bool).copy(),
    ncon=0,
    nefc=0,
  )


_STATE_FIELDS = ("qpos", "qvel", "act", "ctrl", "qacc_warmstart", "qfrc_applied", "xfrc_applied", "mocap_pos", "mocap_quat")


def put_data(
  mjm,
  mjd,
  nworld: int = 1,
  nconmax: Optional[int] = None,
  nccdmax: Optional[int] = None,
  njmax: Optional[int] = None,
  njmax_nnz: Optional[int] = None,
  naconmax: Optional[int] = None,
  naccdmax: Optional[int] = None,
  device=None,
  m: Optional[types.Model] = None,
) -> types.Data:
  """Moves host data to the device, tiled over `nworld` worlds (reference io.py:1016).

  State inputs (qpos, qvel, act, ctrl, qacc_warmstart, qfrc_applied, xfrc_applied, mocap, time,
  eq_active) are copied; derived quantities are recomputed by the first `forward`/`step`
  (the reference additionally copies MuJoCo C's derived arrays, io.py:1109-1238).
  """
  d = make_data(mjm, nworld, nconmax, nccdmax, njmax, njmax_nnz, naconmax, naccdmax, device=device, m=None)
  dev = d.device
  # The reference runs MuJoCo's mj_kinematics on the host data first (io.py:1108), which normalises the quaternions of qpos in
  # place -- an all-zero quaternion (keyframes of benchmarks/aloha_pot) becomes the identity (mju_normalize4).
  if hasattr(mjd, "qpos"):
    q = np.asarray(mjd.qpos, dtype=np.float64)
    for j in range(int(mjm.njnt)):
      jt, qa = int(mjm.jnt_type[j]), int(mjm.jnt_qposadr[j])
      if jt in (C.JointType.FREE, C.JointType.BALL):
        qa += 3 if jt == C.JointType.FREE else 0
        n = float(np.linalg.norm(q[qa : qa + 4]))
        q[qa : qa + 4] = [1.0, 0.0, 0.0, 0.0] if n < 1e-15 else q[qa : qa + 4] / n
    if isinstance(mjd.qpos, np.ndarray):
      mjd.qpos[...] = q
  for name in _STATE_FIELDS:
    if hasattr(mjd, name):
      src = np.asarray(getattr(mjd, name), dtype=np.float32)
      dst = getattr(d, name)
      if dst.numel() == 0:
        continue
      dst.copy_(torch.from_numpy(src.reshape(dst.shape[1:])).to(dev).unsqueeze(0).expand_as(dst))
  d.time.fill_(float(getattr(mjd, "time", 0.0)))
  if hasattr(mjd, "eq_active") and d.eq_active.numel():
    d.eq_active.copy_(torch.from_numpy(np.asarray(mjd.eq_active).astype(np.int32)).to(dev).unsqueeze(0).expand_as(d.eq_active))
  if m is not None:
    bind(m, d)
  return d


def get_data_into(result, mjm, d: types.Data, world_id: int = 0):
  """Copies world `world_id` of device data into a host object (reference io.py:1243).

  Fields are written as attributes with MuJoCo `MjData` names and float64 dtype.  Contacts of
  the world are gathered from the pool; efc rows are already in MuJoCo-C order on device, so
  the reference's re-ordering (io.py:1273-1299) is unnecessary.
  """
  w = world_id
  for name, shape in list(types.DATA_REAL_FIELDS.items()) + list(types.DATA_INT_FIELDS.items()):
    if name.startswith(("contact_", "efc_")) or name in ("nacon", "ncollision", "wcon", "wcon_n"):
      continue
    a = d.field(name)[w].detach().cpu().numpy()
    setattr(result, name, a.astype(np.float64) if a.dtype == np.float32 else a.copy())
  nefc = int(min(int(d.nefc[w]), d.njmax))
  nv = d._sizes["nv"]
  result.nefc = nefc
  for name in ("pos", "margin", "D", "vel", "aref", "frictionloss", "force", "type", "id", "state"):
    a = getattr(d.efc, name)[w, :nefc].detach().cpu().numpy()
    setattr(result, "efc_" + name, a.astype(np.float64) if a.dtype == np.float32 else a.copy())
  result.efc_J = d.efc.J[w, :nefc, :nv].detach().cpu().numpy().astype(np.float64)
  ncon_w = int(d.wcon_n[w])
  ids = d.wcon[w, :ncon_w].long()
  result.ncon = ncon_w
  con = SimpleNamespace()
  for f in dataclasses.fields(types.Contact):
    arr = getattr(d.contact, f.name)
    if arr is None:
      continue
    a = arr[ids].detach().cpu().numpy()
    setattr(con, f.name, a.astype(np.float64) if a.dtype == np.float32 else a)
  result.contact = con
  return result


def reset_data(m: types.Model, d: types.Data, reset=None):
  """Clears data and sets defaults, optionally per world (reference io.py:1458).

  One pass through the C ABI (mjb_reset_data: a contact-pool kernel and a per-world kernel), asynchronous on the current
  stream and graph capturable -- the mask stays on the device.  Like the reference it resets the state (qpos = qpos0, qvel, act,
  ctrl, qacc, qacc_warmstart, applied forces, eq_active, mocap poses, time, energy), the constraint counts, qM and the
  contacts of the reset worlds; derived fields are recomputed by the next `forward` / `step`.

  Args:
    reset: optional bool tensor / array of shape (nworld,); None resets every world.
  """
  from .forward import _ensure_bound

  lib = _lib.load()
  if d.qpos.device.type != "cuda":
    raise RuntimeError("reset_data runs on the CUDA device only: there is no CPU fallback")
  _ensure_bound(m, d)
  ptr, mask = ctypes.c_void_p(0), None
  if reset is not None:
    mask = torch.as_tensor(reset, device=d.qpos.device)
    if mask.shape != (d.nworld,):
      raise ValueError(f"reset must have shape ({d.nworld},), got {tuple(mask.shape)}")
    mask = (mask.view(torch.uint8) if mask.dtype == torch.bool else (mask != 0).view(torch.uint8)).contiguous()
    ptr = ctypes.c_void_p(mask.data_ptr())
  _lib.check(lib, lib.mjb_reset_data(m._handle, d._handle, ptr, _lib.stream_ptr()))


def find_keys(mjm, keyname_prefix: str) -> list:
  """Keyframe ids whose name starts with `keyname_prefix` (reference io.py:2591)."""
  return [i for i, n in enumerate(mjm.key_names) if n.startswith(keyname_prefix)]


def make_trajectory(mjm, keys) -> np.ndarray:
  """Control trajectory through the keyframes `keys`, linearly interpolated at the model timestep (reference io.py:2603)."""
  ctrls = []
  prev_ctrl_key = np.zeros(mjm.nu, dtype=np.float64)
  prev_time, time = 0.0, 0.0
  dt = float(mjm.opt.timestep)
  for key in keys:
    ctrl_key, ctrl_time = np.asarray(mjm.key_ctrl[key], dtype=np.float64), float(mjm.key_time[key])
    if not ctrls and ctrl_time != 0.0:
      raise ValueError("first keyframe must have time 0.0")
    elif ctrls and ctrl_time <= prev_time:
      raise ValueError("keyframes must be in time order")
    while time < ctrl_time:
      frac = (time - prev_time) / (ctrl_time - prev_time)
      ctrls.append(prev_ctrl_key * (1 - frac) + ctrl_key * frac)
      time += dt
    ctrls.append(ctrl_key)
    time += dt
    prev_ctrl_key = ctrl_key
    prev_time = time
  return np.array(ctrls)


def override_model(m: types.Model, overrides: dict):
  """Subset of reference io.py:2498 override_model: `{"opt.solver": "cg", "opt.iterations": 5, ...}`."""
  enums = {
    "opt.solver": C.SolverType, "opt.cone": C.ConeType, "opt.integrator": C.IntegratorType,
    "opt.broadphase": C.BroadphaseType,
  }  # fmt: skip
  for key, val in overrides.items():
    obj, attr = (m.opt, key[4:]) if key.startswith("opt.") else (m, key)
    if isinstance(val, str) and key in enums:
      val = int(enums[key][val.upper()])
    cur = getattr(obj, attr)
    if torch.is_tensor(cur):
      cur.copy_(torch.as_tensor(val, dtype=cur.dtype).reshape(cur.shape))
    else:
      setattr(obj, attr, type(cur)(val) if cur is not None else val)
