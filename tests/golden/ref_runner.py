"""Runs the reference's own `put_model / make_data / forward / step` (source files under /root/reference, unmodified)
inside the numpy Warp interpreter (`warp_interp.py`) with `mujoco_stub.py` standing in for the mujoco module.

The host model is this repo's `MjModelLite` (an MjModel look-alike built by `mujoco_warp_b200.mjcf`); fields of
MjModel that the lite model does not carry (cameras, lights, flex, meshes, tendons, plugins, ...) are presented to the
reference as empty arrays of the right element shape.

Used by the golden-vector generators only (the reference tree does not exist on the GPU box).
"""

from __future__ import annotations

import dataclasses
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (HERE, ROOT):
  if p not in sys.path:
    sys.path.insert(0, p)

import mujoco_stub  # noqa: E402
import warp_interp  # noqa: E402

REFERENCE = "/root/reference"
KERNEL_MODULES = (
  "smooth forward solver constraint collision_driver collision_primitive collision_primitive_core collision_core collision_convex "
  "collision_gjk support passive math derivative util_misc block_cholesky sensor"
).split()

_ref = None


class _KernelBoolOps(__import__("ast").NodeTransformer):
  """Inside @wp.func / @wp.kernel bodies, `a and b` / `a or b` yield a bool in Warp (the operands are converted), whereas Python
  returns one of the operands: `float((v1 & 4) and (v2 & 4))` is 1.0 in Warp and 4.0 in Python (collision_gjk.py:1630-1650 relies on
  it).  Wraps the operands of boolean operators in bool() -- only inside device functions; host code keeps Python semantics
  (`qvel or d.qvel`)."""

  def __init__(self):
    self.depth = 0

  @staticmethod
  def _is_device(node):
    import ast

    for d in node.decorator_list:
      f = d.func if isinstance(d, ast.Call) else d
      if isinstance(f, ast.Attribute) and isinstance(f.value, ast.Name) and f.value.id == "wp" and f.attr in ("func", "kernel"):
        return True
    return False

  def visit_FunctionDef(self, node):
    dev = self._is_device(node)
    self.depth += dev
    self.generic_visit(node)
    self.depth -= dev
    return node

  def visit_BoolOp(self, node):
    import ast

    self.generic_visit(node)
    if self.depth > 0:
      node.values = [ast.Call(func=ast.Name(id="bool", ctx=ast.Load()), args=[v], keywords=[]) for v in node.values]
    return node


def _install_source_transform():
  """Import hook: modules of the reference package are compiled from their (unmodified) source files through _KernelBoolOps."""
  import ast
  import importlib.abc
  import importlib.machinery

  class Loader(importlib.machinery.SourceFileLoader):
    def source_to_code(self, data, path, *, _optimize=-1):
      tree = ast.parse(data, filename=path)
      tree = ast.fix_missing_locations(_KernelBoolOps().visit(tree))
      return compile(tree, path, "exec", dont_inherit=True, optimize=_optimize)

  class Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, name, path, target=None):
      if not (name == "mujoco_warp" or name.startswith("mujoco_warp.")):
        return None
      spec = importlib.machinery.PathFinder.find_spec(name, path)
      if spec is not None and isinstance(spec.loader, importlib.machinery.SourceFileLoader):
        spec.loader = Loader(spec.loader.name, spec.loader.path)
      return spec

  if not any(type(f).__name__ == "Finder" and f.__module__ == __name__ for f in sys.meta_path):
    sys.meta_path.insert(0, Finder())


def load_reference():
  """Imports /root/reference/mujoco_warp under the interpreter; returns (wp, mujoco_warp module)."""
  global _ref
  if _ref is not None:
    return _ref
  wp = warp_interp.install()
  mujoco_stub.install()
  sys.dont_write_bytecode = True
  _install_source_transform()
  if REFERENCE not in sys.path:
    sys.path.insert(0, REFERENCE)
  import importlib

  ref = importlib.import_module("mujoco_warp")
  for name in KERNEL_MODULES:
    mod = sys.modules.get(f"mujoco_warp._src.{name}")
    if mod is not None:
      warp_interp.patch_module(mod)
  _ref = (wp, ref)
  return _ref


_MJMODEL_EXTRA_SHAPES = {  # MjModel arrays the reference reads that are not fields of its Model dataclass
  "cam_output": (0,), "flex_rgba": (0, 4), "flex_selfcollide": (0,), "mesh_facetexcoord": (0, 3), "mesh_texcoord": (0, 2),
  "mesh_texcoordadr": (0,), "numeric_adr": (0,), "numeric_data": (0,), "sensor_needstage": (0,), "sensor_plugin": (0,),
  "actuator_plugin": None, "body_plugin": None, "geom_plugin": None, "plugin_attradr": (0,), "plugin_attr": (0,), "plugin": (0,),
  "name_numericadr": (0,), "names": b"", "tex_data": (0,), "tex_adr": (0,), "tex_height": (0,), "tex_width": (0,), "tex_nchannel": (0,),
}  # fmt: skip


class MjModelAdapter:
  """MjModelLite + empty defaults for every other MjModel field the reference touches."""

  def __init__(self, mjm, types_mod):
    object.__setattr__(self, "_m", mjm)
    object.__setattr__(self, "_fields", {f.name: f.type for f in dataclasses.fields(types_mod.Model)})

  def __getattr__(self, name):
    m = object.__getattribute__(self, "_m")
    if hasattr(m, name):
      return getattr(m, name)
    if name == "nJmom":  # MuJoCo: non-zeros of actuator_moment = dofs of each actuator's joint (joint transmissions only)
      dofs = {0: 6, 1: 3, 2: 1, 3: 1}
      return int(sum(dofs[int(m.jnt_type[int(m.actuator_trnid[a][0])])] for a in range(m.nu)))
    if name in ("actuator_plugin", "body_plugin", "geom_plugin"):
      n = {"actuator_plugin": m.nu, "body_plugin": m.nbody, "geom_plugin": m.ngeom}[name]
      return -np.ones(n, dtype=np.int32)
    if name in _MJMODEL_EXTRA_SHAPES:
      shp = _MJMODEL_EXTRA_SHAPES[name]
      return shp if isinstance(shp, bytes) else np.zeros(shp)
    if name == "vis":
      import types as _t

      return _t.SimpleNamespace(global_=_t.SimpleNamespace(bvactive=0), quality=_t.SimpleNamespace(shadowsize=0))
    fields = object.__getattribute__(self, "_fields")
    if name in fields:
      t = fields[name]
      if t is int or (isinstance(t, str) and t == "int"):
        return 0
      if hasattr(t, "dtype") and getattr(t, "shape", None) is not None:
        dims = []
        for s in t.shape:
          if s == "*":
            continue
          dims.append(int(getattr(self, s)) if isinstance(s, str) else int(s))
        es = warp_interp._elem_shape(t.dtype)
        npdt = warp_interp._np_dtype(t.dtype)
        return np.zeros(tuple(dims) + es, dtype=npdt)
    if name.startswith("n") and name[1:].isalpha():
      return 0
    raise AttributeError(f"MjModelAdapter: '{name}' is neither on the lite model nor a known default")


def make(mjm, nworld, nconmax, njmax):
  """Reference put_model + make_data on the lite host model."""
  wp, ref = load_reference()
  types_mod = sys.modules["mujoco_warp._src.types"]
  adapter = MjModelAdapter(mjm, types_mod)
  m = ref.put_model(adapter)
  d = ref.make_data(adapter, nworld=nworld, nconmax=nconmax, njmax=njmax)
  return m, d


def set_state(d, qpos, qvel, ctrl=None, qacc_warmstart=None):
  d.qpos.numpy()[...] = qpos
  d.qvel.numpy()[...] = qvel
  if ctrl is not None:
    d.ctrl.numpy()[...] = ctrl
  if qacc_warmstart is not None:
    d.qacc_warmstart.numpy()[...] = qacc_warmstart


def snapshot(d, names):
  out = {}
  for n in names:
    o = d
    for part in n.split("."):
      o = getattr(o, part)
    out[n.replace(".", "_")] = np.array(o.numpy() if hasattr(o, "numpy") else o)
  return out
