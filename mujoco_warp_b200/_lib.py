"""ctypes binding of libmjb200.so (the C ABI declared in include/mjb200.h).

The library is built in-tree by `build()` (nvcc, sm_100a).  There is no fallback: if the shared
object is missing or fails to load, every device entry point raises.
"""

from __future__ import annotations

import ctypes
import os
import subprocess
import threading
import time

import torch

from . import types

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libmjb200.so")
INCLUDE = os.path.join(_ROOT, "include")

NVCC_FLAGS = [
  "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math=false",
]  # fmt: skip

_lock = threading.Lock()
_lib = None


def _includes(path, seen=None):
  """Project headers a source file includes, transitively (csrc/ and include/ only)."""
  import re

  seen = set() if seen is None else seen
  with open(path) as fh:
    for name in re.findall(r'^\s*#\s*include\s+"([^"]+)"', fh.read(), flags=re.M):
      for base in (os.path.dirname(path), CSRC, INCLUDE):
        cand = os.path.normpath(os.path.join(base, name))
        if os.path.isfile(cand) and cand not in seen:
          seen.add(cand)
          _includes(cand, seen)
          break
  return seen


def _deps(src):
  """Project headers `src` really includes (the preprocessor decides: the stage-group units see only their own stage headers through
  `MJB_UNIT_GROUP`), falling back to the textual scan if the preprocessor run fails."""
  cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
  res = subprocess.run(["g++", "-MM", "-x", "c++", "-std=c++17", "-D__CUDACC__", "-I", INCLUDE, "-I", CSRC, "-I", cuda_inc, src],
                       capture_output=True, text=True)  # fmt: skip
  if res.returncode != 0:
    return sorted(_includes(src))
  names = res.stdout.replace("\\\n", " ").split(":", 1)[1].split()
  return sorted(os.path.normpath(n) for n in names if os.path.normpath(n) != os.path.normpath(src))


def build(force: bool = False, verbose: bool = False) -> str:
  """Compiles csrc/*.cu into csrc/libmjb200.so for sm_100a (works without a GPU).

  One object per translation unit under csrc/_obj/, rebuilt only when the unit or a header it includes changed: the step
  pipeline (mjb_api.cu, minutes) is not recompiled for an edit of the light units."""
  srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]
  objdir = os.path.join(CSRC, "_obj")
  os.makedirs(objdir, exist_ok=True)
  objs, relink = [], force or not os.path.exists(LIB_PATH)
  common = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xptxas", "-v" if verbose else "-O3",
            "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", CSRC]  # fmt: skip
  jobs = []
  for src in srcs:
    obj = os.path.join(objdir, os.path.splitext(os.path.basename(src))[0] + ".o")
    deps = [src] + (_deps(src) if os.path.exists(obj) and not force else [])
    if force or not os.path.exists(obj) or any(os.path.getmtime(obj) < os.path.getmtime(p) for p in deps):
      jobs.append((src, obj))
    objs.append(obj)

  def compile_one(job):
    src, obj = job
    started = time.time()
    res = subprocess.run(["nvcc"] + common + ["-c", src, "-o", obj], capture_output=True, text=True)
    if res.returncode != 0:
      raise RuntimeError(f"nvcc failed on {os.path.basename(src)}:\n" + res.stdout + res.stderr)
    os.utime(obj, (started, started))  # a header edited while the unit compiled must trigger the next rebuild
    return res.stderr

  if jobs:  # the stage-group kernels are separate units (mjb_g_*.cu): compile them side by side
    from concurrent.futures import ThreadPoolExecutor

    with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 1)) as pool:
      for log in pool.map(compile_one, jobs):
        if verbose:
          print(log)
    relink = True
  if relink or any(os.path.getmtime(LIB_PATH) < os.path.getmtime(o) for o in objs):
    res = subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH] + objs, capture_output=True, text=True)
    if res.returncode != 0:
      raise RuntimeError("nvcc link failed:\n" + res.stdout + res.stderr)
  return LIB_PATH


class MjbOption(ctypes.Structure):
  _fields_ = [(n, ctypes.c_int) for n in types.OPT_INT_FIELDS] + [(n, ctypes.c_float) for n in types.OPT_REAL_FIELDS]


def load():
  """Loads libmjb200.so (raises if it is missing -- there is no CPU / eager fallback)."""
  global _lib
  with _lock:
    if _lib is not None:
      return _lib
    path = os.environ.get("MJB200_LIB", LIB_PATH)  # A/B builds of the same sources (tools/ab_build.py)
    if not os.path.exists(path):
      raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc). There is no CPU fallback.")
    lib = ctypes.CDLL(path)
    vp, ip = ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)
    lib.mjb_field_name.restype = ctypes.c_char_p
    lib.mjb_field_name.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.mjb_field_count.restype = ctypes.c_int
    lib.mjb_field_count.argtypes = [ctypes.c_int]
    lib.mjb_model_create.restype = ctypes.c_int
    lib.mjb_model_create.argtypes = [ip, ctypes.c_int, vp, ctypes.c_size_t, ip, ctypes.c_int, ip, ctypes.c_int, ctypes.POINTER(vp)]
    lib.mjb_model_destroy.argtypes = [vp]
    lib.mjb_data_bind.restype = ctypes.c_int
    lib.mjb_data_bind.argtypes = [vp, ip, ctypes.c_int, ctypes.POINTER(vp), ctypes.c_int, ctypes.POINTER(vp), ctypes.c_int, ctypes.POINTER(vp)]
    lib.mjb_data_destroy.argtypes = [vp]
    for name in ("mjb_step", "mjb_forward", "mjb_step1", "mjb_step2", "mjb_rungekutta4", "mjb_rungekutta4_static"):
      fn = getattr(lib, name)
      fn.restype = ctypes.c_int
      fn.argtypes = [vp, vp, ctypes.POINTER(MjbOption), vp]
    lib.mjb_prepare.restype = ctypes.c_int
    lib.mjb_prepare.argtypes = [vp, vp, ctypes.POINTER(MjbOption)]
    lib.mjb_rk_static_max_floats.restype = ctypes.c_longlong
    lib.mjb_rk_scratch_floats.restype = ctypes.c_longlong
    lib.mjb_rk_scratch_floats.argtypes = [vp, vp]
    lib.mjb_run_stages.restype = ctypes.c_int
    lib.mjb_run_stages.argtypes = [vp, vp, ctypes.POINTER(MjbOption), ctypes.c_uint32, vp]
    lib.mjb_solve_m.restype = ctypes.c_int
    lib.mjb_solve_m.argtypes = [vp, vp, vp, vp, vp]
    lib.mjb_mul_m.restype = ctypes.c_int
    lib.mjb_mul_m.argtypes = [vp, vp, vp, vp, vp]
    lib.mjb_xfrc_accumulate.restype = ctypes.c_int
    lib.mjb_xfrc_accumulate.argtypes = [vp, vp, vp, vp]
    lib.mjb_deriv_smooth_vel.restype = ctypes.c_int
    lib.mjb_deriv_smooth_vel.argtypes = [vp, vp, vp, vp, ctypes.c_int, vp]
    lib.mjb_ctrl_noise.restype = ctypes.c_int
    lib.mjb_ctrl_noise.argtypes = [vp, vp, ctypes.POINTER(MjbOption), vp, ctypes.c_int, ctypes.c_float, ctypes.c_float, vp]
    lib.mjb_state_size.restype = ctypes.c_int
    lib.mjb_state_size.argtypes = [vp, vp, ctypes.c_int]
    for name in ("mjb_get_state", "mjb_set_state"):
      fn = getattr(lib, name)
      fn.restype = ctypes.c_int
      fn.argtypes = [vp, vp, vp, ctypes.c_int, ctypes.c_int, vp, vp]
    lib.mjb_reset_data.restype = ctypes.c_int
    lib.mjb_reset_data.argtypes = [vp, vp, vp, vp]
    lib.mjb_contact_force.restype = ctypes.c_int
    lib.mjb_contact_force.argtypes = [vp, vp, ctypes.c_int, vp, ctypes.c_int, ctypes.c_int, vp, vp]
    lib.mjb_launch_count.restype = ctypes.c_longlong
    lib.mjb_timing_enable.argtypes = [ctypes.c_int]
    lib.mjb_timing_ms.restype = ctypes.c_double
    lib.mjb_timing_launches.restype = ctypes.c_longlong
    lib.mjb_timing_group_ms.restype = ctypes.c_double
    lib.mjb_timing_group_ms.argtypes = [ctypes.c_int]
    lib.mjb_group_count.restype = ctypes.c_int
    lib.mjb_group_name.restype = ctypes.c_char_p
    lib.mjb_group_name.argtypes = [ctypes.c_int]
    lib.mjb_last_error.restype = ctypes.c_char_p
    lib.mjb_version.restype = ctypes.c_char_p
    _check_tables(lib)
    _lib = lib
    return lib


def field_names(lib, kind: int):
  return [lib.mjb_field_name(kind, i).decode() for i in range(lib.mjb_field_count(kind))]


def _check_tables(lib):
  """The Python shape tables must list exactly the fields of include/mjb_fields.h, in order."""
  want = [
    types.MODEL_SIZES, list(types.MODEL_INT_FIELDS), list(types.MODEL_REAL_FIELDS), types.OPT_INT_FIELDS,
    types.OPT_REAL_FIELDS, types.DATA_SIZES, list(types.DATA_INT_FIELDS), list(types.DATA_REAL_FIELDS),
  ]  # fmt: skip
  for kind, names in enumerate(want):
    got = field_names(lib, kind)
    if got != list(names):
      raise RuntimeError(f"field table {kind} mismatch between include/mjb_fields.h and types.py: {set(got) ^ set(names)}")


def check(lib, rc: int):
  if rc != 0:
    raise RuntimeError(f"libmjb200 error {rc}: {lib.mjb_last_error().decode()}")


_OPT_TENSORS = ("timestep", "tolerance", "ls_tolerance", "gravity", "impratio_invsqrt", "ccd_tolerance")
_OPT_INTS = ("integrator", "cone", "solver", "iterations", "ls_iterations", "disableflags", "enableflags", "broadphase", "broadphase_filter",
             "run_collision_detection", "ccd_iterations", "ls_parallel")  # fmt: skip


def make_option(m) -> MjbOption:
  """Snapshot of m.opt / m.stat for one call (options are passed by value).

  Users mutate m.opt between steps, so the struct is rebuilt whenever an int field, a tensor object or a
  tensor's in-place version counter changed; otherwise the cached struct is reused (this sits on the
  per-step host path)."""
  o = m.opt
  mi = m.stat.meaninertia
  for n in _OPT_TENSORS:  # the reference indexes these as [worldid % shape[0]]; here one value serves all worlds
    if getattr(o, n).numel() != (3 if n == "gravity" else 1):
      raise NotImplementedError(f"per-world Option.{n} (shape {tuple(getattr(o, n).shape)}) is not implemented: the kernels take one value for all worlds")
  if mi.numel() != 1:
    raise NotImplementedError("per-world stat.meaninertia is not implemented")
  if bool(o.ls_parallel):  # reference solver.py:331-566: N step sizes evaluated at once; the in-kernel search here is the iterative one
    raise NotImplementedError("opt.ls_parallel (parallel linesearch) is not implemented: use the iterative linesearch (ls_parallel=False)")
  key = tuple(getattr(o, n) for n in _OPT_INTS) + tuple(x for n in _OPT_TENSORS for x in (id(getattr(o, n)), getattr(o, n)._version)) + (id(mi), mi._version)
  cached = getattr(m, "_opt_cache", None)
  if cached is not None and cached[0] == key:
    return cached[1]
  g = o.gravity.reshape(-1)
  opt = MjbOption(
    integrator=int(o.integrator), cone=int(o.cone), solver=int(o.solver), iterations=int(o.iterations),
    ls_iterations=int(o.ls_iterations), disableflags=int(o.disableflags), enableflags=int(o.enableflags),
    broadphase=int(o.broadphase), broadphase_filter=int(o.broadphase_filter),
    run_collision_detection=int(bool(o.run_collision_detection)), ccd_iterations=int(o.ccd_iterations),
    timestep=float(o.timestep.reshape(-1)[0]), tolerance=float(o.tolerance.reshape(-1)[0]),
    ls_tolerance=float(o.ls_tolerance.reshape(-1)[0]), gravity_x=float(g[0]), gravity_y=float(g[1]), gravity_z=float(g[2]),
    impratio_invsqrt=float(o.impratio_invsqrt.reshape(-1)[0]), meaninertia=float(mi.reshape(-1)[0]),
    ccd_tolerance=float(o.ccd_tolerance.reshape(-1)[0]),
  )  # fmt: skip
  m._opt_cache = (key, opt)
  return opt


def model_create(lib, m, int_names, real_names):
  sizes = (ctypes.c_int * len(types.MODEL_SIZES))(*[int(m.sizes[n]) for n in types.MODEL_SIZES])
  ioff = (ctypes.c_int * len(int_names))(*[m._offsets[n] for n in int_names])
  roff = (ctypes.c_int * len(real_names))(*[m._offsets[n] for n in real_names])
  out = ctypes.c_void_p()
  check(lib, lib.mjb_model_create(sizes, len(types.MODEL_SIZES), ctypes.c_void_p(m._blob.data_ptr()), m._blob.numel(), ioff,
                                  len(int_names), roff, len(real_names), ctypes.byref(out)))  # fmt: skip
  return out


def data_ptr_signature(d):
  return tuple(d.field(n).data_ptr() for n in list(types.DATA_INT_FIELDS) + list(types.DATA_REAL_FIELDS))


def data_bind(lib, m, d):
  int_names, real_names = list(types.DATA_INT_FIELDS), list(types.DATA_REAL_FIELDS)
  sizes_all = dict(m.sizes)
  sizes_all.update(d._sizes)
  for names, dtype in ((int_names, torch.int32), (real_names, torch.float32)):
    for n in names:
      t = d.field(n)
      if t.dtype != dtype or not t.is_contiguous() or t.device != m.device:
        raise ValueError(f"Data.{n}: expected a contiguous {dtype} tensor on {m.device}, got {t.dtype} on {t.device} (contiguous={t.is_contiguous()})")
      want = types.data_field_numel(n, sizes_all)
      if want is not None and t.numel() != want:
        raise ValueError(f"Data.{n}: expected {want} elements for this model / nworld, got shape {tuple(t.shape)}")
  sizes = (ctypes.c_int * len(types.DATA_SIZES))(*[int(d._sizes[n]) for n in types.DATA_SIZES])
  ip = (ctypes.c_void_p * len(int_names))(*[d.field(n).data_ptr() for n in int_names])
  rp = (ctypes.c_void_p * len(real_names))(*[d.field(n).data_ptr() for n in real_names])
  out = ctypes.c_void_p()
  check(lib, lib.mjb_data_bind(m._handle, sizes, len(types.DATA_SIZES), ip, len(int_names), rp, len(real_names), ctypes.byref(out)))
  return out


def data_destroy(lib, handle):
  lib.mjb_data_destroy(handle)


def stream_ptr():
  return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
