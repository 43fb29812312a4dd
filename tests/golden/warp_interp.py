"""A float64 numpy interpreter for the part of the `warp` API that the reference's step path uses.

Purpose: `warp-lang` and `mujoco` cannot be installed here, so the reference cannot run as shipped.  Its kernels are,
however, plain Python functions decorated with `@wp.kernel` / `@wp.func`, launched by `wp.launch(kernel, dim, inputs,
outputs)`.  With this module registered as `warp` (and `mujoco_stub.py` as `mujoco`) the reference's OWN source files
under /root/reference import unmodified and `mujoco_warp.step(m, d)` executes thread by thread in the Python
interpreter, in float64.  The generator scripts in this directory use that to produce golden vectors that pin the CPU
oracle to the reference's algorithm (the float32 rounding of the real Warp build is a separate matter, covered by the
device-vs-oracle tolerances).

Only the golden-vector generators import this file; nothing in the product, the tests or the bench does.

Semantics reproduced
  * kernels run one thread at a time in launch order (a valid schedule of a data-race-free kernel; atomics are then
    plain read-modify-writes).  Tiled launches run with a block of ONE thread (`wp.block_dim() == 1`), which makes every
    cooperative loop `for i in range(tid, n, wp.block_dim())` a full loop and every tile reduction exact.
  * vectors / matrices are VALUES: indexing an array element or a matrix row copies, augmented assignment rebinds.
  * integers follow C: `int / int` truncates (reference constraint.py:1781 relies on it).  `wint` carries that through
    arithmetic; kernel modules get `int` / `range` rebound to the `wint` versions by `patch_module`.
  * `wp.tid()` returns as many indices as the call site unpacks (read from the caller's bytecode), as Warp's code
    generator does.
"""

from __future__ import annotations

import dis
import itertools
import math as _math
import sys
import types as _types

import numpy as np

_F = np.float64  # scalars are numpy float64: IEEE semantics (x / 0 = inf, 0 / 0 = nan) like the compiled kernels, no Python exceptions

# ------------------------------------------------------------------------------------------------
# integers with C semantics
# ------------------------------------------------------------------------------------------------


class wint(int):
  __slots__ = ()

  def __truediv__(self, o):
    if isinstance(o, int) and not isinstance(o, bool):
      a, b = int(self), int(o)
      q = abs(a) // abs(b)
      return wint(q if (a >= 0) == (b >= 0) else -q)
    return int.__truediv__(self, o)

  def __rtruediv__(self, o):
    if isinstance(o, int) and not isinstance(o, bool):
      return wint(o).__truediv__(self)
    return int.__rtruediv__(self, o)

  def __repr__(self):
    return int.__repr__(self)


def _wrap_int_op(name):
  base = getattr(int, name)

  def f(self, *a):
    r = base(self, *a)
    return wint(r) if type(r) is int else r

  f.__name__ = name
  return f


for _n in "add sub mul floordiv mod radd rsub rmul rfloordiv rmod neg pos abs and or xor rand ror rxor lshift rshift invert".split():
  setattr(wint, f"__{_n}__", _wrap_int_op(f"__{_n}__"))


class wrange:
  """range() whose items are wints."""

  def __init__(self, *a):
    self._r = range(*[int(x) for x in a])

  def __iter__(self):
    return (wint(i) for i in self._r)

  def __reversed__(self):
    return (wint(i) for i in reversed(self._r))

  def __len__(self):
    return len(self._r)

  def __getitem__(self, i):
    return wint(self._r[i])


def _wint_ctor(x=0):
  if isinstance(x, (bool, np.bool_, str, bytes)):
    return wint(int(x))
  return wint(_math.trunc(x))


def patch_module(mod):
  """Rebinds `int` and `range` inside a reference kernel module to the C-semantics versions."""
  mod.int = _IntMeta
  mod.range = wrange
  mod.float = _F


class _IntType(type):
  def __instancecheck__(cls, o):
    return isinstance(o, int)

  def __call__(cls, x=0):
    return _wint_ctor(x)


class _IntMeta(metaclass=_IntType):
  """Callable like int() (returns wint) and usable in isinstance checks."""


# ------------------------------------------------------------------------------------------------
# vector / matrix value types
# ------------------------------------------------------------------------------------------------


def _scalar(v, isint):
  if isint:
    return wint(int(v))
  return _F(v)


class _Arr(np.ndarray):
  _shape = None
  _isint = False

  def __new__(cls, *args):
    shape = cls._shape
    dt = np.int64 if cls._isint else np.float64
    a = np.zeros(shape, dtype=dt).view(cls)
    if len(args) == 1 and np.ndim(args[0]) == 0:
      a[...] = args[0]
    elif len(args) == 1:
      a[...] = np.asarray(args[0], dtype=dt).reshape(shape)
    elif len(args) > 1:
      flat = []
      for x in args:  # scalars, or rows / sub-vectors
        flat.extend(np.asarray(x, dtype=dt).reshape(-1).tolist())
      a[...] = np.asarray(flat, dtype=dt).reshape(shape)
    return a

  def __array_finalize__(self, obj):
    pass

  def __getitem__(self, idx):
    r = np.ndarray.__getitem__(self.view(np.ndarray), idx)
    if isinstance(r, np.ndarray) and r.ndim > 0:
      return _wrap(r.copy(), self._isint)
    return _scalar(r, self._isint)

  def __setitem__(self, idx, val):
    np.ndarray.__setitem__(self.view(np.ndarray), idx, val)

  def __mul__(self, o):
    # Warp: matrix * vector and matrix * matrix are products, everything else is element-wise / scalar
    if self.ndim == 2 and isinstance(o, np.ndarray) and o.ndim >= 1:
      return _wrap(np.asarray(self) @ np.asarray(o))
    return _wrap(np.multiply(np.asarray(self), np.asarray(o)), self._isint and _is_intlike(o))

  def __rmul__(self, o):
    return _wrap(np.multiply(np.asarray(o), np.asarray(self)), self._isint and _is_intlike(o))

  def __matmul__(self, o):
    return _wrap(np.asarray(self) @ np.asarray(o))

  def __add__(self, o):
    return _wrap(np.add(np.asarray(self), np.asarray(o)), self._isint and _is_intlike(o))

  __radd__ = __add__

  def __sub__(self, o):
    return _wrap(np.subtract(np.asarray(self), np.asarray(o)), self._isint and _is_intlike(o))

  def __rsub__(self, o):
    return _wrap(np.subtract(np.asarray(o), np.asarray(self)), self._isint and _is_intlike(o))

  def __truediv__(self, o):
    return _wrap(np.true_divide(np.asarray(self), np.asarray(o)))

  def __neg__(self):
    return _wrap(np.negative(np.asarray(self)), self._isint)

  def __pos__(self):
    return _wrap(np.asarray(self).copy(), self._isint)

  # value semantics for augmented assignment
  def __iadd__(self, o):
    return self + o

  def __isub__(self, o):
    return self - o

  def __imul__(self, o):
    return self * o

  def __itruediv__(self, o):
    return self / o

  def __eq__(self, o):
    return bool(np.array_equal(np.asarray(self), np.asarray(o)))

  def __ne__(self, o):
    return not self.__eq__(o)

  __hash__ = None

  def __len__(self):
    return self.shape[0]

  def __iter__(self):
    for i in range(self.shape[0]):
      yield self[i]

  def __repr__(self):
    return f"{type(self).__name__}({np.asarray(self).tolist()})"

  x = property(lambda s: s[0])
  y = property(lambda s: s[1])
  z = property(lambda s: s[2])
  w = property(lambda s: s[3])


def _is_intlike(o):
  if isinstance(o, bool):
    return False
  if isinstance(o, int):
    return True
  return isinstance(o, np.ndarray) and o.dtype.kind in "iu"


_TYPE_CACHE = {}


def _make(shape, isint=False):
  key = (tuple(int(s) for s in shape), bool(isint))
  if key not in _TYPE_CACHE:
    nm = ("vec%d" % key[0][0] if len(key[0]) == 1 else "mat%d%d" % key[0]) + ("i" if isint else "")
    _TYPE_CACHE[key] = type(nm, (_Arr,), {"_shape": key[0], "_isint": key[1]})
  return _TYPE_CACHE[key]


def _is_int_dtype(dtype):
  return dtype in (int, wint, _IntMeta, np.int32, np.int64) or getattr(dtype, "_wp_int", False)


def _vector(length, dtype=float):
  return _make((length,), _is_int_dtype(dtype))


def _matrix(shape, dtype=float):
  return _make(tuple(shape), _is_int_dtype(dtype))


def _wrap(v, isint=False):
  if isinstance(v, np.ndarray) and v.ndim > 0:
    isint = isint and v.dtype.kind in "iu"
    v = np.ascontiguousarray(v, dtype=np.int64 if isint else np.float64)
    return v.view(_make(v.shape, isint))
  if isinstance(v, (bool, np.bool_)):
    return bool(v)
  if isint or isinstance(v, (int, np.integer)):
    return wint(int(v))
  return _F(v)


# ------------------------------------------------------------------------------------------------
# arrays
# ------------------------------------------------------------------------------------------------


def _elem_shape(dtype):
  if isinstance(dtype, type) and issubclass(dtype, _Arr):
    return tuple(dtype._shape)
  return ()


def _np_dtype(dtype):
  if isinstance(dtype, type) and issubclass(dtype, _Arr):
    return np.int64 if dtype._isint else np.float64
  if dtype in (bool, np.bool_):
    return np.bool_
  if _is_int_dtype(dtype) or getattr(dtype, "_wp_uint", False):
    return np.int64
  return np.float64


class warray:
  """wp.array: numpy storage of shape (logical shape) + (element shape), logical indexing."""

  _default_ndim = 1

  def __init__(self, data=None, dtype=None, shape=None, ndim=None, device=None, copy=True, **kw):
    self.device = "cpu"
    if data is None and shape is None:  # annotation placeholder (reference types.py:692)
      self.dtype = dtype
      self.ndim = ndim if ndim is not None else self._default_ndim
      self.shape = (0,)
      self.strides = (0,)
      self._a = None
      return
    if isinstance(data, warray):
      if dtype is None:
        dtype = data.dtype
      data = data._a
    if dtype is None:
      dtype = float
      if data is not None:
        k = np.asarray(data).dtype.kind
        dtype = bool if k == "b" else (int if k in "iu" else float)
    es = _elem_shape(dtype)
    npdt = _np_dtype(dtype)
    if data is None:
      shape = (shape,) if np.ndim(shape) == 0 else tuple(shape)
      a = np.zeros(tuple(int(s) for s in shape) + es, dtype=npdt)
    else:
      a = np.array(data, dtype=npdt)
      if shape is not None:
        shape = (shape,) if np.ndim(shape) == 0 else tuple(shape)
        a = a.reshape(tuple(int(s) for s in shape) + es)
      elif es:
        # trailing dims hold the element; if they do not match, the data is flat per element
        if a.shape[a.ndim - len(es):] != es:
          a = a.reshape(a.shape[:-1] + (-1,) + es) if a.ndim >= 1 else a.reshape((-1,) + es)
    self._bind(a, dtype)

  def _bind(self, a, dtype):
    self._a = a
    self.dtype = dtype
    es = _elem_shape(dtype)
    lshape = a.shape[: a.ndim - len(es)]
    self.shape = tuple(wint(s) for s in lshape)
    self.ndim = len(lshape)
    self.strides = a.strides[: self.ndim]
    return self

  @classmethod
  def _view(cls, a, dtype):
    o = warray.__new__(warray)
    o.device = "cpu"
    return o._bind(a, dtype)

  def __class_getitem__(cls, dtype):
    return cls(dtype=dtype, ndim=cls._default_ndim)

  # ---- element access -------------------------------------------------------------------------
  def _elem(self, v):
    dt = self.dtype
    if isinstance(dt, type) and issubclass(dt, _Arr):
      return np.array(v).view(dt)
    if dt in (bool, np.bool_):
      return bool(v)
    if self._a.dtype.kind in "iu":
      return wint(int(v))
    return _F(v)

  def __getitem__(self, idx):
    if not isinstance(idx, tuple):
      idx = (idx,)
    nint = sum(1 for i in idx if not isinstance(i, slice))
    if len(idx) == self.ndim and nint == self.ndim:
      for k, i in enumerate(idx):
        if not (0 <= i < self.shape[k]):
          raise IndexError(f"index {tuple(int(j) for j in idx)} out of bounds for array of shape {self.shape}")
      dt = self.dtype
      if isinstance(dt, type) and issubclass(dt, _Arr) and _store_through(sys._getframe(1)):
        return self._a[tuple(int(i) for i in idx)].view(dt)  # `a[i, j][k] = v` writes the array element in Warp
      return self._elem(self._a[tuple(int(i) for i in idx)])
    if len(idx) > self.ndim:
      raise IndexError("too many indices")
    for k, i in enumerate(idx):
      if not isinstance(i, slice) and not (0 <= i < self.shape[k]):
        raise IndexError(f"index {idx} out of bounds for array of shape {self.shape}")
    return warray._view(self._a[tuple(i if isinstance(i, slice) else int(i) for i in idx)], self.dtype)

  def __setitem__(self, idx, val):
    if not isinstance(idx, tuple):
      idx = (idx,)
    if len(idx) == self.ndim:
      for k, i in enumerate(idx):
        if not isinstance(i, slice) and not (0 <= i < self.shape[k]):
          raise IndexError(f"index {tuple(int(j) for j in idx)} out of bounds for array of shape {self.shape}")
    if isinstance(val, warray):
      val = val._a
    self._a[tuple(i if isinstance(i, slice) else int(i) for i in idx)] = val

  def __len__(self):
    return int(self.shape[0])

  # ---- host API -------------------------------------------------------------------------------
  @property
  def size(self):
    return int(np.prod([int(s) for s in self.shape])) if self._a is not None else 0

  @property
  def ptr(self):
    return self._a.ctypes.data if self._a is not None else 0

  def numpy(self):
    return self._a

  def list(self):
    return self._a.tolist()

  def zero_(self):
    self._a[...] = 0
    return self

  def fill_(self, v):
    self._a[...] = v
    return self

  def assign(self, src):
    self._a[...] = src._a if isinstance(src, warray) else np.asarray(src)
    return self

  def reshape(self, shape):
    shape = (shape,) if np.ndim(shape) == 0 else tuple(shape)
    return warray._view(self._a.reshape(tuple(int(s) for s in shape) + _elem_shape(self.dtype)), self.dtype)

  def flatten(self):
    return self.reshape((-1,))

  def view(self, dtype):
    return warray._view(self._a, dtype)

  def contiguous(self):
    return self

  def to(self, device):
    return self

  def __repr__(self):
    return f"warray(shape={self.shape}, dtype={getattr(self.dtype, '__name__', self.dtype)})"


def _arr_cls(nd):
  return type(f"array{nd}d", (warray,), {"_default_ndim": nd})


def _full_like_ctor(fillv):
  def f(shape=None, dtype=float, device=None, **kw):
    a = warray(shape=shape if shape is not None else 0, dtype=dtype)
    if fillv is not None:
      a._a[...] = fillv
    return a

  return f


def _full(shape=None, value=0, dtype=float, device=None, **kw):
  a = warray(shape=shape, dtype=dtype)
  a._a[...] = value
  return a


def _clone(a, **kw):
  return warray._view(a._a.copy(), a.dtype)


def _copy(dest, src, dest_offset=0, src_offset=0, count=0, **kw):
  if count or dest_offset or src_offset:
    d = dest._a.reshape((-1,) + _elem_shape(dest.dtype))
    s = src._a.reshape((-1,) + _elem_shape(src.dtype))
    n = count if count else min(len(d) - dest_offset, len(s) - src_offset)
    d[dest_offset : dest_offset + n] = s[src_offset : src_offset + n]
  else:
    dest._a[...] = src._a.reshape(dest._a.shape)


# ------------------------------------------------------------------------------------------------
# kernels and launches
# ------------------------------------------------------------------------------------------------

_STATE = {"tid": (), "block_dim": 1, "launches": []}
_ARITY = {}


class Kernel:
  def __init__(self, func):
    self.func = func
    self.key = func.__qualname__
    self.__name__ = func.__name__
    self.__wrapped__ = func

  def __call__(self, *a, **k):
    return self.func(*a, **k)


def _kernel(f=None, **kw):
  if f is not None and callable(f):
    return Kernel(f)
  return lambda g: Kernel(g)


def _copy_struct(o):
  """Warp passes structs to functions BY VALUE (reference collision_gjk.py:2226 relies on it: ccd() shrinks its Geom arguments
  and returns without restoring them).  Arrays inside a struct stay shared, vectors are copied."""
  c = type(o).__new__(type(o))
  for k, v in o.__dict__.items():
    if isinstance(v, _Arr):
      v = v.copy().view(type(v))
    elif getattr(type(v), "_wp_struct", False):
      v = _copy_struct(v)
    c.__dict__[k] = v
  return c


def _func(f=None, **kw):
  def wrap(g):
    import functools

    @functools.wraps(g)
    def call(*args, **kwargs):
      if any(getattr(type(a), "_wp_struct", False) for a in args):
        args = tuple(_copy_struct(a) if getattr(type(a), "_wp_struct", False) else a for a in args)
      return g(*args, **kwargs)

    return call

  if f is not None and callable(f):
    return wrap(f)
  return wrap


def _func_native(snippet=None, **kw):
  def deco(g):
    def noop(*a, **k):
      return None

    return noop

  return deco


_STORE_THROUGH = {}


def _store_through(f):
  """True when the subscript being evaluated in frame `f` is the target of a nested element store `a[i, j][k] = v`
  (bytecode: BINARY_SUBSCR, simple loads of k, STORE_SUBSCR): Warp writes the array element, a copy would lose the write."""
  key = (f.f_code, f.f_lasti)
  r = _STORE_THROUGH.get(key)
  if r is None:
    ins = list(dis.get_instructions(f.f_code))
    r = False
    for k, i in enumerate(ins):
      if i.offset <= f.f_lasti and (k + 1 == len(ins) or ins[k + 1].offset > f.f_lasti):
        j = k + 1
        while j < len(ins) and ins[j].opname in ("LOAD_FAST", "LOAD_CONST", "LOAD_DEREF", "LOAD_GLOBAL", "LOAD_FAST_CHECK"):
          j += 1
        r = i.opname == "BINARY_SUBSCR" and j > k + 1 and j < len(ins) and ins[j].opname == "STORE_SUBSCR"
        break
    _STORE_THROUGH[key] = r
  return r


def _tid():
  f = sys._getframe(1)
  key = (f.f_code, f.f_lasti)
  n = _ARITY.get(key)
  if n is None:
    ins = list(dis.get_instructions(f.f_code))
    n = 1
    for k, i in enumerate(ins):  # f_lasti may point into the CALL's inline cache entries
      if i.offset <= f.f_lasti and (k + 1 == len(ins) or ins[k + 1].offset > f.f_lasti):
        nxt = ins[k + 1]
        if nxt.opname == "UNPACK_SEQUENCE":
          n = nxt.arg
        break
    _ARITY[key] = n
  t = _STATE["tid"]
  if n == 1:
    return t[0]
  return t[:n]


def _launch(kernel, dim, inputs=(), outputs=(), device=None, block_dim=None, stream=None, tiled=False, **kw):
  dims = (int(dim),) if np.ndim(dim) == 0 else tuple(int(x) for x in dim)
  if tiled:
    dims = dims + (1,)
  func = kernel.func if isinstance(kernel, Kernel) else kernel
  args = list(inputs) + list(outputs)
  _STATE["launches"].append(getattr(func, "__qualname__", str(func)))
  saved = (_STATE["tid"], _STATE["block_dim"])
  _STATE["block_dim"] = 1
  try:
    for t in itertools.product(*[range(d) for d in dims]):
      _STATE["tid"] = tuple(wint(i) for i in t)
      func(*args)
  finally:
    _STATE["tid"], _STATE["block_dim"] = saved


def _launch_tiled(kernel, dim, inputs=(), outputs=(), block_dim=None, **kw):
  return _launch(kernel, dim, inputs, outputs, tiled=True, **kw)


def _atomic(op):
  def f(arr, *a):
    *idx, val = a
    if isinstance(arr, _Arr):  # atomic on a local vector component
      old = arr[idx[0]]
      arr[idx[0]] = op(old, val)
      return old
    idx = tuple(idx)
    old = arr[idx]
    arr[idx] = op(old, val)
    return old

  return f


# ------------------------------------------------------------------------------------------------
# tiles (block of one thread): plain float64 / int64 ndarrays with scalar-returning indexing
# ------------------------------------------------------------------------------------------------


class Tile(np.ndarray):
  def __getitem__(self, idx):
    r = np.ndarray.__getitem__(self.view(np.ndarray), idx)
    if isinstance(r, np.ndarray) and r.ndim > 0:
      et = getattr(self, "_etype", None)
      if et is not None and r.shape == tuple(et._shape):
        return r.copy().view(et)
      return r.view(Tile)
    if isinstance(r, np.ndarray):
      r = r[()]
    if isinstance(r, (np.integer,)):
      return wint(int(r))
    if isinstance(r, np.bool_):
      return bool(r)
    return _F(r)

  def __setitem__(self, idx, val):
    np.ndarray.__setitem__(self.view(np.ndarray), idx, val)

  def __array_finalize__(self, obj):
    self._etype = getattr(obj, "_etype", None)


def _tile(a, etype=None):
  t = np.ascontiguousarray(a).view(Tile)
  t._etype = etype
  return t


def _shape_t(shape):
  return (int(shape),) if np.ndim(shape) == 0 else tuple(int(s) for s in shape)


def _tile_load(arr, shape, offset=None, storage=None, bounds_check=True, **kw):
  shape = _shape_t(shape)
  offset = (0,) * len(shape) if offset is None else _shape_t(offset)
  src = arr._a
  out = np.zeros(shape + _elem_shape(arr.dtype), dtype=src.dtype)
  sl_src, sl_dst = [], []
  for k, (o, n) in enumerate(zip(offset, shape)):
    hi = min(src.shape[k], o + n)
    if hi <= o:
      return _tile(out)
    sl_src.append(slice(o, hi))
    sl_dst.append(slice(0, hi - o))
  out[tuple(sl_dst)] = src[tuple(sl_src)]
  return _tile(out)


def _tile_store(arr, tile, offset=None, bounds_check=True, **kw):
  t = np.asarray(tile)
  nd = arr.ndim
  offset = (0,) * nd if offset is None else _shape_t(offset)
  dst = arr._a
  sl_src, sl_dst = [], []
  for k, o in enumerate(offset):
    n = t.shape[k]
    hi = min(dst.shape[k], o + n)
    if hi <= o:
      return
    sl_dst.append(slice(o, hi))
    sl_src.append(slice(0, hi - o))
  dst[tuple(sl_dst)] = t[tuple(sl_src)]


def _cholesky(a):
  """Lower Cholesky factor by the textbook column recurrence (tolerates the zero padding of nv_pad tiles)."""
  a = np.array(a, dtype=np.float64)
  n = a.shape[0]
  L = np.zeros_like(a)
  with np.errstate(all="ignore"):
    for j in range(n):
      s = a[j, j] - np.dot(L[j, :j], L[j, :j])
      L[j, j] = np.sqrt(s)
      for i in range(j + 1, n):
        L[i, j] = (a[i, j] - np.dot(L[i, :j], L[j, :j])) / L[j, j]
  return _tile(L)


def _lower_solve(L, b):
  L, x = np.asarray(L), np.array(b, dtype=np.float64)
  n = L.shape[0]
  with np.errstate(all="ignore"):
    for i in range(n):
      x[i] = (x[i] - L[i, :i] @ x[:i]) / L[i, i]
  return x


def _upper_solve(U, b):
  U, x = np.asarray(U), np.array(b, dtype=np.float64)
  n = U.shape[0]
  with np.errstate(all="ignore"):
    for i in range(n - 1, -1, -1):
      x[i] = (x[i] - U[i, i + 1 :] @ x[i + 1 :]) / U[i, i]
  return x


def _tile_cholesky_solve(L, y):
  z = _lower_solve(L, y)
  return _tile(_upper_solve(np.asarray(L).T, z))


def _tile_lower_solve_inplace(L, x):
  np.asarray(x)[...] = _lower_solve(L, x)


def _tile_upper_solve_inplace(U, x):
  np.asarray(x)[...] = _upper_solve(U, x)


def _tile_matmul(a, b, out=None, alpha=1.0, beta=1.0):
  p = np.asarray(a) @ np.asarray(b)
  if out is None:
    return _tile(p)
  o = np.asarray(out)
  o[...] = beta * o + alpha * p
  return None


def _tile_map(fn, *tiles):
  arrs = [np.asarray(t) for t in tiles]
  out = np.empty(arrs[0].shape, dtype=object)
  for idx in np.ndindex(arrs[0].shape):
    out[idx] = fn(*[_wrap(a[idx]) for a in arrs])
  k = "i" if all(isinstance(v, int) and not isinstance(v, bool) for v in out.reshape(-1)) and out.size else "f"
  return _tile(out.astype(np.int64 if k == "i" else np.float64))


def _tile_reduce(op, t):
  a = np.asarray(t)
  et = getattr(t, "_etype", None)
  if et is not None:
    es = len(et._shape)
    flat = a.reshape((-1,) + tuple(et._shape))
    acc = flat[0].copy()
    for r in flat[1:]:
      acc = np.asarray(op(_wrap(acc), _wrap(r)))
    return _tile(acc.reshape((1,) + tuple(et._shape)), et)
  flat = a.reshape(-1)
  acc = _wrap(flat[0])
  for r in flat[1:]:
    acc = op(acc, _wrap(r))
  return _tile(np.asarray([acc]))


def _tile_from_value(x, preserve_type=False):
  if isinstance(x, _Arr):
    if preserve_type:
      return _tile(np.asarray(x).reshape((1,) + x.shape), type(x))
    return _tile(np.asarray(x).reshape((1,) + x.shape))
  return _tile(np.asarray([x]))


def _tile_sort(keys, values):
  k, v = np.asarray(keys), np.asarray(values)
  order = np.argsort(k, kind="stable")
  k[...] = k[order]
  v[...] = v[order]


def _tile_ctor(fillv):
  def f(shape, dtype=float, storage=None, **kw):
    return _tile(np.full(_shape_t(shape), fillv, dtype=_np_dtype(dtype)))

  return f


def _tile_view(t, offset, shape=None, **kw):
  offset = _shape_t(offset)
  a = np.asarray(t)
  shape = _shape_t(shape) if shape is not None else None
  sl = []
  for k, o in enumerate(offset):
    n = shape[k] if shape is not None else a.shape[k] - o
    sl.append(slice(o, o + n))
  return a[tuple(sl)].view(Tile)


def _tile_transpose(t):
  """A VIEW of the tile: block_cholesky.py:60 solves in place into `wp.tile_transpose(A_ik_tile)` and then stores A_ik_tile."""
  v = np.asarray(t).T.view(Tile)
  v._etype = getattr(t, "_etype", None)
  return v


def _tile_assign(dst, src, offset):
  offset = _shape_t(offset)
  d, s = np.asarray(dst), np.asarray(src)
  sl = tuple(slice(o, o + n) for o, n in zip(offset, s.shape))
  d[sl] = s


def _tile_diag_add(A, d):
  a = np.array(A, dtype=np.float64)
  a[np.diag_indices(a.shape[0])] += np.asarray(d)
  return _tile(a)


# ------------------------------------------------------------------------------------------------
# scalar / vector builtins
# ------------------------------------------------------------------------------------------------


def _minmax(fn, pyfn):
  def f(a, b=None):
    if b is None:
      return _wrap(fn.reduce(np.asarray(a)))
    if np.ndim(a) == 0 and np.ndim(b) == 0:
      if _is_intlike(a) and _is_intlike(b):
        return wint(pyfn(int(a), int(b)))
      return _F(fn(_F(a), _F(b)))  # numpy propagates NaN like fmin/fmax do not; inputs are finite
    return _wrap(fn(np.asarray(a), np.asarray(b)), _is_intlike(np.asarray(a)) and _is_intlike(np.asarray(b)))

  return f


def _normalize(v):
  n = float(np.sqrt(np.dot(v, v)))
  return _wrap(np.asarray(v) / n) if n > 0.0 else _wrap(np.zeros_like(np.asarray(v), dtype=np.float64))


def _f1(fn):
  def f(v):
    with np.errstate(all="ignore"):
      return _wrap(fn(np.asarray(v, dtype=np.float64))) if np.ndim(v) else _F(fn(_F(v)))

  return f


def _abs(v):
  if np.ndim(v):
    return _wrap(np.abs(np.asarray(v)), _is_intlike(np.asarray(v)))
  return wint(abs(int(v))) if _is_intlike(v) else _F(abs(v))


def _where(c, a, b):
  return a if c else b


def _struct(cls):
  ann = dict(getattr(cls, "__annotations__", {}))

  def __init__(self):
    for k, t in ann.items():
      if isinstance(t, type) and issubclass(t, _Arr):
        v = t()
      elif t in (bool,):
        v = False
      elif _is_int_dtype(t):
        v = wint(0)
      elif t is float:
        v = 0.0
      elif isinstance(t, type) and hasattr(t, "_wp_struct"):
        v = t()
      else:
        v = None
      object.__setattr__(self, k, v)

  cls.__init__ = __init__
  cls._wp_struct = True
  return cls


class _Device:
  is_cuda = False
  is_cpu = True
  alias = "cpu"
  sm_count = 1

  def __str__(self):
    return "cpu"


class _Scoped:
  def __init__(self, *a, **k):
    pass

  def __enter__(self):
    return self

  def __exit__(self, *a):
    return False


def _capture_while(cond, while_body, **kwargs):
  while int(cond.numpy().reshape(-1)[0]) != 0:
    while_body(**kwargs)


def _array_scan(src, dst, inclusive=True):
  c = np.cumsum(src._a.reshape(-1))
  if not inclusive:
    c = np.concatenate(([0], c[:-1]))
  dst._a.reshape(-1)[...] = c


def _segmented_sort_pairs(keys, values, count, segment_start_indices, segment_end_indices=None):
  k, v = keys._a.reshape(-1), values._a.reshape(-1)
  st = segment_start_indices._a.reshape(-1)
  nseg = len(st) if segment_end_indices is not None else len(st) - 1
  for s in range(nseg):
    lo = int(st[s])
    hi = int(segment_end_indices._a.reshape(-1)[s]) if segment_end_indices is not None else int(st[s + 1])
    order = np.argsort(k[lo:hi], kind="stable")
    k[lo:hi] = k[lo:hi][order]
    v[lo:hi] = v[lo:hi][order]


def build():
  wp = _types.ModuleType("warp")
  wp.func, wp.kernel, wp.func_native, wp.struct = _func, _kernel, _func_native, _struct
  wp.Kernel = Kernel
  wp.static = lambda x: x
  wp.constant = lambda x: x
  wp.set_module_options = lambda *a, **k: None
  wp.init = lambda *a, **k: None
  wp.config = _types.SimpleNamespace(enable_mathdx_gemm=False, quiet=True, verbose=False)
  wp.float32 = wp.float64 = wp.float = _F

  class _I32(metaclass=_IntType):
    _wp_int = True

  class _U32(metaclass=_IntType):
    _wp_int = True
    _wp_uint = True

  wp.int32 = wp.int64 = wp.int = wp.int8 = wp.int16 = _I32
  wp.uint32 = wp.uint64 = wp.uint8 = wp.uint16 = _U32
  wp.bool = bool
  wp.inf = float("inf")
  wp.pi = _math.pi
  wp.types = _types.SimpleNamespace(vector=_vector, matrix=_matrix)
  wp.vec = lambda length, dtype=float: _vector(length, dtype)
  wp.mat = lambda shape, dtype=float: _matrix(shape, dtype)
  wp.vector, wp.matrix = _vector, _matrix
  wp.vec2, wp.vec3, wp.vec4 = _vector(2), _vector(3), _vector(4)
  wp.vec2f, wp.vec3f, wp.vec4f = wp.vec2, wp.vec3, wp.vec4
  wp.vec2i, wp.vec3i, wp.vec4i = _vector(2, int), _vector(3, int), _vector(4, int)
  wp.quat = wp.quatf = wp.vec4
  wp.mat22, wp.mat33, wp.mat44 = _matrix((2, 2)), _matrix((3, 3)), _matrix((4, 4))
  wp.mat22f, wp.mat33f, wp.mat44f = wp.mat22, wp.mat33, wp.mat44
  wp.spatial_vector = wp.spatial_vectorf = _vector(6)
  wp.spatial_top = lambda v: _wrap(np.asarray(v)[:3].copy())
  wp.spatial_bottom = lambda v: _wrap(np.asarray(v)[3:].copy())
  wp.where = _where
  wp.abs = _abs
  wp.sqrt = _f1(np.sqrt)
  wp.sin, wp.cos, wp.tan = _f1(np.sin), _f1(np.cos), _f1(np.tan)
  wp.asin, wp.acos, wp.atan = _f1(np.arcsin), _f1(np.arccos), _f1(np.arctan)
  wp.atan2 = lambda a, b: _F(np.arctan2(a, b))
  wp.exp, wp.log = _f1(np.exp), _f1(np.log)
  wp.ceil, wp.floor, wp.round = _f1(np.ceil), _f1(np.floor), (lambda v: _F(np.sign(v) * np.floor(abs(v) + 0.5)))
  wp.pow = lambda a, b: _F(np.power(float(a), float(b)))
  wp.sign = lambda v: _wrap(np.where(np.asarray(v) < 0, -1.0, 1.0)) if np.ndim(v) else (_F(-1.0) if v < 0 else _F(1.0))
  wp.min, wp.max = _minmax(np.minimum, min), _minmax(np.maximum, max)
  wp.clamp = lambda v, lo, hi: _wrap(np.clip(v, lo, hi), _is_intlike(v) and _is_intlike(lo))
  wp.dot = lambda a, b: _F(np.dot(np.asarray(a), np.asarray(b)))
  wp.cross = lambda a, b: _wrap(np.cross(np.asarray(a), np.asarray(b)))
  wp.length = lambda v: _F(np.sqrt(np.dot(np.asarray(v), np.asarray(v))))
  wp.length_sq = lambda v: _F(np.dot(np.asarray(v), np.asarray(v)))
  wp.norm_l2 = wp.length
  wp.normalize = _normalize
  wp.transpose = lambda m: _wrap(np.asarray(m).T.copy())
  wp.cw_mul = lambda a, b: _wrap(np.asarray(a) * np.asarray(b))
  wp.cw_div = lambda a, b: _wrap(np.asarray(a) / np.asarray(b))
  wp.outer = lambda a, b: _wrap(np.outer(a, b))
  wp.diag = lambda v: _wrap(np.diag(np.asarray(v)))
  wp.trace = lambda m: _F(np.trace(np.asarray(m)))
  wp.determinant = lambda m: _F(np.linalg.det(np.asarray(m)))
  wp.inverse = lambda m: _wrap(np.linalg.inv(np.asarray(m)))
  wp.identity = lambda n, dtype=float: _wrap(np.eye(n))
  wp.matrix_from_rows = lambda *rows: _wrap(np.stack([np.asarray(r, dtype=np.float64) for r in rows]))
  wp.matrix_from_cols = lambda *cols: _wrap(np.stack([np.asarray(c, dtype=np.float64) for c in cols], axis=1))
  wp.isnan = lambda v: bool(np.any(np.isnan(v)))
  wp.isinf = lambda v: bool(np.any(np.isinf(v)))
  wp.isfinite = lambda v: bool(np.all(np.isfinite(v)))
  wp.add = lambda a, b: a + b
  wp.sub = lambda a, b: a - b
  wp.mul = lambda a, b: a * b
  wp.printf = lambda *a, **k: None
  wp.print = lambda *a, **k: None

  wp.array = warray
  wp.array1d, wp.array2d, wp.array3d, wp.array4d = _arr_cls(1), _arr_cls(2), _arr_cls(3), _arr_cls(4)
  wp.zeros = _full_like_ctor(0)
  wp.empty = _full_like_ctor(None)
  wp.ones = _full_like_ctor(1)
  wp.full = _full
  wp.zeros_like = lambda a, **k: warray._view(np.zeros_like(a._a), a.dtype)
  wp.empty_like = wp.zeros_like
  wp.ones_like = lambda a, **k: warray._view(np.ones_like(a._a), a.dtype)
  wp.clone = _clone
  wp.copy = _copy
  wp.from_numpy = lambda a, dtype=None, **k: warray(a, dtype=dtype)
  wp.launch, wp.launch_tiled = _launch, _launch_tiled
  wp.tid = _tid
  wp.block_dim = lambda: wint(_STATE["block_dim"])
  wp.atomic_add = _atomic(lambda a, b: a + b)
  wp.atomic_sub = _atomic(lambda a, b: a - b)
  wp.atomic_max = _atomic(lambda a, b: max(a, b))
  wp.atomic_min = _atomic(lambda a, b: min(a, b))

  wp.tile = _tile_from_value
  wp.tile_load, wp.tile_store = _tile_load, _tile_store
  wp.tile_cholesky, wp.tile_cholesky_solve = _cholesky, _tile_cholesky_solve
  wp.tile_lower_solve_inplace, wp.tile_upper_solve_inplace = _tile_lower_solve_inplace, _tile_upper_solve_inplace
  wp.tile_matmul = _tile_matmul
  wp.tile_transpose = _tile_transpose
  wp.tile_map, wp.tile_reduce = _tile_map, _tile_reduce
  wp.tile_sum = lambda t: _tile(np.asarray([np.asarray(t).sum()]))
  wp.tile_zeros, wp.tile_ones = _tile_ctor(0), _tile_ctor(1)
  wp.tile_arange = lambda *a, dtype=int, **k: _tile(np.arange(*[int(x) for x in a], dtype=_np_dtype(dtype)))
  wp.tile_broadcast = lambda t, shape: _tile(np.broadcast_to(np.asarray(t), _shape_t(shape)).copy())
  wp.tile_view, wp.tile_assign, wp.tile_diag_add, wp.tile_sort = _tile_view, _tile_assign, _tile_diag_add, _tile_sort

  wp.get_device = lambda *a, **k: _Device()
  wp.get_preferred_device = wp.get_device
  wp.ScopedDevice = wp.ScopedStream = wp.ScopedCapture = wp.ScopedTimer = _Scoped
  wp.is_conditional_graph_supported = lambda: False
  wp.capture_while = _capture_while
  wp.synchronize = lambda *a, **k: None
  wp.Event = lambda *a, **k: None
  wp.record_event = lambda *a, **k: None
  wp.get_event_elapsed_time = lambda *a, **k: 0.0
  wp.Texture = wp.Bvh = wp.Mesh = type("Opaque", (), {})
  wp.Texture2D = wp.Texture
  wp.utils = _types.SimpleNamespace(array_scan=_array_scan, segmented_sort_pairs=_segmented_sort_pairs)
  wp.context = _types.SimpleNamespace(Device=_Device, runtime=None)
  wp.__version__ = "0.0-numpy-interpreter"
  wp._is_shim = True
  wp._state = _STATE
  return wp


def install():
  """Registers the interpreter as `warp` for the duration of a generator script."""
  if "warp" in sys.modules and not getattr(sys.modules["warp"], "_is_shim", False):
    raise RuntimeError("a real warp is importable: use it instead of the interpreter")
  wp = build()
  np.seterr(all="ignore")
  sys.modules["warp"] = wp
  sys.modules["warp.types"] = wp.types
  sys.modules["warp.utils"] = wp.utils
  return wp
