"""A numpy stand-in for the small part of the `warp` API that the reference's *pure* `@wp.func` geometry / math
functions use, so that those functions can be executed here (no GPU, no warp-lang wheel) exactly as written, from
their files under /root/reference, to produce golden vectors for the oracle.

Only used by the golden-vector generators in this directory; nothing in the product or in the tests imports it.

Semantics reproduced:
  * vectors / matrices are VALUES in Warp: `a = M[i]` copies a row, `a += b` rebinds `a` to a new value (it never
    mutates another name bound to the old value).  `_Arr.__getitem__` therefore returns copies and the augmented
    operators return new objects.
  * `wp.func` is the identity decorator; `wp.static(x)` is `x`; scalar constructors are Python scalars.
  * arithmetic runs in float64: the vectors pin the ALGORITHM (the float64 oracle must agree to ~1e-9); float32
    rounding of the real reference is a separate matter covered by the device-vs-oracle tolerances.
"""

import sys
import types as _types

import numpy as np


class _Arr(np.ndarray):
  _shape = None

  def __new__(cls, *args):
    shape = cls._shape
    a = np.zeros(shape, dtype=np.float64).view(cls)
    if len(args) == 1 and np.ndim(args[0]) == 0:
      a[...] = float(args[0])
    elif len(args) == 1:
      a[...] = np.asarray(args[0], dtype=np.float64).reshape(shape)
    elif len(args) > 1:
      flat = []
      for x in args:  # scalars, or rows given as vectors
        flat.extend(np.asarray(x, dtype=np.float64).reshape(-1).tolist())
      a[...] = np.asarray(flat).reshape(shape)
    return a

  def __array_finalize__(self, obj):
    pass

  def __getitem__(self, idx):
    r = np.ndarray.__getitem__(self, idx)
    if isinstance(r, np.ndarray) and r.ndim > 0:
      return r.copy()
    return float(r)

  def __setitem__(self, idx, val):
    # on the plain-ndarray view: numpy's subclass path would go through the copying __getitem__ above
    np.ndarray.__setitem__(self.view(np.ndarray), idx, val)

  def __mul__(self, o):
    # Warp: matrix * vector and matrix * matrix are products, everything else is element-wise / scalar
    if self.ndim == 2 and isinstance(o, np.ndarray) and o.ndim >= 1:
      return _wrap(np.asarray(self) @ np.asarray(o))
    return _wrap(np.multiply(np.asarray(self), np.asarray(o)))

  # value semantics for augmented assignment
  def __iadd__(self, o):
    return self + o

  def __isub__(self, o):
    return self - o

  def __imul__(self, o):
    return self * o

  def __itruediv__(self, o):
    return self / o

  x = property(lambda s: float(np.ndarray.__getitem__(s, 0)))
  y = property(lambda s: float(np.ndarray.__getitem__(s, 1)))
  z = property(lambda s: float(np.ndarray.__getitem__(s, 2)))
  w = property(lambda s: float(np.ndarray.__getitem__(s, 3)))


def _make(shape, name):
  return type(name, (_Arr,), {"_shape": tuple(shape)})


def _vector(length, dtype=None):
  return _make((length,), f"vec{length}")


def _matrix(shape, dtype=None):
  return _make(shape, f"mat{shape[0]}{shape[1]}")


def _wrap(v):
  return v.view(_Arr) if isinstance(v, np.ndarray) and v.ndim > 0 else float(v)


def _normalize(v):
  n = float(np.sqrt(np.dot(v, v)))
  return _wrap(np.asarray(v) / n) if n > 0.0 else _wrap(np.zeros_like(np.asarray(v)))


def _minmax(fn):
  def f(a, b=None):
    if b is None:
      return float(fn.reduce(np.asarray(a)))
    return _wrap(fn(np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)))

  return f


def build():
  wp = _types.ModuleType("warp")
  wp.func = lambda f=None, **kw: f if f is not None else (lambda g: g)
  wp.kernel = wp.func
  wp.static = lambda x: x
  wp.set_module_options = lambda *a, **k: None
  wp.float32 = wp.float64 = wp.float = float
  wp.int32 = wp.int64 = wp.int = int
  wp.bool = bool
  wp.inf = float("inf")
  wp.types = _types.SimpleNamespace(vector=_vector, matrix=_matrix)
  wp.vec2, wp.vec3, wp.vec4 = _vector(2), _vector(3), _vector(4)
  wp.vec2i = wp.vec2
  wp.quat = _vector(4)
  wp.mat22, wp.mat33, wp.mat44 = _matrix((2, 2)), _matrix((3, 3)), _matrix((4, 4))
  wp.spatial_vector = _vector(6)
  wp.where = lambda c, a, b: a if c else b
  wp.abs = lambda v: _wrap(np.abs(v))
  wp.sqrt = lambda v: float(np.sqrt(v))
  wp.sin, wp.cos, wp.acos, wp.atan2 = (lambda v: float(np.sin(v))), (lambda v: float(np.cos(v))), (lambda v: float(np.arccos(v))), (lambda a, b: float(np.arctan2(a, b)))
  wp.exp, wp.log, wp.pow = (lambda v: float(np.exp(v))), (lambda v: float(np.log(v))), (lambda a, b: float(np.power(a, b)))
  wp.sign = lambda v: -1.0 if v < 0 else 1.0
  wp.min, wp.max = _minmax(np.minimum), _minmax(np.maximum)
  wp.clamp = lambda v, lo, hi: _wrap(np.clip(v, lo, hi))
  wp.dot = lambda a, b: float(np.dot(a, b))
  wp.cross = lambda a, b: _wrap(np.cross(a, b))
  wp.length = lambda v: float(np.sqrt(np.dot(v, v)))
  wp.length_sq = lambda v: float(np.dot(v, v))
  wp.normalize = _normalize
  wp.transpose = lambda m: _wrap(np.asarray(m).T.copy())
  wp.cw_mul = lambda a, b: _wrap(np.asarray(a) * np.asarray(b))
  wp.outer = lambda a, b: _wrap(np.outer(a, b))
  wp.matrix_from_rows = lambda *rows: _wrap(np.stack([np.asarray(r, dtype=np.float64) for r in rows]))
  wp.matrix_from_cols = lambda *cols: _wrap(np.stack([np.asarray(c, dtype=np.float64) for c in cols], axis=1))
  wp.isnan = lambda v: bool(np.isnan(v))
  wp.array = wp.array2d = wp.array3d = _types.SimpleNamespace  # annotations only
  return wp


def install():
  """Registers the stand-in as `warp` for the duration of a generator script."""
  if "warp" in sys.modules and not getattr(sys.modules["warp"], "_is_shim", False):
    raise RuntimeError("a real warp is importable: use it instead of the shim")
  wp = build()
  wp._is_shim = True
  sys.modules["warp"] = wp
  return wp
