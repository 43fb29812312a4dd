"""Minimal MJCF -> MjModel-like compiler (host, numpy float64).

The reference obtains its host model from MuJoCo C (`mujoco.MjSpec.from_file(...).compile()`,
reference mujoco_warp/testspeed.py:82-99) and `put_model` only reads attributes off that object
(reference mujoco_warp/_src/io.py:77-647).  MuJoCo is not installed in this environment, so this
module provides the subset of the MJCF compiler needed by the benchmark models: defaults/classes,
bodies/joints/geoms/sites, primitive-geom inertia, actuators (motor/position/velocity/general),
equality(joint/connect/weld), contact pair/exclude, keyframes, options/flags, plus the
`mj_setConst`-style derived constants (invweight0, meaninertia, subtreemass, ...).

The result, `MjModelLite`, exposes MuJoCo's `MjModel` attribute names so that `put_model` can treat
it exactly like a `mujoco.MjModel` (duck typing).  If a real `mujoco.MjModel` is available it can be
passed to `put_model` instead.

Compiler-derived constants are produced from MuJoCo's documented conventions; they are NOT pinned
against MuJoCo C in this container ("parity unpinned" for these constants).
"""

from __future__ import annotations

import math
import os
import xml.etree.ElementTree as ET
from types import SimpleNamespace

import numpy as np

from . import constants as C

mjMINVAL = 1e-15

_GEOM_TYPES = {
  "plane": C.GeomType.PLANE,
  "hfield": C.GeomType.HFIELD,
  "sphere": C.GeomType.SPHERE,
  "capsule": C.GeomType.CAPSULE,
  "ellipsoid": C.GeomType.ELLIPSOID,
  "cylinder": C.GeomType.CYLINDER,
  "box": C.GeomType.BOX,
  "mesh": C.GeomType.MESH,
}
_JNT_TYPES = {"free": C.JointType.FREE, "ball": C.JointType.BALL, "slide": C.JointType.SLIDE, "hinge": C.JointType.HINGE}


# ----------------------------------------------------------------------------- small math helpers
def _vec(s, n=None, default=None):
  if s is None:
    return None if default is None else np.array(default, dtype=np.float64)
  v = np.array([float(x) for x in s.split()], dtype=np.float64)
  if n is not None and v.size != n:
    raise ValueError(f"expected {n} values, got {v.size}: {s!r}")
  return v


def quat_mul(u, v):
  return np.array(
    [
      u[0] * v[0] - u[1] * v[1] - u[2] * v[2] - u[3] * v[3],
      u[0] * v[1] + u[1] * v[0] + u[2] * v[3] - u[3] * v[2],
      u[0] * v[2] - u[1] * v[3] + u[2] * v[0] + u[3] * v[1],
      u[0] * v[3] + u[1] * v[2] - u[2] * v[1] + u[3] * v[0],
    ]
  )


def quat_to_mat(q):
  w, x, y, z = q
  return np.array(
    [
      [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
      [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
      [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z],
    ]
  )


def rot_vec_quat(v, q):
  return quat_to_mat(q) @ v


def mat_to_quat(m):
  """Rotation matrix -> unit quaternion (w,x,y,z)."""
  tr = m[0, 0] + m[1, 1] + m[2, 2]
  if tr > 0:
    s = math.sqrt(tr + 1.0) * 2
    q = [0.25 * s, (m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s]
  elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
    s = math.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
    q = [(m[2, 1] - m[1, 2]) / s, 0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s]
  elif m[1, 1] > m[2, 2]:
    s = math.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
    q = [(m[0, 2] - m[2, 0]) / s, (m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s]
  else:
    s = math.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
    q = [(m[1, 0] - m[0, 1]) / s, (m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s]
  q = np.array(q)
  return q / np.linalg.norm(q)


def z2quat(vec):
  """Quaternion rotating the z-axis onto vec (MuJoCo mjuu_z2quat convention)."""
  vec = np.asarray(vec, dtype=np.float64)
  n = np.linalg.norm(vec)
  if n < mjMINVAL:
    return np.array([1.0, 0, 0, 0])
  vec = vec / n
  axis = np.cross([0.0, 0.0, 1.0], vec)
  s = np.linalg.norm(axis)
  if s < 1e-10:
    axis = np.array([1.0, 0.0, 0.0])
  else:
    axis = axis / s
  ang = math.atan2(s, vec[2])
  q = np.concatenate([[math.cos(ang / 2)], axis * math.sin(ang / 2)])
  return q / np.linalg.norm(q)


def eig3(mat):
  """Eigen-decomposition of a symmetric 3x3 by quaternion Jacobi iteration.

  Returns (eigval desc, quat) with mat = R diag(eigval) R^T, R = quat_to_mat(quat).
  Follows the structure of MuJoCo's documented mju_eig3 (Jacobi rotations accumulated in a
  quaternion, eigenvalues sorted in decreasing order).
  """
  eps = 1e-12
  quat = np.array([1.0, 0, 0, 0])
  eigval = np.zeros(3)
  for _ in range(500):
    R = quat_to_mat(quat)
    D = R.T @ mat @ R
    eigval = np.array([D[0, 0], D[1, 1], D[2, 2]])
    if abs(D[0, 1]) > abs(D[0, 2]) and abs(D[0, 1]) > abs(D[1, 2]):
      rk, ck, rotk = 0, 1, 2
    elif abs(D[0, 2]) > abs(D[1, 2]):
      rk, ck, rotk = 0, 2, 1
    else:
      rk, ck, rotk = 1, 2, 0
    if abs(D[rk, ck]) < eps:
      break
    tau = (D[ck, ck] - D[rk, rk]) / (2 * D[rk, ck])
    if tau >= 0:
      t = 1.0 / (tau + math.sqrt(1 + tau * tau))
    else:
      t = -1.0 / (-tau + math.sqrt(1 + tau * tau))
    c = 1.0 / math.sqrt(1 + t * t)
    if c > 1.0 - eps:
      break
    tmp = np.zeros(4)
    tmp[rotk + 1] = -math.sqrt(0.5 - 0.5 * c) if tau >= 0 else math.sqrt(0.5 - 0.5 * c)
    if rotk == 1:
      tmp[rotk + 1] = -tmp[rotk + 1]
    tmp[0] = math.sqrt(1.0 - tmp[rotk + 1] ** 2)
    tmp /= np.linalg.norm(tmp)
    quat = quat_mul(quat, tmp)
    quat /= np.linalg.norm(quat)
  # sort decreasing (bubble: 0,1,0), rotating the quaternion accordingly
  for j in range(3):
    j1 = j % 2
    if eigval[j1] < eigval[j1 + 1]:
      eigval[j1], eigval[j1 + 1] = eigval[j1 + 1], eigval[j1]
      tmp = np.zeros(4)
      tmp[0] = 0.707106781186548
      tmp[(j1 + 2) % 3 + 1] = tmp[0]
      quat = quat_mul(quat, tmp)
      quat /= np.linalg.norm(quat)
  return eigval, quat


# ----------------------------------------------------------------------------- defaults handling
_ACT_TAGS = ("general", "motor", "position", "velocity", "intvelocity", "damper", "cylinder", "muscle", "adhesion")


class _Defaults:
  """Default-class tree: class name -> {tag -> attrib dict}, with inheritance from the parent."""

  def __init__(self):
    self.classes = {"main": {}}
    self.parent = {"main": None}

  def parse(self, elem, parent=None):
    name = elem.get("class", "main" if parent is None else None)
    if name is None:
      raise ValueError("nested <default> requires a class name")
    if name not in self.classes:
      self.classes[name] = {}
    if name != "main" or parent is not None:
      self.parent[name] = parent
    # inherit from parent
    if parent is not None:
      for tag, attrs in self.classes[parent].items():
        merged = dict(attrs)
        merged.update(self.classes[name].get(tag, {}))
        self.classes[name][tag] = merged
    for child in elem:
      if child.tag == "default":
        continue
      d = self.classes[name].setdefault(child.tag, {})
      d.update(child.attrib)
    for child in elem:
      if child.tag == "default":
        self.parse(child, name)

  def get(self, cls, tag):
    return dict(self.classes.get(cls or "main", {}).get(tag, {}))


def _resolve(defaults, elem, childclass, tags=None):
  """Effective attributes of elem: class defaults overlaid with explicit attributes."""
  cls = elem.get("class", childclass or "main")
  if cls not in defaults.classes:
    raise ValueError(f"unknown default class {cls!r}")
  attrs = {}
  for tag in tags or (elem.tag,):
    attrs.update(defaults.get(cls, tag))
  explicit = dict(elem.attrib)
  # an explicit orientation/placement spec overrides a defaulted alternative one
  orient = ("quat", "axisangle", "euler", "xyaxes", "zaxis")
  if any(k in explicit for k in orient):
    for k in orient:
      attrs.pop(k, None)
  if "fromto" in explicit:
    attrs.pop("pos", None)
    for k in orient:
      attrs.pop(k, None)
  attrs.update(explicit)
  return attrs


def _orientation(attrs, comp):
  """Parse quat / axisangle / euler / xyaxes / zaxis into a unit quaternion."""
  if "quat" in attrs:
    q = _vec(attrs["quat"], 4)
    return q / np.linalg.norm(q)
  if "axisangle" in attrs:
    aa = _vec(attrs["axisangle"], 4)
    ang = aa[3] * (math.pi / 180.0 if comp.degree else 1.0)
    ax = aa[:3] / np.linalg.norm(aa[:3])
    return np.concatenate([[math.cos(ang / 2)], ax * math.sin(ang / 2)])
  if "euler" in attrs:
    e = _vec(attrs["euler"], 3) * (math.pi / 180.0 if comp.degree else 1.0)
    q = np.array([1.0, 0, 0, 0])
    for i, ch in enumerate(comp.eulerseq):
      ax = np.zeros(3)
      ax["xyz".index(ch.lower())] = 1.0
      qi = np.concatenate([[math.cos(e[i] / 2)], ax * math.sin(e[i] / 2)])
      # lower-case: intrinsic (rotating frame) -> post-multiply; upper-case: extrinsic
      q = quat_mul(q, qi) if ch.islower() else quat_mul(qi, q)
    return q / np.linalg.norm(q)
  if "xyaxes" in attrs:
    v = _vec(attrs["xyaxes"], 6)
    x = v[:3] / np.linalg.norm(v[:3])
    y = v[3:] - x * np.dot(x, v[3:])
    y = y / np.linalg.norm(y)
    z = np.cross(x, y)
    return mat_to_quat(np.stack([x, y, z], axis=1))
  if "zaxis" in attrs:
    return z2quat(_vec(attrs["zaxis"], 3))
  return np.array([1.0, 0, 0, 0])


# ----------------------------------------------------------------------------- geom mass properties
def _geom_volume_inertia(gtype, size):
  """Volume and unit-density inertia diagonal (about the geom centre, geom frame)."""
  if gtype == C.GeomType.SPHERE:
    r = size[0]
    vol = 4.0 / 3.0 * math.pi * r**3
    i = 0.4 * r * r
    return vol, np.array([i, i, i]) * vol
  if gtype == C.GeomType.CAPSULE:
    r, h2 = size[0], 2 * size[1]
    vol_c = math.pi * r * r * h2
    vol_s = 4.0 / 3.0 * math.pi * r**3
    vol = vol_c + vol_s
    ix = vol_c * (3 * r * r + h2 * h2) / 12.0
    iz = vol_c * r * r / 2.0
    sph = 2.0 * vol_s * r * r / 5.0
    ix += sph + vol_s * h2 * (3 * r + 2 * h2) / 8.0
    iz += sph
    return vol, np.array([ix, ix, iz])
  if gtype == C.GeomType.CYLINDER:
    r, h2 = size[0], 2 * size[1]
    vol = math.pi * r * r * h2
    ix = vol * (3 * r * r + h2 * h2) / 12.0
    return vol, np.array([ix, ix, vol * r * r / 2.0])
  if gtype == C.GeomType.ELLIPSOID:
    a, b, c = size
    vol = 4.0 / 3.0 * math.pi * a * b * c
    return vol, vol / 5.0 * np.array([b * b + c * c, a * a + c * c, a * a + b * b])
  if gtype == C.GeomType.BOX:
    a, b, c = size
    vol = 8.0 * a * b * c
    return vol, vol / 3.0 * np.array([b * b + c * c, a * a + c * c, a * a + b * b])
  return 0.0, np.zeros(3)


def _geom_bounds(gtype, size):
  """(rbound, aabb[6]=center,halfsize) for a primitive geom."""
  big = C.MJ_MAXVAL
  if gtype == C.GeomType.PLANE:
    return 0.0, np.array([0, 0, 0, big, big, big], dtype=np.float64)
  if gtype == C.GeomType.SPHERE:
    r = size[0]
    return r, np.array([0, 0, 0, r, r, r])
  if gtype == C.GeomType.CAPSULE:
    r, h = size[0], size[1]
    return r + h, np.array([0, 0, 0, r, r, r + h])
  if gtype == C.GeomType.CYLINDER:
    r, h = size[0], size[1]
    return math.sqrt(r * r + h * h), np.array([0, 0, 0, r, r, h])
  if gtype == C.GeomType.ELLIPSOID:
    return float(max(size)), np.array([0, 0, 0, size[0], size[1], size[2]])
  if gtype == C.GeomType.BOX:
    return float(np.linalg.norm(size)), np.array([0, 0, 0, size[0], size[1], size[2]])
  raise NotImplementedError(f"geom type {gtype} bounds")


# ----------------------------------------------------------------------------- mesh assets
def _load_mesh_vertices(path):
  """Vertex positions of a binary / ASCII STL or an OBJ file (faces are not needed: collision uses the convex hull)."""
  ext = os.path.splitext(path)[1].lower()
  with open(path, "rb") as f:
    raw = f.read()
  if ext == ".obj":
    rows = [ln.split()[1:4] for ln in raw.decode("utf-8", "replace").splitlines() if ln.startswith("v ")]
    return np.array(rows, dtype=np.float64).reshape(-1, 3)
  if ext == ".stl":
    ntri = int(np.frombuffer(raw[80:84], dtype="<u4")[0]) if len(raw) >= 84 else 0
    if len(raw) == 84 + 50 * ntri:  # binary: 80-byte header, count, 50-byte records (normal, 3 vertices, attribute)
      rec = np.frombuffer(raw[84:], dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", 9), ("a", "<u2")]))
      return rec["v"].reshape(-1, 3).astype(np.float64)
    rows = [ln.split()[1:4] for ln in raw.decode("utf-8", "replace").splitlines() if ln.strip().startswith("vertex")]
    return np.array(rows, dtype=np.float64).reshape(-1, 3)
  raise NotImplementedError(f"mesh file format {ext}")


def _convex_hull_vertices(v):
  """Vertices of the convex hull (in first-appearance order).  MuJoCo collides meshes through their convex hulls; the support
  function over the hull vertices equals the support function over all vertices."""
  v = np.asarray(v, dtype=np.float64).reshape(-1, 3)
  _, first = np.unique(np.round(v, 12), axis=0, return_index=True)
  v = v[np.sort(first)]
  if len(v) < 4:
    return v
  try:
    from scipy.spatial import ConvexHull

    return v[np.sort(ConvexHull(v).vertices)]
  except Exception:  # degenerate (flat) point sets: keep all points
    return v


def _hull_graph(verts):
  """MuJoCo-layout convex graph of hull vertices `verts` (already reduced to the hull); None for degenerate sets."""
  v = np.asarray(verts, dtype=np.float64).reshape(-1, 3)
  if len(v) < 4:
    return None
  try:
    from scipy.spatial import ConvexHull

    hull = ConvexHull(v)
  except Exception:
    return None
  if len(hull.vertices) != len(v):  # a vertex inside a facet: not every point is a graph node; keep the exhaustive search
    return None
  nvert, nface = len(v), len(hull.simplices)
  nbr = [[] for _ in range(nvert)]
  for a, b, c in hull.simplices:
    for x, y in ((a, b), (b, c), (c, a), (b, a), (c, b), (a, c)):
      if y not in nbr[x]:
        nbr[x].append(int(y))
  edgeadr, edges = [], []
  for i in range(nvert):
    edgeadr.append(len(edges))
    edges.extend(sorted(nbr[i]))
    edges.append(-1)
  edges += [-1] * (nvert + 3 * nface - len(edges))  # MuJoCo sizes the block as numvert + 3 numface
  return np.concatenate(([nvert, nface], edgeadr, np.arange(nvert), edges, np.asarray(hull.simplices).reshape(-1))).astype(np.int32)


def _hull_mass_properties(verts):
  """(volume, centre of mass, 3x3 inertia about the centre of mass) of the convex hull of `verts` at unit density: signed
  tetrahedra between the origin and the outward-oriented hull triangles.  For a convex mesh this is what MuJoCo's mesh
  compiler integrates over the faces; a non-convex mesh is treated as its hull (MuJoCo's `inertia="convex"`)."""
  from scipy.spatial import ConvexHull

  v = np.asarray(verts, dtype=np.float64).reshape(-1, 3)
  hull = ConvexHull(v)
  vol, first, second = 0.0, np.zeros(3), np.zeros((3, 3))
  canon = (np.eye(3) + np.ones((3, 3))) / 120.0  # integral of x x^T over the unit tetrahedron with a vertex at the origin
  for tri, eq in zip(hull.simplices, hull.equations):
    a, b, c = v[tri]
    if np.dot(np.cross(b - a, c - a), eq[:3]) < 0:  # orient outward
      b, c = c, b
    A = np.stack([a, b, c], axis=1)
    det = np.linalg.det(A)  # 6 x signed volume
    vol += det / 6.0
    first += det * (a + b + c) / 24.0
    second += det * (A @ canon @ A.T)
  com = first / vol
  cov = second - vol * np.outer(com, com)  # second moment about the centre of mass
  inertia = np.trace(cov) * np.eye(3) - cov
  return float(vol), com, inertia


def _parse_meshes(root, comp, defaults, basedir):
  """name -> hull vertices (scaled, in the mesh file's frame) for every <asset><mesh>; None when the file is absent."""
  meshes = {}
  for asset in root.findall("asset"):
    for e in asset.findall("mesh"):
      a = _resolve(defaults, e, None)
      name = a.get("name") or os.path.splitext(os.path.basename(a.get("file", "")))[0]
      scale = _vec(a.get("scale"), 3, [1, 1, 1])
      if "vertex" in a:
        verts = _vec(a["vertex"]).reshape(-1, 3)
      else:
        path = os.path.join(basedir, comp.meshdir, a.get("file", ""))
        verts = _load_mesh_vertices(path) if os.path.isfile(path) else None
      meshes[name] = None if verts is None else _convex_hull_vertices(verts * scale)
  return meshes


def _read_png_gray(path):
  """8 / 16-bit greyscale (optionally with alpha / RGB: first channel) non-interlaced PNG -> float array (rows, cols).  Enough of the
  format for height-field images; no third-party imaging library is available here."""
  import struct
  import zlib

  data = open(path, "rb").read()
  if data[:8] != b"\x89PNG\r\n\x1a\n":
    raise ValueError(f"{path}: not a PNG file")
  pos, idat, hdr = 8, b"", None
  while pos < len(data):
    n, kind = struct.unpack(">I4s", data[pos : pos + 8])
    body = data[pos + 8 : pos + 8 + n]
    if kind == b"IHDR":
      hdr = struct.unpack(">IIBBBBB", body)
    elif kind == b"IDAT":
      idat += body
    pos += 12 + n
  w, h, depth, ctype, _, _, interlace = hdr
  if interlace or depth not in (8, 16) or ctype not in (0, 2, 4, 6):
    raise NotImplementedError(f"{path}: PNG variant not supported (depth {depth}, colour type {ctype}, interlace {interlace})")
  nch = {0: 1, 2: 3, 4: 2, 6: 4}[ctype]
  bpp = nch * depth // 8
  raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + w * bpp)
  out = np.zeros((h, w * bpp), dtype=np.uint8)
  prev = np.zeros(w * bpp, dtype=np.int32)
  for r in range(h):
    f, line = int(raw[r, 0]), raw[r, 1:].astype(np.int32)
    cur = np.zeros(w * bpp, dtype=np.int32)
    if f == 0:
      cur = line
    elif f == 2:
      cur = (line + prev) & 255
    else:  # sub / average / paeth depend on the pixel to the left: byte-serial
      for i in range(w * bpp):
        a = cur[i - bpp] if i >= bpp else 0
        b = prev[i]
        c = prev[i - bpp] if i >= bpp else 0
        if f == 1:
          pred = a
        elif f == 3:
          pred = (a + b) >> 1
        else:
          pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
          pred = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
        cur[i] = (line[i] + pred) & 255
    out[r] = cur
    prev = cur
  px = out.reshape(h, w, nch, depth // 8)
  val = px[..., 0].astype(np.float64) if depth == 8 else px[..., 0].astype(np.float64) * 256 + px[..., 1]
  return val[:, :, 0]


def _parse_hfields(root, comp, defaults, basedir):
  """name -> (size[4], nrow, ncol, data) for every <asset><hfield>.  Like MuJoCo's compiler, elevation data are shifted and scaled
  to [0, 1] and image rows are flipped (the first image row is the far edge, +y)."""
  out = {}
  for asset in root.findall("asset"):
    for e in asset.findall("hfield"):
      a = _resolve(defaults, e, None)
      name = a.get("name") or os.path.splitext(os.path.basename(a.get("file", "")))[0]
      size = _vec(a.get("size"), 4)
      if "elevation" in a:
        nrow, ncol = int(a["nrow"]), int(a["ncol"])
        data = _vec(a["elevation"]).reshape(nrow, ncol)[::-1]
      elif "file" in a:
        path = os.path.join(basedir, comp.meshdir if not os.path.isfile(os.path.join(basedir, a["file"])) else "", a["file"])
        data = _read_png_gray(path)[::-1]
        nrow, ncol = data.shape
      else:
        nrow, ncol = int(a["nrow"]), int(a["ncol"])
        data = np.zeros((nrow, ncol))
      data = np.asarray(data, dtype=np.float64)
      lo, hi = float(data.min()), float(data.max())
      data = (data - lo) / (hi - lo) if hi > lo else np.zeros_like(data)
      out[name] = (size, nrow, ncol, np.ascontiguousarray(data))
  return out


# ----------------------------------------------------------------------------- parse tree
class _Compiler:
  def __init__(self):
    self.degree = True
    self.eulerseq = "xyz"
    self.autolimits = True
    self.meshdir = ""
    self.inertiagrouprange = (0, 5)
    self.boundmass = 0.0
    self.boundinertia = 0.0


def _expand_includes(root, basedir):
  """Inline <include file=.../> elements (recursively)."""
  changed = True
  while changed:
    changed = False
    for parent in root.iter():
      for i, child in enumerate(list(parent)):
        if child.tag == "include":
          path = os.path.join(basedir, child.get("file"))
          sub = ET.parse(path).getroot()
          _expand_includes(sub, os.path.dirname(path))
          idx = list(parent).index(child)
          parent.remove(child)
          for j, e in enumerate(list(sub)):
            parent.insert(idx + j, e)
          changed = True


_NAME_REFS = ("joint", "joint1", "joint2", "body", "body1", "body2", "geom", "geom1", "geom2", "site", "site1", "site2", "target",
              "jointinparent", "tendon", "objname", "refname")  # attributes that hold the name of another element


def _rename_tree(elem, fn_name, fn_class):
  """Applies fn_name to every name / name reference and fn_class to every class reference below (and at) elem."""
  for e in elem.iter():
    if "name" in e.attrib:
      e.set("name", fn_name(e.get("name")))
    for k in _NAME_REFS:
      if k in e.attrib:
        e.set(k, fn_name(e.get(k)))
    for k in ("class", "childclass"):
      if k in e.attrib:
        e.set(k, fn_class(e.get(k)))


def _expand_replicate(root):
  """<replicate count offset euler sep>: N copies of the children, copy i inside a <frame> moved by the accumulated offset / rotation,
  names (and the references to them inside the copy) suffixed with sep + i.  <attach> children remember the suffix for _expand_attach."""
  import copy

  while True:
    hit = next(((p, c) for p in root.iter() for c in p if c.tag == "replicate"), None)
    if hit is None:
      return
    parent, rep = hit
    count = int(rep.get("count"))
    sep = rep.get("sep", "")
    width = len(str(count - 1))
    idx = list(parent).index(rep)
    parent.remove(rep)
    local = {e.get("name") for e in rep.iter() if "name" in e.attrib}
    for i in range(count):
      suffix = sep + str(i).zfill(width)
      fr = ET.Element("frame")
      fr.set("_rep_offset", rep.get("offset", "0 0 0"))
      fr.set("_rep_euler", rep.get("euler", "0 0 0"))
      fr.set("_rep_index", str(i))
      for child in rep:
        c = copy.deepcopy(child)
        _rename_tree(c, lambda n: n + suffix if n in local else n, lambda k: k)
        for a in c.iter("attach"):  # includes c itself
          a.set("_suffix", a.get("_suffix", "") + suffix)
        fr.append(c)
      parent.insert(idx + i, fr)


def _expand_attach(root, basedir):
  """<attach model body prefix>: the named body of a sub-model (<asset><model file>) is copied in place with prefixed names, together
  with the sub-model's default classes and the actuators / contact pairs and excludes / equalities that refer to the copied subtree."""
  import copy

  models = {m.get("name", os.path.splitext(os.path.basename(m.get("file")))[0]): m.get("file") for a in root.findall("asset") for m in a.findall("model")}
  subs = {}
  done_defaults = set()
  while True:
    hit = next(((p, c) for p in root.iter() for c in p if c.tag == "attach"), None)
    if hit is None:
      break
    parent, att = hit
    mname, prefix, suffix = att.get("model"), att.get("prefix", ""), att.get("_suffix", "")
    if mname not in subs:
      path = os.path.join(basedir, models[mname])
      sub = ET.parse(path).getroot()
      _expand_includes(sub, os.path.dirname(path))
      _expand_structure(sub, os.path.dirname(path))
      for tag in ("compiler",):
        if [dict(c.attrib) for c in sub.findall(tag)] != [dict(c.attrib) for c in root.findall(tag)]:
          raise NotImplementedError("<attach>: sub-model and parent must use the same <compiler> settings")
      subs[mname] = sub
    sub = subs[mname]
    body = next((b for b in sub.iter("body") if b.get("name") == att.get("body")), None)
    if body is None:
      raise ValueError(f"<attach>: body {att.get('body')!r} not found in model {mname!r}")
    inside = {e.get("name") for e in body.iter() if "name" in e.attrib}
    fn_name = lambda n: prefix + n + suffix if n in inside else n  # noqa: E731
    fn_class = lambda k: prefix + k  # noqa: E731
    main = prefix + "main"
    if (mname, prefix) not in done_defaults:  # the sub-model's class tree, once, below the parent's main class
      done_defaults.add((mname, prefix))
      top = root.find("default")
      if top is None:
        top = ET.Element("default")
        root.insert(0, top)
      holder = ET.SubElement(top, "default")
      holder.set("class", main)
      for d in sub.findall("default"):
        for e in d:
          c = copy.deepcopy(e)
          for n in c.iter("default"):  # includes c itself
            n.set("class", prefix + n.get("class"))
          for n in c.iter():
            if n.tag != "default" and "class" in n.attrib:
              n.set("class", prefix + n.get("class"))
          holder.append(c)
    b = copy.deepcopy(body)
    _rename_tree(b, fn_name, fn_class)
    if "childclass" not in b.attrib:
      b.set("childclass", main)
    idx = list(parent).index(att)
    parent.remove(att)
    parent.insert(idx, b)
    for sec in ("actuator", "contact", "equality"):
      for src in sub.findall(sec):
        for e in src:
          refs = [e.get(k) for k in _NAME_REFS if k in e.attrib]
          if not refs or not all(r in inside or r == "world" for r in refs) or not any(r in inside for r in refs):
            continue
          c = copy.deepcopy(e)
          if "name" in c.attrib:
            c.set("name", prefix + c.get("name") + suffix)
          for k in _NAME_REFS:
            if k in c.attrib:
              c.set(k, fn_name(c.get(k)))
          c.set("class", prefix + c.get("class") if "class" in c.attrib else main)
          dst = root.find(sec)
          if dst is None:
            dst = ET.SubElement(root, sec)
          dst.append(c)
  for a in root.findall("asset"):
    for m in a.findall("model"):
      a.remove(m)


def _expand_frames(root, comp):
  """<frame pos quat|euler|...>: a pure coordinate transform, folded into the pose of its children (bodies, geoms, sites, nested
  frames; cameras and lights are not compiled).  Replicate frames carry their accumulated offset / rotation."""
  while True:
    hit = next(((p, c) for p in root.iter() for c in p if c.tag == "frame"), None)
    if hit is None:
      return
    parent, fr = hit
    fpos = _vec(fr.get("pos"), 3, [0, 0, 0])
    fquat = _orientation(fr.attrib, comp)
    if "_rep_index" in fr.attrib:  # copy i of a replicate: i accumulated steps of (offset, euler)
      dq = _orientation({"euler": fr.get("_rep_euler")}, comp)
      off = _vec(fr.get("_rep_offset"), 3)
      fpos, fquat = np.zeros(3), np.array([1.0, 0, 0, 0])
      for _ in range(int(fr.get("_rep_index"))):
        fpos = fpos + rot_vec_quat(off, fquat)
        fquat = quat_mul(fquat, dq)
    idx = list(parent).index(fr)
    parent.remove(fr)
    for j, c in enumerate(list(fr)):
      if c.tag in ("body", "geom", "site", "frame", "camera", "light"):
        if "fromto" in c.attrib:
          ft = _vec(c.get("fromto"), 6)
          a, b = fpos + rot_vec_quat(ft[:3], fquat), fpos + rot_vec_quat(ft[3:], fquat)
          c.set("fromto", " ".join(repr(float(x)) for x in np.concatenate([a, b])))
        else:
          cpos = _vec(c.get("pos"), 3, [0, 0, 0])
          cquat = _orientation(c.attrib, comp)
          for k in ("quat", "axisangle", "euler", "xyaxes", "zaxis"):
            c.attrib.pop(k, None)
          c.set("pos", " ".join(repr(float(x)) for x in fpos + rot_vec_quat(cpos, fquat)))
          c.set("quat", " ".join(repr(float(x)) for x in quat_mul(fquat, cquat)))
        if "childclass" in fr.attrib and c.tag in ("body", "frame") and "childclass" not in c.attrib:
          c.set("childclass", fr.get("childclass"))
      parent.insert(idx + j, c)


def _expand_structure(root, basedir):
  """Model-composition elements of MJCF, rewritten into plain bodies before compilation: <replicate>, <attach>, <frame>."""
  if not any(e.tag in ("replicate", "attach", "frame") for e in root.iter()):
    return
  comp = _Compiler()
  for c in root.findall("compiler"):
    if "angle" in c.attrib:
      comp.degree = c.get("angle") == "degree"
    if "eulerseq" in c.attrib:
      comp.eulerseq = c.get("eulerseq")
  _expand_replicate(root)
  _expand_attach(root, basedir)
  _expand_frames(root, comp)


def _bool(s, default=False):
  if s is None:
    return default
  return s.strip().lower() == "true"


def _limited(attr, has_range, autolimits):
  """MJCF tri-state limited attribute: 'true' / 'false' / 'auto'."""
  if attr is None or attr == "auto":
    return bool(has_range) if autolimits else False
  return attr.strip().lower() == "true"


class MjModelLite(SimpleNamespace):
  """Host model with MuJoCo `MjModel` attribute names (numpy, float64/int32)."""

  def keyframe(self, name_or_id):
    kid = self.key_names.index(name_or_id) if isinstance(name_or_id, str) else int(name_or_id)
    return SimpleNamespace(
      qpos=self.key_qpos[kid].copy(), qvel=self.key_qvel[kid].copy(), ctrl=self.key_ctrl[kid].copy(), act=self.key_act[kid].copy()
    )


def save_model(mjm: MjModelLite, path: str) -> None:
  """Writes a compiled host model to one .npz file (the on-disk model cache: compiling aloha_pot's 133 meshes -- hulls, mass
  properties, hill-climbing graphs -- takes seconds, loading the cache milliseconds).  Arrays are stored as they are; scalars, name
  lists and the option / statistic blocks travel in one JSON entry.  `load_model` restores an identical MjModelLite."""
  import json

  arrays, meta = {}, {"scalars": {}, "lists": {}, "ns": {}}
  for k, v in vars(mjm).items():
    if isinstance(v, np.ndarray):
      arrays["a/" + k] = v
    elif isinstance(v, SimpleNamespace):
      meta["ns"][k] = {}
      for kk, vv in vars(v).items():
        if isinstance(vv, np.ndarray):
          arrays[f"n/{k}/{kk}"] = vv
        else:
          meta["ns"][k][kk] = vv
    elif isinstance(v, (list, tuple)):
      meta["lists"][k] = list(v)
    elif isinstance(v, (int, float, str, bool, np.integer, np.floating)) or v is None:
      meta["scalars"][k] = v.item() if isinstance(v, (np.integer, np.floating)) else v
    else:
      raise TypeError(f"save_model: field {k} of type {type(v).__name__} is not serialisable")
  arrays["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
  with open(path, "wb") as fh:
    np.savez_compressed(fh, **arrays)


def load_model(path: str) -> MjModelLite:
  """Reads a model written by `save_model`."""
  import json

  with np.load(path) as z:
    meta = json.loads(bytes(z["meta"]).decode())
    m = MjModelLite()
    for k, v in meta["scalars"].items():
      setattr(m, k, v)
    for k, v in meta["lists"].items():
      setattr(m, k, v)
    for k, d in meta["ns"].items():
      setattr(m, k, SimpleNamespace(**d))
    for name in z.files:
      if name.startswith("a/"):
        setattr(m, name[2:], z[name])
      elif name.startswith("n/"):
        _, k, kk = name.split("/", 2)
        setattr(getattr(m, k), kk, z[name])
  return m


def load_mjcf(path: str) -> MjModelLite:
  """Parse and compile an MJCF file."""
  tree = ET.parse(path)
  root = tree.getroot()
  _expand_includes(root, os.path.dirname(os.path.abspath(path)))
  _expand_structure(root, os.path.dirname(os.path.abspath(path)))
  return _compile(root, os.path.dirname(os.path.abspath(path)))


def load_mjcf_string(xml: str, basedir: str = ".") -> MjModelLite:
  root = ET.fromstring(xml)
  _expand_includes(root, basedir)
  _expand_structure(root, basedir)
  return _compile(root, basedir)


def _compile(root, basedir) -> MjModelLite:
  comp = _Compiler()
  for c in root.findall("compiler"):
    if "angle" in c.attrib:
      comp.degree = c.get("angle") == "degree"
    if "eulerseq" in c.attrib:
      comp.eulerseq = c.get("eulerseq")
    if "autolimits" in c.attrib:
      comp.autolimits = _bool(c.get("autolimits"))
    if "meshdir" in c.attrib:
      comp.meshdir = c.get("meshdir")
    if "boundmass" in c.attrib:
      comp.boundmass = float(c.get("boundmass"))
    if "boundinertia" in c.attrib:
      comp.boundinertia = float(c.get("boundinertia"))

  defaults = _Defaults()
  for dflt in root.findall("default"):
    defaults.parse(dflt, None)

  # ---- option
  opt = SimpleNamespace(
    timestep=0.002,
    impratio=1.0,
    tolerance=1e-8,
    ls_tolerance=0.01,
    noslip_tolerance=1e-6,
    ccd_tolerance=1e-6,
    gravity=np.array([0.0, 0.0, -9.81]),
    wind=np.zeros(3),
    magnetic=np.array([0.0, -0.5, 0.0]),
    density=0.0,
    viscosity=0.0,
    integrator=int(C.IntegratorType.EULER),
    cone=int(C.ConeType.PYRAMIDAL),
    jacobian=int(C.JacobianType.AUTO),
    solver=int(C.SolverType.NEWTON),
    iterations=100,
    ls_iterations=50,
    noslip_iterations=0,
    ccd_iterations=35,
    disableflags=0,
    enableflags=0,
    sdf_initpoints=40,
    sdf_iterations=10,
  )
  _enum = {
    "integrator": {"Euler": 0, "RK4": 1, "implicit": 2, "implicitfast": 3},
    "cone": {"pyramidal": 0, "elliptic": 1},
    "jacobian": {"dense": 0, "sparse": 1, "auto": 2},
    "solver": {"PGS": 0, "CG": 1, "Newton": 2},
  }
  for o in root.findall("option"):
    for k, v in o.attrib.items():
      if k in _enum:
        setattr(opt, k, _enum[k][v])
      elif k in ("gravity", "wind", "magnetic"):
        setattr(opt, k, _vec(v, 3))
      elif k in ("iterations", "ls_iterations", "noslip_iterations", "ccd_iterations", "sdf_initpoints", "sdf_iterations"):
        setattr(opt, k, int(v))
      elif hasattr(opt, k):
        setattr(opt, k, float(v))
    for f in o.findall("flag"):
      for k, v in f.attrib.items():
        key = k.upper()
        if key in C.DisableBit.__members__:
          bit = int(C.DisableBit[key])
          if v == "disable":
            opt.disableflags |= bit
          else:
            opt.disableflags &= ~bit
        elif key in C.EnableBit.__members__:
          bit = int(C.EnableBit[key])
          if v == "enable":
            opt.enableflags |= bit
          else:
            opt.enableflags &= ~bit

  # ---- traverse bodies (DFS pre-order, MuJoCo id assignment)
  bodies, joints, geoms, sites = [], [], [], []

  def add_body(elem, parent_id, childclass):
    bid = len(bodies)
    if elem.tag == "worldbody":
      b = dict(name="world", parent=0, pos=np.zeros(3), quat=np.array([1.0, 0, 0, 0]), mocap=False, gravcomp=0.0, inertial=None)
    else:
      a = dict(elem.attrib)
      childclass = a.get("childclass", childclass)
      b = dict(
        name=a.get("name", f"body{bid}"),
        parent=parent_id,
        pos=_vec(a.get("pos"), 3, [0, 0, 0]),
        quat=_orientation(a, comp),
        mocap=_bool(a.get("mocap")),
        gravcomp=float(a.get("gravcomp", 0.0)),
        inertial=None,
      )
    bodies.append(b)
    b["jnt"], b["geom"] = [], []
    for child in elem:
      if child.tag == "inertial":
        ia = child.attrib
        inert = dict(pos=_vec(ia.get("pos"), 3, [0, 0, 0]), quat=_orientation(ia, comp), mass=float(ia["mass"]))
        if "fullinertia" in ia:
          f = _vec(ia["fullinertia"], 6)
          full = np.array([[f[0], f[3], f[4]], [f[3], f[1], f[5]], [f[4], f[5], f[2]]])
          ev, q = eig3(full)
          inert["inertia"] = ev
          inert["quat"] = quat_mul(inert["quat"], q)
        else:
          inert["inertia"] = _vec(ia.get("diaginertia"), 3, [0, 0, 0])
        b["inertial"] = inert
      elif child.tag in ("joint", "freejoint"):
        if child.tag == "freejoint":
          a = dict(child.attrib)
          a["type"] = "free"
        else:
          a = _resolve(defaults, child, childclass)
        a["_body"] = bid
        b["jnt"].append(len(joints))
        joints.append(a)
      elif child.tag == "geom":
        a = _resolve(defaults, child, childclass)
        a["_body"] = bid
        b["geom"].append(len(geoms))
        geoms.append(a)
      elif child.tag == "site":
        a = _resolve(defaults, child, childclass)
        a["_body"] = bid
        sites.append(a)
    for child in elem:
      if child.tag == "body":
        add_body(child, bid, childclass)

  wb = root.findall("worldbody")
  add_body(wb[0], 0, None)
  for extra in wb[1:]:  # additional worldbody sections (from includes) merge into the world
    for child in extra:
      wb[0].append(child)
  if len(wb) > 1:
    bodies.clear(), joints.clear(), geoms.clear(), sites.clear()
    add_body(wb[0], 0, None)

  return _finalize(root, comp, defaults, opt, bodies, joints, geoms, sites, basedir)


def _finalize(root, comp, defaults, opt, bodies, joints, geoms, sites, basedir) -> MjModelLite:
  m = MjModelLite()
  m.opt = opt
  nbody, njnt, ngeom, nsite = len(bodies), len(joints), len(geoms), len(sites)
  f64, i32 = np.float64, np.int32

  # ---------------- bodies
  m.nbody = nbody
  m.body_names = [b["name"] for b in bodies]
  m.body_parentid = np.array([b["parent"] for b in bodies], dtype=i32)
  m.body_pos = np.array([b["pos"] for b in bodies], dtype=f64).reshape(nbody, 3)
  m.body_quat = np.array([b["quat"] for b in bodies], dtype=f64).reshape(nbody, 4)
  m.body_gravcomp = np.array([b["gravcomp"] for b in bodies], dtype=f64)
  m.body_jntnum = np.array([len(b["jnt"]) for b in bodies], dtype=i32)
  m.body_jntadr = np.array([b["jnt"][0] if b["jnt"] else -1 for b in bodies], dtype=i32)
  m.body_geomnum = np.array([len(b["geom"]) for b in bodies], dtype=i32)
  m.body_geomadr = np.array([b["geom"][0] if b["geom"] else -1 for b in bodies], dtype=i32)
  mocap = [i for i, b in enumerate(bodies) if b["mocap"]]
  m.nmocap = len(mocap)
  m.body_mocapid = np.full(nbody, -1, dtype=i32)
  for k, i in enumerate(mocap):
    m.body_mocapid[i] = k

  # ---------------- joints / dofs
  m.njnt = njnt
  m.jnt_names = [j.get("name", f"jnt{i}") for i, j in enumerate(joints)]
  m.jnt_type = np.zeros(njnt, dtype=i32)
  m.jnt_bodyid = np.zeros(njnt, dtype=i32)
  m.jnt_qposadr = np.zeros(njnt, dtype=i32)
  m.jnt_dofadr = np.zeros(njnt, dtype=i32)
  m.jnt_pos = np.zeros((njnt, 3))
  m.jnt_axis = np.zeros((njnt, 3))
  m.jnt_stiffness = np.zeros(njnt)
  m.jnt_range = np.zeros((njnt, 2))
  m.jnt_actfrcrange = np.zeros((njnt, 2))
  m.jnt_margin = np.zeros(njnt)
  m.jnt_limited = np.zeros(njnt, dtype=bool)
  m.jnt_actfrclimited = np.zeros(njnt, dtype=bool)
  m.jnt_actgravcomp = np.zeros(njnt, dtype=i32)
  m.jnt_solref = np.tile([0.02, 1.0], (njnt, 1))
  m.jnt_solimp = np.tile([0.9, 0.95, 0.001, 0.5, 2.0], (njnt, 1))
  qpos0, qpos_spring = [], []
  dof_jntid, dof_bodyid, dof_armature, dof_damping, dof_frictionloss = [], [], [], [], []
  dof_solref, dof_solimp = [], []
  nq = nv = 0
  ang = math.pi / 180.0 if comp.degree else 1.0
  for i, j in enumerate(joints):
    jt = _JNT_TYPES[j.get("type", "hinge")]
    bid = j["_body"]
    m.jnt_type[i] = jt
    m.jnt_bodyid[i] = bid
    m.jnt_qposadr[i] = nq
    m.jnt_dofadr[i] = nv
    m.jnt_pos[i] = _vec(j.get("pos"), 3, [0, 0, 0])
    axis = _vec(j.get("axis"), 3, [0, 0, 1])
    m.jnt_axis[i] = axis / max(np.linalg.norm(axis), mjMINVAL)
    m.jnt_stiffness[i] = float(j.get("stiffness", 0))
    m.jnt_margin[i] = float(j.get("margin", 0))
    rng = _vec(j.get("range"), 2, [0, 0])
    has_range = "range" in j and (rng[0] != 0 or rng[1] != 0)
    if jt in (C.JointType.HINGE, C.JointType.BALL):
      rng = rng * ang
    m.jnt_range[i] = rng
    m.jnt_limited[i] = _limited(j.get("limited"), has_range, comp.autolimits)
    afr = _vec(j.get("actuatorfrcrange"), 2, [0, 0])
    m.jnt_actfrcrange[i] = afr
    m.jnt_actfrclimited[i] = _limited(j.get("actuatorfrclimited"), "actuatorfrcrange" in j, comp.autolimits)
    m.jnt_actgravcomp[i] = int(_bool(j.get("actuatorgravcomp")))
    if "solreflimit" in j:
      v = _vec(j["solreflimit"])
      m.jnt_solref[i, : v.size] = v
    if "solimplimit" in j:
      v = _vec(j["solimplimit"])
      m.jnt_solimp[i, : v.size] = v
    ref = float(j.get("ref", 0)) * (ang if jt == C.JointType.HINGE else 1.0)
    sref = float(j.get("springref", 0)) * (ang if jt == C.JointType.HINGE else 1.0)
    if jt == C.JointType.FREE:
      q0 = np.concatenate([bodies[bid]["pos"], bodies[bid]["quat"]])
      qpos0 += list(q0)
      qpos_spring += list(q0)
      nqj, nvj = 7, 6
      m.jnt_axis[i] = [0, 0, 1]
    elif jt == C.JointType.BALL:
      qpos0 += [1.0, 0, 0, 0]
      qpos_spring += [1.0, 0, 0, 0]
      nqj, nvj = 4, 3
    else:
      qpos0 += [ref]
      qpos_spring += [sref]
      nqj, nvj = 1, 1
    for _ in range(nvj):
      dof_jntid.append(i)
      dof_bodyid.append(bid)
      dof_armature.append(float(j.get("armature", 0)))
      dof_damping.append(float(j.get("damping", 0)))
      dof_frictionloss.append(float(j.get("frictionloss", 0)))
      sr = np.array([0.02, 1.0])
      si = np.array([0.9, 0.95, 0.001, 0.5, 2.0])
      if "solreffriction" in j:
        v = _vec(j["solreffriction"])
        sr[: v.size] = v
      if "solimpfriction" in j:
        v = _vec(j["solimpfriction"])
        si[: v.size] = v
      dof_solref.append(sr)
      dof_solimp.append(si)
    nq += nqj
    nv += nvj
  m.nq, m.nv = nq, nv
  m.qpos0 = np.array(qpos0, dtype=f64)
  m.qpos_spring = np.array(qpos_spring, dtype=f64)
  m.dof_jntid = np.array(dof_jntid, dtype=i32)
  m.dof_bodyid = np.array(dof_bodyid, dtype=i32)
  m.dof_armature = np.array(dof_armature, dtype=f64)
  m.dof_damping = np.array(dof_damping, dtype=f64)
  m.dof_frictionloss = np.array(dof_frictionloss, dtype=f64)
  m.dof_solref = np.array(dof_solref, dtype=f64).reshape(nv, 2)
  m.dof_solimp = np.array(dof_solimp, dtype=f64).reshape(nv, 5)
  m.body_dofnum = np.zeros(nbody, dtype=i32)
  m.body_dofadr = np.full(nbody, -1, dtype=i32)
  for d in range(nv):
    b = m.dof_bodyid[d]
    if m.body_dofadr[b] < 0:
      m.body_dofadr[b] = d
    m.body_dofnum[b] += 1
  # dof_parentid: previous dof in the same body, else last dof of nearest ancestor with dofs
  m.dof_parentid = np.full(nv, -1, dtype=i32)
  for d in range(nv):
    b = m.dof_bodyid[d]
    if d > m.body_dofadr[b]:
      m.dof_parentid[d] = d - 1
    else:
      p = m.body_parentid[b]
      while p > 0 and m.body_dofnum[p] == 0:
        p = m.body_parentid[p]
      if p > 0:
        m.dof_parentid[d] = m.body_dofadr[p] + m.body_dofnum[p] - 1
  # body_rootid / weldid / treeid
  m.body_rootid = np.zeros(nbody, dtype=i32)
  m.body_weldid = np.zeros(nbody, dtype=i32)
  for b in range(1, nbody):
    p = m.body_parentid[b]
    m.body_rootid[b] = b if p == 0 else m.body_rootid[p]
    m.body_weldid[b] = b if (m.body_jntnum[b] > 0 or bodies[b]["mocap"]) else m.body_weldid[p]
  m.dof_treeid = np.zeros(nv, dtype=i32)
  roots = [d for d in range(nv) if m.dof_parentid[d] == -1]
  m.ntree = len(roots)
  tid = -1
  for d in range(nv):
    if m.dof_parentid[d] == -1:
      tid += 1
    m.dof_treeid[d] = tid
  # legacy sparse M addressing (diag then ancestors)
  m.dof_Madr = np.zeros(nv, dtype=i32)
  adr = 0
  for d in range(nv):
    m.dof_Madr[d] = adr
    k = d
    while k >= 0:
      adr += 1
      k = m.dof_parentid[k]
  m.nM = adr
  # CSR lower-triangular layout (row = dof, columns = ancestors ascending, diagonal last)
  m.M_rownnz = np.zeros(nv, dtype=i32)
  m.M_rowadr = np.zeros(nv, dtype=i32)
  colind, mapM2M = [], []
  for d in range(nv):
    chain = []
    k, a = d, m.dof_Madr[d]
    while k >= 0:
      chain.append((k, a))
      a += 1
      k = m.dof_parentid[k]
    chain.reverse()
    m.M_rowadr[d] = len(colind)
    m.M_rownnz[d] = len(chain)
    colind += [c for c, _ in chain]
    mapM2M += [a for _, a in chain]
  m.M_colind = np.array(colind, dtype=i32)
  m.mapM2M = np.array(mapM2M, dtype=i32)
  m.nC = len(colind)

  # ---------------- geoms
  m.ngeom = ngeom
  m.geom_names = [g.get("name", f"geom{i}") for i, g in enumerate(geoms)]
  m.geom_type = np.zeros(ngeom, dtype=i32)
  m.geom_bodyid = np.array([g["_body"] for g in geoms], dtype=i32)
  m.geom_contype = np.ones(ngeom, dtype=i32)
  m.geom_conaffinity = np.ones(ngeom, dtype=i32)
  m.geom_condim = np.full(ngeom, 3, dtype=i32)
  m.geom_priority = np.zeros(ngeom, dtype=i32)
  m.geom_group = np.zeros(ngeom, dtype=i32)
  m.geom_dataid = np.full(ngeom, -1, dtype=i32)
  m.geom_size = np.zeros((ngeom, 3))
  m.geom_pos = np.zeros((ngeom, 3))
  m.geom_quat = np.tile([1.0, 0, 0, 0], (ngeom, 1))
  m.geom_friction = np.tile([1.0, 0.005, 0.0001], (ngeom, 1))
  m.geom_solmix = np.ones(ngeom)
  m.geom_solref = np.tile([0.02, 1.0], (ngeom, 1))
  m.geom_solimp = np.tile([0.9, 0.95, 0.001, 0.5, 2.0], (ngeom, 1))
  m.geom_margin = np.zeros(ngeom)
  m.geom_gap = np.zeros(ngeom)
  m.geom_rbound = np.zeros(ngeom)
  m.geom_aabb = np.zeros((ngeom, 6))
  geom_mass = np.zeros(ngeom)
  geom_inertia = np.zeros((ngeom, 3))
  geom_mesh_com = np.zeros((ngeom, 3))  # mesh geoms: centre of mass in the geom (= mesh file) frame and the full tensor about it
  geom_mesh_inertia = np.zeros((ngeom, 3, 3))
  meshes = _parse_meshes(root, comp, defaults, basedir)
  hfields = _parse_hfields(root, comp, defaults, basedir)
  hfield_names = list(hfields)  # every height-field asset enters the model, in file order (geom_dataid indexes this list)
  mesh_ids = {}  # only meshes referenced by a geom enter the model, in order of first use
  hfield_geoms = {}
  for i, g in enumerate(geoms):
    gt = _GEOM_TYPES[g.get("type", "mesh" if "mesh" in g else "sphere")]
    if gt == C.GeomType.HFIELD and g.get("hfield") not in hfields:
      raise NotImplementedError(f"geom type hfield needs an <asset><hfield> named '{g.get('hfield')}'")
    if gt == C.GeomType.HFIELD:
      hfield_geoms[i] = hfield_names.index(g["hfield"])
      hs = hfields[g["hfield"]][0]
      g = dict(g, size=f"{hs[0]} {hs[1]} {hs[2]}")  # bounds below; the collision code reads hfield_size
    if gt == C.GeomType.MESH:
      # Collision uses the convex hull of the asset's vertices, kept in the mesh file's frame: MuJoCo re-centres a mesh on its
      # inertial frame and folds that transform into geom_pos / geom_quat -- the world-space shape is the same, only the
      # reported geom_xpos / geom_xmat of mesh geoms differ.  Mesh mass properties are not computed: the body needs <inertial>.
      verts = meshes.get(g.get("mesh"))
      collides = int(g.get("contype", 1)) != 0 or int(g.get("conaffinity", 1)) != 0
      if verts is None and collides:
        raise NotImplementedError(f"mesh asset '{g.get('mesh')}' of a colliding geom is missing (mesh files are loaded from meshdir)")
      m.geom_type[i] = gt
      m.geom_pos[i], m.geom_quat[i] = _vec(g.get("pos"), 3, [0, 0, 0]), _orientation(g, comp)
      for key, arr, cast in (("contype", m.geom_contype, int), ("conaffinity", m.geom_conaffinity, int), ("condim", m.geom_condim, int),
                             ("priority", m.geom_priority, int), ("group", m.geom_group, int), ("solmix", m.geom_solmix, float),
                             ("margin", m.geom_margin, float), ("gap", m.geom_gap, float)):  # fmt: skip
        if key in g:
          arr[i] = cast(g[key])
      for key, arr in (("friction", m.geom_friction), ("solref", m.geom_solref), ("solimp", m.geom_solimp)):
        if key in g:
          v = _vec(g[key])
          arr[i, : v.size] = v
      if verts is not None:
        if g.get("mesh") not in mesh_ids:
          mesh_ids[g.get("mesh")] = len(mesh_ids)
        m.geom_dataid[i] = mesh_ids[g.get("mesh")]
        lo, hi = verts.min(axis=0), verts.max(axis=0)
        m.geom_size[i] = 0.5 * (hi - lo)
        m.geom_rbound[i] = float(np.linalg.norm(verts, axis=1).max())
        m.geom_aabb[i] = np.concatenate((0.5 * (lo + hi), 0.5 * (hi - lo)))
        if bodies[g["_body"]]["inertial"] is None and g["_body"] != 0 and len(verts) >= 4:
          vol, com, inert = _hull_mass_properties(verts)
          dens = float(g["mass"]) / max(vol, mjMINVAL) if "mass" in g else float(g.get("density", 1000.0))
          geom_mass[i] = dens * vol
          geom_mesh_com[i], geom_mesh_inertia[i] = com, inert * dens
      elif bodies[g["_body"]]["inertial"] is None and g["_body"] != 0 and comp.inertiagrouprange[0] <= m.geom_group[i] <= comp.inertiagrouprange[1]:
        raise NotImplementedError(f"mesh asset '{g.get('mesh')}' is missing and its body has no <inertial>: the body's mass properties cannot be computed")
      continue
    m.geom_type[i] = gt
    size = np.zeros(3)
    sv = _vec(g.get("size"))
    if sv is not None:
      size[: sv.size] = sv
    pos = _vec(g.get("pos"), 3, [0, 0, 0])
    quat = _orientation(g, comp)
    if "fromto" in g:
      ft = _vec(g["fromto"], 6)
      vec = ft[0:3] - ft[3:6]
      size[1] = np.linalg.norm(vec) / 2
      quat = z2quat(vec)
      pos = 0.5 * (ft[0:3] + ft[3:6])
    m.geom_size[i], m.geom_pos[i], m.geom_quat[i] = size, pos, quat
    for key, arr, cast in (
      ("contype", m.geom_contype, int),
      ("conaffinity", m.geom_conaffinity, int),
      ("condim", m.geom_condim, int),
      ("priority", m.geom_priority, int),
      ("group", m.geom_group, int),
      ("solmix", m.geom_solmix, float),
      ("margin", m.geom_margin, float),
      ("gap", m.geom_gap, float),
    ):
      if key in g:
        arr[i] = cast(g[key])
    for key, arr in (("friction", m.geom_friction), ("solref", m.geom_solref), ("solimp", m.geom_solimp)):
      if key in g:
        v = _vec(g[key])
        arr[i, : v.size] = v
    if gt not in (C.GeomType.PLANE, C.GeomType.HFIELD):
      vol, inert = _geom_volume_inertia(gt, size)
      if "mass" in g:
        mass = float(g["mass"])
        dens = mass / max(vol, mjMINVAL)
      else:
        dens = float(g.get("density", 1000.0))
        mass = dens * vol
      geom_mass[i] = mass
      geom_inertia[i] = inert * dens
    if gt == C.GeomType.HFIELD:
      # bounding volume of the field: the box [-x, x] x [-y, y] x [-z_bottom, z_top] in the geom frame (MuJoCo's compiler is not
      # available to compare with; these feed the broadphase filters only, the narrowphase reads hfield_size)
      hs = hfields[hfield_names[hfield_geoms[i]]][0]
      m.geom_dataid[i] = hfield_geoms[i]
      m.geom_size[i] = [hs[0], hs[1], hs[2]]
      m.geom_rbound[i] = float(np.sqrt(hs[0] ** 2 + hs[1] ** 2 + max(hs[2], hs[3]) ** 2))
      m.geom_aabb[i] = [0, 0, 0.5 * (hs[2] - hs[3]), hs[0], hs[1], 0.5 * (hs[2] + hs[3])]
    else:
      m.geom_rbound[i], m.geom_aabb[i] = _geom_bounds(gt, size)

  # ---------------- mesh tables (convex hull vertices; no hill-climbing graph: mesh_graphadr = -1 selects the exhaustive support
  # search, reference collision_gjk.py:142)
  m.nhfield = len(hfield_names)
  m.hfield_names = hfield_names
  m.hfield_size = np.array([hfields[n][0] for n in hfield_names], dtype=np.float64).reshape(-1, 4)
  m.hfield_nrow = np.array([hfields[n][1] for n in hfield_names], dtype=i32)
  m.hfield_ncol = np.array([hfields[n][2] for n in hfield_names], dtype=i32)
  m.hfield_adr = np.concatenate(([0], np.cumsum(m.hfield_nrow * m.hfield_ncol)[:-1])).astype(i32) if hfield_names else np.zeros(0, dtype=i32)
  m.hfield_data = np.concatenate([hfields[n][3].ravel() for n in hfield_names]) if hfield_names else np.zeros(0)
  m.nhfielddata = int(m.hfield_data.size)
  names = sorted(mesh_ids, key=mesh_ids.get)
  m.nmesh = len(names)
  m.mesh_names = names
  m.mesh_vertnum = np.array([len(meshes[n]) for n in names], dtype=i32)
  m.mesh_vertadr = np.concatenate(([0], np.cumsum(m.mesh_vertnum)[:-1])).astype(i32) if names else np.zeros(0, dtype=i32)
  m.mesh_vert = np.concatenate([meshes[n] for n in names]).reshape(-1, 3) if names else np.zeros((0, 3))
  m.nmeshvert = int(m.mesh_vert.shape[0])
  # hill-climbing graph of each hull in MuJoCo's mesh_graph layout (the reference's support functions read it, collision_gjk.py:153-176,
  # collision_primitive.py:141-146): numvert, numface, vert_edgeadr[numvert], vert_globalid[numvert],
  # edge_localid[numvert + 3 numface] (neighbour lists, each closed by -1), face_globalid[3 numface]
  graphs, adr = [], []
  for n in names:
    g = _hull_graph(meshes[n])
    adr.append(-1 if g is None else sum(len(x) for x in graphs))
    if g is not None:
      graphs.append(g)
  m.mesh_graphadr = np.array(adr, dtype=i32).reshape(m.nmesh)
  m.mesh_graph = np.concatenate(graphs).astype(i32) if graphs else np.zeros(0, dtype=i32)
  m.nmeshgraph = int(m.mesh_graph.size)
  m.mesh_polyadr = np.full(m.nmesh, -1, dtype=i32)

  # ---------------- body inertial properties
  m.body_mass = np.zeros(nbody)
  m.body_inertia = np.zeros((nbody, 3))
  m.body_ipos = np.zeros((nbody, 3))
  m.body_iquat = np.tile([1.0, 0, 0, 0], (nbody, 1))
  for b in range(1, nbody):
    inertial = bodies[b]["inertial"]
    if inertial is not None:
      m.body_mass[b] = inertial["mass"]
      m.body_inertia[b] = inertial["inertia"]
      m.body_ipos[b] = inertial["pos"]
      m.body_iquat[b] = inertial["quat"]
      continue
    sel = [
      g for g in bodies[b]["geom"] if comp.inertiagrouprange[0] <= m.geom_group[g] <= comp.inertiagrouprange[1] and geom_mass[g] > 0
    ]
    has_mesh = any(m.geom_type[g] == C.GeomType.MESH for g in sel)
    if len(sel) == 1 and not has_mesh:
      g = sel[0]
      m.body_mass[b] = geom_mass[g]
      m.body_inertia[b] = geom_inertia[g]
      m.body_ipos[b] = m.geom_pos[g]
      m.body_iquat[b] = m.geom_quat[g]
    elif len(sel) >= 1:
      mass = geom_mass[sel].sum()
      gcom = {g: m.geom_pos[g] + quat_to_mat(m.geom_quat[g]) @ geom_mesh_com[g] for g in sel}  # geom centre of mass in the body frame
      com = sum(geom_mass[g] * gcom[g] for g in sel) / mass
      tot = np.zeros((3, 3))
      for g in sel:
        R = quat_to_mat(m.geom_quat[g])
        d = gcom[g] - com
        local = geom_mesh_inertia[g] if m.geom_type[g] == C.GeomType.MESH else np.diag(geom_inertia[g])
        tot += R @ local @ R.T + geom_mass[g] * (np.dot(d, d) * np.eye(3) - np.outer(d, d))
      ev, q = eig3(tot)
      m.body_mass[b] = mass
      m.body_inertia[b] = ev
      m.body_ipos[b] = com
      m.body_iquat[b] = q
    if comp.boundmass > 0:
      m.body_mass[b] = max(m.body_mass[b], comp.boundmass)
    if comp.boundinertia > 0:
      m.body_inertia[b] = np.maximum(m.body_inertia[b], comp.boundinertia)
  m.body_subtreemass = m.body_mass.copy()
  for b in range(nbody - 1, 0, -1):
    m.body_subtreemass[m.body_parentid[b]] += m.body_subtreemass[b]

  # ---------------- sites
  m.nsite = nsite
  m.site_names = [s.get("name", f"site{i}") for i, s in enumerate(sites)]
  m.site_bodyid = np.array([s["_body"] for s in sites], dtype=i32)
  m.site_pos = np.array([_vec(s.get("pos"), 3, [0, 0, 0]) for s in sites], dtype=f64).reshape(nsite, 3)
  m.site_quat = np.array([_orientation(s, comp) for s in sites], dtype=f64).reshape(nsite, 4)

  _finalize_actuators(m, root, comp, defaults)
  _finalize_constraints(m, root, comp, defaults)
  _finalize_keyframes(m, root)
  _set_const(m)
  return m


def _finalize_actuators(m, root, comp, defaults):
  f64, i32 = np.float64, np.int32
  acts = []
  for sec in root.findall("actuator"):
    for e in sec:
      if e.tag not in _ACT_TAGS:
        continue
      a = _resolve(defaults, e, None, tags=("general", e.tag))
      a["_tag"] = e.tag
      acts.append(a)
  nu = len(acts)
  m.nu = nu
  m.actuator_names = [a.get("name", f"act{i}") for i, a in enumerate(acts)]
  m.actuator_trntype = np.zeros(nu, dtype=i32)
  m.actuator_dyntype = np.zeros(nu, dtype=i32)
  m.actuator_gaintype = np.zeros(nu, dtype=i32)
  m.actuator_biastype = np.zeros(nu, dtype=i32)
  m.actuator_trnid = np.full((nu, 2), -1, dtype=i32)
  m.actuator_actadr = np.full(nu, -1, dtype=i32)
  m.actuator_actnum = np.zeros(nu, dtype=i32)
  m.actuator_ctrllimited = np.zeros(nu, dtype=bool)
  m.actuator_forcelimited = np.zeros(nu, dtype=bool)
  m.actuator_actlimited = np.zeros(nu, dtype=bool)
  m.actuator_actearly = np.zeros(nu, dtype=bool)
  m.actuator_dynprm = np.zeros((nu, 10))
  m.actuator_gainprm = np.zeros((nu, 10))
  m.actuator_biasprm = np.zeros((nu, 10))
  m.actuator_ctrlrange = np.zeros((nu, 2))
  m.actuator_forcerange = np.zeros((nu, 2))
  m.actuator_actrange = np.zeros((nu, 2))
  m.actuator_gear = np.zeros((nu, 6))
  m.actuator_cranklength = np.zeros(nu)
  m.actuator_acc0 = np.zeros(nu)
  m.actuator_lengthrange = np.zeros((nu, 2))
  dyn = {"none": 0, "integrator": 1, "filter": 2, "filterexact": 3, "muscle": 4, "user": 5}
  gain = {"fixed": 0, "affine": 1, "muscle": 2, "user": 3}
  bias = {"none": 0, "affine": 1, "muscle": 2, "user": 3}
  na = 0
  for i, a in enumerate(acts):
    tag = a["_tag"]
    m.actuator_gainprm[i, 0] = 1.0
    m.actuator_dynprm[i, 0] = 1.0
    m.actuator_gear[i, 0] = 1.0
    if "joint" in a:
      m.actuator_trntype[i] = C.TrnType.JOINT
      m.actuator_trnid[i, 0] = m.jnt_names.index(a["joint"])
    elif "jointinparent" in a:
      m.actuator_trntype[i] = C.TrnType.JOINTINPARENT
      m.actuator_trnid[i, 0] = m.jnt_names.index(a["jointinparent"])
    else:
      raise NotImplementedError("actuator transmission other than joint is not supported by mjcf lite")
    if "gear" in a:
      v = _vec(a["gear"])
      m.actuator_gear[i, :] = 0
      m.actuator_gear[i, : v.size] = v
    if "dyntype" in a:
      m.actuator_dyntype[i] = dyn[a["dyntype"]]
    if "gaintype" in a:
      m.actuator_gaintype[i] = gain[a["gaintype"]]
    if "biastype" in a:
      m.actuator_biastype[i] = bias[a["biastype"]]
    for key, arr in (("dynprm", m.actuator_dynprm), ("gainprm", m.actuator_gainprm), ("biasprm", m.actuator_biasprm)):
      if key in a:
        v = _vec(a[key])
        arr[i, :] = 0
        arr[i, : v.size] = v
    if tag == "position":
      kp = float(a.get("kp", 1.0))
      kv = float(a.get("kv", 0.0))
      m.actuator_gainprm[i, 0] = kp
      m.actuator_biastype[i] = C.BiasType.AFFINE
      m.actuator_biasprm[i, :3] = [0, -kp, -kv]
      if "dampratio" in a or "timeconst" in a:
        raise NotImplementedError("position actuator dampratio/timeconst")
    elif tag == "velocity":
      kv = float(a.get("kv", 1.0))
      m.actuator_gainprm[i, 0] = kv
      m.actuator_biastype[i] = C.BiasType.AFFINE
      m.actuator_biasprm[i, :3] = [0, 0, -kv]
    elif tag == "damper":
      kv = float(a.get("kv", 1.0))
      m.actuator_gaintype[i] = C.GainType.AFFINE
      m.actuator_gainprm[i, :3] = [0, 0, -kv]
    elif tag not in ("motor", "general"):
      raise NotImplementedError(f"actuator shortcut <{tag}>")
    for key, rng, lim in (
      ("ctrl", m.actuator_ctrlrange, m.actuator_ctrllimited),
      ("force", m.actuator_forcerange, m.actuator_forcelimited),
      ("act", m.actuator_actrange, m.actuator_actlimited),
    ):
      if key + "range" in a:
        rng[i] = _vec(a[key + "range"], 2)
      lim[i] = _limited(a.get(key + "limited"), key + "range" in a, comp.autolimits)
    m.actuator_actearly[i] = _bool(a.get("actearly"))
    if m.actuator_dyntype[i] != C.DynType.NONE:
      m.actuator_actadr[i] = na
      m.actuator_actnum[i] = int(a.get("actdim", 1))
      na += m.actuator_actnum[i]
  m.na = na


def _finalize_constraints(m, root, comp, defaults):
  f64, i32 = np.float64, np.int32
  # equality
  eqs = []
  for sec in root.findall("equality"):
    for e in sec:
      a = _resolve(defaults, e, None)
      a["_tag"] = e.tag
      eqs.append(a)
  neq = len(eqs)
  m.neq = neq
  m.eq_type = np.zeros(neq, dtype=i32)
  m.eq_obj1id = np.full(neq, -1, dtype=i32)
  m.eq_obj2id = np.full(neq, -1, dtype=i32)
  m.eq_objtype = np.zeros(neq, dtype=i32)
  m.eq_active0 = np.ones(neq, dtype=bool)
  m.eq_solref = np.tile([0.02, 1.0], (neq, 1))
  m.eq_solimp = np.tile([0.9, 0.95, 0.001, 0.5, 2.0], (neq, 1))
  m.eq_data = np.zeros((neq, 11))
  for i, a in enumerate(eqs):
    tag = a["_tag"]
    m.eq_active0[i] = _bool(a.get("active"), True)
    for key, arr in (("solref", m.eq_solref), ("solimp", m.eq_solimp)):
      if key in a:
        v = _vec(a[key])
        arr[i, : v.size] = v
    if tag == "joint":
      m.eq_type[i] = C.EqType.JOINT
      m.eq_objtype[i] = 3  # mjOBJ_JOINT
      m.eq_obj1id[i] = m.jnt_names.index(a["joint1"])
      if "joint2" in a:
        m.eq_obj2id[i] = m.jnt_names.index(a["joint2"])
      pc = _vec(a.get("polycoef"), None, [0, 1, 0, 0, 0])
      m.eq_data[i, : pc.size] = pc
    elif tag in ("connect", "weld"):
      # body1 / body2 (world when absent); the qpos0-dependent half of eq_data is filled by _set_const like mj_setConst does
      m.eq_type[i] = C.EqType.CONNECT if tag == "connect" else C.EqType.WELD
      if "site1" in a:  # the site poses define the constraint frames (reference constraint.py:203-208, :868-877); eq_data is unused
        m.eq_objtype[i] = 6  # mjOBJ_SITE
        m.eq_obj1id[i] = m.site_names.index(a["site1"])
        m.eq_obj2id[i] = m.site_names.index(a["site2"])
      else:
        m.eq_objtype[i] = 1  # mjOBJ_BODY
        m.eq_obj1id[i] = m.body_names.index(a["body1"])
        m.eq_obj2id[i] = m.body_names.index(a["body2"]) if "body2" in a else 0
      m.eq_data[i, :3] = _vec(a.get("anchor"), 3, [0, 0, 0])
      if tag == "weld":
        rel = _vec(a.get("relpose"), 7, [0, 1, 0, 0, 0, 0, 0])
        if np.any(rel[3:] != 0):
          raise NotImplementedError("weld with an explicit relpose quaternion is not supported by mjcf lite yet")
        m.eq_data[i, 10] = float(a.get("torquescale", 1.0))
    else:
      raise NotImplementedError(f"equality <{tag}> is not supported by mjcf lite yet")

  # contact pairs / excludes
  pairs, excludes = [], []
  for sec in root.findall("contact"):
    for e in sec:
      if e.tag == "pair":
        pairs.append(_resolve(defaults, e, None))
      elif e.tag == "exclude":
        excludes.append(dict(e.attrib))
  m.nexclude = len(excludes)
  m.exclude_signature = np.array(
    [
      (min(m.body_names.index(x["body1"]), m.body_names.index(x["body2"])) << 16)
      + max(m.body_names.index(x["body1"]), m.body_names.index(x["body2"]))
      for x in excludes
    ],
    dtype=i32,
  )
  npair = len(pairs)
  m.npair = npair
  m.pair_dim = np.zeros(npair, dtype=i32)
  m.pair_geom1 = np.zeros(npair, dtype=i32)
  m.pair_geom2 = np.zeros(npair, dtype=i32)
  m.pair_solref = np.zeros((npair, 2))
  m.pair_solreffriction = np.zeros((npair, 2))
  m.pair_solimp = np.zeros((npair, 5))
  m.pair_margin = np.zeros(npair)
  m.pair_gap = np.zeros(npair)
  m.pair_friction = np.zeros((npair, 5))
  for i, p in enumerate(pairs):
    g1, g2 = m.geom_names.index(p["geom1"]), m.geom_names.index(p["geom2"])
    m.pair_geom1[i], m.pair_geom2[i] = g1, g2
    # unspecified pair parameters are mixed from the geoms (MuJoCo compiler behaviour)
    m.pair_dim[i] = int(p.get("condim", max(m.geom_condim[g1], m.geom_condim[g2])))
    mix = m.geom_solmix[g1] / max(m.geom_solmix[g1] + m.geom_solmix[g2], mjMINVAL)
    m.pair_solref[i] = _vec(p.get("solref"), None, mix * m.geom_solref[g1] + (1 - mix) * m.geom_solref[g2])
    m.pair_solimp[i] = mix * m.geom_solimp[g1] + (1 - mix) * m.geom_solimp[g2]
    if "solimp" in p:
      v = _vec(p["solimp"])
      m.pair_solimp[i, : v.size] = v
    m.pair_solreffriction[i] = _vec(p.get("solreffriction"), 2, [0, 0])
    m.pair_margin[i] = float(p.get("margin", max(m.geom_margin[g1], m.geom_margin[g2])))
    m.pair_gap[i] = float(p.get("gap", max(m.geom_gap[g1], m.geom_gap[g2])))
    fr = np.maximum(m.geom_friction[g1], m.geom_friction[g2])
    m.pair_friction[i] = [fr[0], fr[0], fr[1], fr[2], fr[2]]
    if "friction" in p:
      v = _vec(p["friction"])
      m.pair_friction[i, : v.size] = v

  # absent features (sizes the device model expects)
  m.ntendon = m.nflex = m.nsensor = m.nsensordata = m.ncam = m.nlight = 0
  m.nwrap = m.nJten = m.nflexvert = m.nflexedge = m.nflexelem = 0
  m.ngravcomp = int((m.body_gravcomp != 0).sum())


def _finalize_keyframes(m, root):
  keys = []
  for sec in root.findall("keyframe"):
    keys += list(sec.findall("key"))
  nkey = len(keys)
  m.nkey = nkey
  m.key_names = [k.get("name", f"key{i}") for i, k in enumerate(keys)]
  m.key_time = np.zeros(nkey)
  m.key_qpos = np.tile(m.qpos0, (nkey, 1)).reshape(nkey, m.nq)
  m.key_qvel = np.zeros((nkey, m.nv))
  m.key_act = np.zeros((nkey, m.na))
  m.key_ctrl = np.zeros((nkey, m.nu))
  for i, k in enumerate(keys):
    if "time" in k.attrib:
      m.key_time[i] = float(k.get("time"))
    for name, arr in (("qpos", m.key_qpos), ("qvel", m.key_qvel), ("act", m.key_act), ("ctrl", m.key_ctrl)):
      if name in k.attrib:
        arr[i] = _vec(k.get(name), arr.shape[1])


# ----------------------------------------------------------------------------- qpos0 constants
def host_kinematics(m, qpos):
  """Float64 forward kinematics (numpy, one world). Returns a namespace of frames."""
  nb = m.nbody
  xpos, xquat = np.zeros((nb, 3)), np.tile([1.0, 0, 0, 0], (nb, 1))
  xanchor, xaxis = np.zeros((m.njnt, 3)), np.zeros((m.njnt, 3))
  for b in range(1, nb):
    p = m.body_parentid[b]
    jadr, jnum = m.body_jntadr[b], m.body_jntnum[b]
    if jnum == 1 and m.jnt_type[jadr] == C.JointType.FREE:
      qa = m.jnt_qposadr[jadr]
      xpos[b] = qpos[qa : qa + 3]
      q = qpos[qa + 3 : qa + 7]
      xquat[b] = q / np.linalg.norm(q)
      xanchor[jadr], xaxis[jadr] = xpos[b], m.jnt_axis[jadr]
      continue
    pos = rot_vec_quat(m.body_pos[b], xquat[p]) + xpos[p]
    quat = quat_mul(xquat[p], m.body_quat[b])
    for j in range(jadr, jadr + jnum):
      qa = m.jnt_qposadr[j]
      anchor = rot_vec_quat(m.jnt_pos[j], quat) + pos
      axis = rot_vec_quat(m.jnt_axis[j], quat)
      jt = m.jnt_type[j]
      if jt == C.JointType.BALL:
        ql = qpos[qa : qa + 4]
        quat = quat_mul(quat, ql / np.linalg.norm(ql))
        pos = anchor - rot_vec_quat(m.jnt_pos[j], quat)
      elif jt == C.JointType.SLIDE:
        pos = pos + axis * (qpos[qa] - m.qpos0[qa])
      elif jt == C.JointType.HINGE:
        a = qpos[qa] - m.qpos0[qa]
        ql = np.concatenate([[math.cos(a / 2)], m.jnt_axis[j] * math.sin(a / 2)])
        quat = quat_mul(quat, ql)
        pos = anchor - rot_vec_quat(m.jnt_pos[j], quat)
      xanchor[j], xaxis[j] = anchor, axis
    xpos[b], xquat[b] = pos, quat / np.linalg.norm(quat)
  xmat = np.array([quat_to_mat(q) for q in xquat])
  xipos = np.array([xpos[b] + xmat[b] @ m.body_ipos[b] for b in range(nb)])
  ximat = np.array([quat_to_mat(quat_mul(xquat[b], m.body_iquat[b])) for b in range(nb)])
  return SimpleNamespace(xpos=xpos, xquat=xquat, xmat=xmat, xipos=xipos, ximat=ximat, xanchor=xanchor, xaxis=xaxis)


def host_mass_matrix(m, qpos):
  """Float64 dense joint-space inertia (CRB) plus the com-frame quantities used to build it."""
  k = host_kinematics(m, qpos)
  nb, nv = m.nbody, m.nv
  sub = k.xipos * m.body_mass[:, None]
  for b in range(nb - 1, 0, -1):
    sub[m.body_parentid[b]] += sub[b]
  subtree_com = np.where(m.body_subtreemass[:, None] > 0, sub / np.maximum(m.body_subtreemass[:, None], mjMINVAL), k.xipos)
  # 6x6 spatial inertia about the root subtree com (rot, lin ordering)
  I6 = np.zeros((nb, 6, 6))
  for b in range(1, nb):
    R = k.ximat[b]
    Ic = R @ np.diag(m.body_inertia[b]) @ R.T
    d = k.xipos[b] - subtree_com[m.body_rootid[b]]
    mass = m.body_mass[b]
    dx = np.array([[0, -d[2], d[1]], [d[2], 0, -d[0]], [-d[1], d[0], 0]])
    I6[b, :3, :3] = Ic - mass * dx @ dx
    I6[b, :3, 3:] = mass * dx
    I6[b, 3:, :3] = -mass * dx
    I6[b, 3:, 3:] = mass * np.eye(3)
  crb = I6.copy()
  for b in range(nb - 1, 0, -1):
    p = m.body_parentid[b]
    if p > 0:
      crb[p] += crb[b]
  cdof = np.zeros((nv, 6))
  for j in range(m.njnt):
    b, d0, jt = m.jnt_bodyid[j], m.jnt_dofadr[j], m.jnt_type[j]
    off = subtree_com[m.body_rootid[b]] - k.xanchor[j]
    if jt == C.JointType.FREE:
      cdof[d0 : d0 + 3, 3:] = np.eye(3)
      for a in range(3):
        ax = k.xmat[b][:, a]
        cdof[d0 + 3 + a] = np.concatenate([ax, np.cross(ax, off)])
    elif jt == C.JointType.BALL:
      for a in range(3):
        ax = k.xmat[b][:, a]
        cdof[d0 + a] = np.concatenate([ax, np.cross(ax, off)])
    elif jt == C.JointType.SLIDE:
      cdof[d0, 3:] = k.xaxis[j]
    else:
      cdof[d0] = np.concatenate([k.xaxis[j], np.cross(k.xaxis[j], off)])
  M = np.zeros((nv, nv))
  for i in range(nv):
    buf = crb[m.dof_bodyid[i]] @ cdof[i]
    M[i, i] = m.dof_armature[i] + cdof[i] @ buf
    j = m.dof_parentid[i]
    while j >= 0:
      M[i, j] = M[j, i] = cdof[j] @ buf
      j = m.dof_parentid[j]
  return M, SimpleNamespace(kin=k, subtree_com=subtree_com, cdof=cdof, crb=crb)


def _set_const(m):
  """qpos0-dependent constants (MuJoCo mj_setConst; reference io.py:1750-1958, 2078)."""
  nv = m.nv
  m.stat = SimpleNamespace(meaninertia=1.0)
  m.dof_invweight0 = np.zeros(nv)
  m.body_invweight0 = np.zeros((m.nbody, 2))
  if nv == 0:
    return
  M, aux = host_mass_matrix(m, m.qpos0)
  m.stat.meaninertia = float(np.mean(np.diag(M)))
  Minv = np.linalg.inv(M)
  A = np.diag(Minv).copy()
  for j in range(m.njnt):
    d0, jt = m.jnt_dofadr[j], m.jnt_type[j]
    if jt == C.JointType.FREE:
      m.dof_invweight0[d0 : d0 + 3] = A[d0 : d0 + 3].mean()
      m.dof_invweight0[d0 + 3 : d0 + 6] = A[d0 + 3 : d0 + 6].mean()
    elif jt == C.JointType.BALL:
      m.dof_invweight0[d0 : d0 + 3] = A[d0 : d0 + 3].mean()
    else:
      m.dof_invweight0[d0] = A[d0]
  for b in range(1, m.nbody):
    if m.body_weldid[b] == 0:
      continue
    off = aux.kin.xipos[b] - aux.subtree_com[m.body_rootid[b]]
    J = np.zeros((6, nv))
    bb = b
    while bb > 0 and m.body_dofnum[bb] == 0:
      bb = m.body_parentid[bb]
    d = m.body_dofadr[bb] + m.body_dofnum[bb] - 1 if bb > 0 else -1
    while d >= 0:
      J[:3, d] = aux.cdof[d, 3:] + np.cross(aux.cdof[d, :3], off)
      J[3:, d] = aux.cdof[d, :3]
      d = m.dof_parentid[d]
    Ab = np.diag(J @ Minv @ J.T)
    tr, rot = Ab[:3].mean(), Ab[3:].mean()
    if tr < mjMINVAL and rot > mjMINVAL:
      tr = rot
    elif rot < mjMINVAL and tr > mjMINVAL:
      rot = tr
    m.body_invweight0[b] = [tr, rot]
  # connect / weld: anchors and relative pose at qpos0 (MuJoCo engine_setconst.c set0)
  k = aux.kin
  for i in range(m.neq):
    b1, b2 = int(m.eq_obj1id[i]), int(m.eq_obj2id[i])
    if m.eq_objtype[i] != 1:
      continue
    if m.eq_type[i] == C.EqType.CONNECT:  # data[0:3] anchor in body1, data[3:6] the same point in body2
      world = k.xpos[b1] + k.xmat[b1] @ m.eq_data[i, :3]
      m.eq_data[i, 3:6] = k.xmat[b2].T @ (world - k.xpos[b2])
    elif m.eq_type[i] == C.EqType.WELD:  # data[0:3] anchor in body2, data[3:6] the same point in body1, data[6:10] = q1^-1 q2
      world = k.xpos[b2] + k.xmat[b2] @ m.eq_data[i, :3]
      m.eq_data[i, 3:6] = k.xmat[b1].T @ (world - k.xpos[b1])
      q1 = k.xquat[b1] * np.array([1.0, -1.0, -1.0, -1.0])
      m.eq_data[i, 6:10] = quat_mul(q1, k.xquat[b2])
  # actuator_acc0 = |M^-1 moment| for joint transmissions
  for i in range(m.nu):
    j = m.actuator_trnid[i, 0]
    mom = np.zeros(nv)
    d0 = m.jnt_dofadr[j]
    if m.jnt_type[j] in (C.JointType.HINGE, C.JointType.SLIDE):
      mom[d0] = m.actuator_gear[i, 0]
    elif m.jnt_type[j] == C.JointType.BALL:
      mom[d0 : d0 + 3] = m.actuator_gear[i, :3]
    else:
      mom[d0 : d0 + 6] = m.actuator_gear[i]
    m.actuator_acc0[i] = np.linalg.norm(Minv @ mom)
