"""Host-side logic (no GPU): MJCF-lite compiler, put_model derived tables, make_data sizing and argument
validation (same exception types as the reference, io.py:100-144 and :893-918), put_data /
get_data_into / reset_data on host tensors."""

import numpy as np
import pytest
import torch

import mujoco_warp_b200 as mjw
from mujoco_warp_b200 import constants as C
from mujoco_warp_b200 import io as mio
from mujoco_warp_b200 import types


def test_humanoid_sizes(humanoid):
  m = humanoid
  assert (m.nq, m.nv, m.nu, m.nbody, m.njnt, m.ngeom) == (28, 27, 21, 17, 22, 20)
  assert m.nkey >= 1 and np.asarray(m.key_qpos).shape[1] == 28
  assert abs(float(np.sum(m.body_mass)) - 40.84) < 0.01


def test_put_model_derived_tables(humanoid):
  m = mjw.put_model(humanoid, device="cpu")
  s = m.sizes
  assert s["nv_pad"] == 28 and s["nlevel"] == 7
  assert s["nxn"] == 161  # 190 geom pairs - same body - parent/child (reference io.py:270-358)
  hm = m._host
  # DFS pre-order: every body's subtree is the id range [b, subtreeend[b])
  end = hm.ints["body_subtreeend"]
  parent = hm.ints["body_parentid"]
  for b in range(1, s["nbody"]):
    assert parent[b] < b and end[parent[b]] >= end[b] > b
  # tree levels partition the bodies
  tree = hm.ints["body_tree"]
  assert sorted(tree.tolist()) == list(range(s["nbody"]))
  # tensors are float32 / int32 views with the reference's logical shapes
  assert m.body_mass.shape == (1, 17) and m.body_mass.dtype == torch.float32
  assert m.geom_size.shape == (1, 20, 3)
  assert m.jnt_type.dtype == torch.int32
  assert float(m.opt.tolerance) >= np.float32(1e-6)  # io.py: tolerance floored for float32


def test_is_sparse_and_padding_rules(humanoid):
  assert not mio.is_sparse(humanoid)
  assert mio._get_padded_sizes(27, 0, False, 16)[1] == 28
  assert mio._get_padded_sizes(40, 0, False, 16)[1] == 48


def test_make_data_shapes_and_defaults(humanoid):
  d = mjw.make_data(humanoid, nworld=3, nconmax=24, njmax=64, device="cpu")
  assert d.nworld == 3 and d.njmax == 64 and d.naconmax == 72
  assert d.qpos.shape == (3, 28) and d.qvel.shape == (3, 27) and d.ctrl.shape == (3, 21)
  assert d.xquat.shape == (3, 17, 4) and d.xmat.shape == (3, 17, 3, 3) and d.cdof.shape == (3, 27, 6)
  assert d.qM.shape == (3, 28, 28) and d.qLD.shape == (3, 27, 27)
  assert d.efc.J.shape == (3, 64, 28) and d.efc.force.shape == (3, 64) and d.efc.state.dtype == torch.int32
  assert d.contact.dist.shape == (72,) and d.contact.frame.shape == (72, 3, 3) and d.contact.geom.shape == (72, 2)
  assert d.nefc.shape == (3,) and d.nacon.shape == (1,)
  np.testing.assert_allclose(d.qpos[0].numpy(), np.asarray(humanoid.qpos0), rtol=1e-6)
  assert float(d.xquat[1, 5, 0]) == 1.0  # identity quaternions
  # the flat-name accessor and the nested views alias the same storage
  assert d.field("efc_J").data_ptr() == d.efc.J.data_ptr()
  assert d.field("contact_dist").data_ptr() == d.contact.dist.data_ptr()


@pytest.mark.parametrize(
  "kwargs,msg",
  [
    ({"nworld": 0}, "nworld"),
    ({"nconmax": -1}, "nconmax"),
    ({"njmax": -1}, "njmax"),
    ({"naconmax": -1}, "naconmax"),
    ({"nccdmax": 30, "nconmax": 24}, "ccdmax"),
    ({"naccdmax": 9999, "naconmax": 10}, "naccdmax"),
  ],
)
def test_make_data_rejects_bad_sizes(humanoid, kwargs, msg):
  args = dict(nworld=2, nconmax=24, njmax=64, device="cpu")
  args.update(kwargs)
  with pytest.raises(ValueError, match=msg):
    mjw.make_data(humanoid, **args)


UNSUPPORTED = [
  ('<option solver="PGS"/>', "PGS"),
  ('<option integrator="implicit"/>', "IMPLICIT"),
  ('<option noslip_iterations="2"/>', "noslip"),
]


@pytest.mark.parametrize("snippet,msg", UNSUPPORTED)
def test_put_model_rejects_unsupported_features(snippet, msg):
  xml = f"""<mujoco>{snippet}<worldbody><body><joint type="slide"/><geom type="sphere" size=".1"/></body></worldbody></mujoco>"""
  mjm = mjw.load_mjcf_string(xml)
  with pytest.raises(NotImplementedError, match=msg):
    mjw.put_model(mjm, device="cpu")


def test_sparse_eligible_models_are_accepted_and_stored_dense():
  """Models the reference stores sparse (jacobian=sparse, or nv > 32 under auto: io.py:67) keep the dense shapes of qM / qLD / efc.J."""
  xml = """<mujoco><option jacobian="sparse"/><worldbody><body><joint type="slide"/><geom type="sphere" size=".1"/></body></worldbody></mujoco>"""
  mjm = mjw.load_mjcf_string(xml)
  assert mio.is_sparse(mjm)
  m = mjw.put_model(mjm, device="cpu")
  assert m.is_sparse is False
  d = mjw.make_data(mjm, nworld=2, nconmax=4, njmax=8, device="cpu")
  assert tuple(d.qM.shape) == (2, m.sizes["nv_pad"], m.sizes["nv_pad"]) and d.efc.J.dim() == 3 and d.efc.J.shape[2] == m.sizes["nv_pad"]


def test_replicate_attach_frame_are_expanded():
  """benchmarks/humanoid/three_humanoids.xml: <replicate> of a <frame> holding an <attach> of humanoid.xml's torso.  Three copies of the
  humanoid's tree, defaults, actuators; copy i rotated by i * 16.36 degrees about z at radius 4."""
  import os

  from tests import helpers as H

  one = mjw.load_mjcf(H.HUMANOID)
  three = mjw.load_mjcf(os.path.join(os.path.dirname(H.HUMANOID), "three_humanoids.xml"))
  assert (three.nv, three.nq, three.nu, three.nbody, three.ngeom) == (3 * one.nv, 3 * one.nq, 3 * one.nu, 3 * (one.nbody - 1) + 1, 3 * (one.ngeom - 1) + 1)
  assert mio.is_sparse(three) and not mio.is_sparse(one)
  nb = one.nbody - 1
  for i in range(3):
    a = np.deg2rad(16.36 * i)
    np.testing.assert_allclose(three.body_pos[1 + nb * i], [4 * np.sin(a), -4 * np.cos(a), 1.282], atol=1e-12)
    np.testing.assert_allclose(three.body_quat[1 + nb * i], [np.cos(a / 2), 0, 0, np.sin(a / 2)], atol=1e-12)
    np.testing.assert_allclose(three.body_mass[1 + nb * i : 1 + nb * (i + 1)], one.body_mass[1:], rtol=1e-12)
    np.testing.assert_allclose(three.dof_damping[one.nv * i : one.nv * (i + 1)], one.dof_damping)
    np.testing.assert_allclose(three.dof_invweight0[one.nv * i : one.nv * (i + 1)], one.dof_invweight0, rtol=1e-9)
    np.testing.assert_allclose(three.actuator_gear[one.nu * i : one.nu * (i + 1)], one.actuator_gear)
    np.testing.assert_array_equal(three.actuator_trnid[one.nu * i : one.nu * (i + 1), 0], one.actuator_trnid[:, 0] + one.njnt * i)
    np.testing.assert_allclose(three.jnt_range[one.njnt * i : one.njnt * (i + 1)], one.jnt_range)
  assert len(set(three.names_body)) == three.nbody if hasattr(three, "names_body") else True
  # an offset replicate of plain geoms, and a nested frame
  xml = """<mujoco><worldbody>
    <replicate count="3" offset="1 0 0" sep="_"><geom name="g" type="sphere" size=".1" pos="0 2 0"/></replicate>
    <frame pos="0 0 1" euler="0 0 90"><frame pos="1 0 0"><body name="b" pos="0 1 0"><joint type="slide"/><geom type="sphere" size=".1"/></body></frame></frame>
  </worldbody></mujoco>"""
  mjm = mjw.load_mjcf_string(xml)
  np.testing.assert_allclose(mjm.geom_pos[:3], [[0, 2, 0], [1, 2, 0], [2, 2, 0]], atol=1e-12)
  np.testing.assert_allclose(mjm.body_pos[1], [-1, 1, 1], atol=1e-12)  # R_z(90) (1, 1, 0) + (0, 0, 1)


def test_mesh_geoms_without_assets_are_rejected_and_every_pair_is_accepted():
  """A colliding mesh geom needs its asset (hull vertices) and its body an explicit <inertial>: rejected up front, like the
  reference rejects what it does not implement (io.py:100-144).  All pairs of plane / sphere / capsule / ellipsoid /
  cylinder / box / mesh have a narrowphase (primitive or GJK / EPA)."""
  xml = """<mujoco><worldbody>
    <body><freejoint/><geom type="mesh" mesh="m"/></body>
    <body pos="1 0 0"><freejoint/><geom type="sphere" size=".1"/></body>
  </worldbody></mujoco>"""
  with pytest.raises(NotImplementedError, match="mesh"):
    mjw.load_mjcf_string(xml)
  geoms = ['type="sphere" size=".1"', 'type="capsule" size=".1 .2"', 'type="ellipsoid" size=".1 .2 .3"', 'type="cylinder" size=".1 .2"',
           'type="box" size=".1 .2 .3"', 'type="mesh" mesh="t"']  # fmt: skip
  inertial = '<inertial pos="0 0 0" mass="1" diaginertia=".1 .1 .1"/>'
  bodies = "".join(f'<body pos="{i} 0 1">{inertial}<joint type="slide"/><geom {g}/></body>' for i, g in enumerate(geoms))
  asset = '<asset><mesh name="t" vertex="0 0 0 1 0 0 0 1 0 0 0 1 .1 .1 .1"/></asset>'
  mjm = mjw.load_mjcf_string(f'<mujoco>{asset}<worldbody><geom type="plane" size="1 1 .1"/>{bodies}</worldbody></mujoco>')
  assert mjm.nmesh == 1 and mjm.mesh_vertnum.tolist() == [4]  # the interior point is not a hull vertex
  m = mjw.put_model(mjm, device="cpu")
  assert m.sizes["nxn"] == 21 and m.sizes["nmesh"] == 1 and m.sizes["nmeshvert"] == 4
  np.testing.assert_allclose(m.mesh_vert.numpy().reshape(-1, 3), mjm.mesh_vert, rtol=1e-6)
  assert int(m.geom_dataid[6]) == 0 and int(m.geom_dataid[1]) == -1


def test_mesh_mass_properties_come_from_the_convex_hull():
  """Bodies without <inertial> get mass, centre of mass and principal inertia from their mesh geoms (MuJoCo mesh inertia="convex":
  exact polyhedral integration over the hull): an off-centre, rotated box given as a vertex cloud equals the analytic box."""
  cube = np.array([[x, y, z] for x in (-0.1, 0.3) for y in (-0.2, 0.2) for z in (0.0, 0.6)])
  verts = " ".join(f"{v:.6f}" for v in cube.ravel())
  xml = f"""<mujoco><asset><mesh name="b" vertex="{verts}"/></asset><worldbody>
    <body><freejoint/><geom type="mesh" mesh="b" density="500" pos="0.5 0 0" euler="0 0 90"/></body>
    <body pos="2 0 0"><freejoint/><geom type="box" size=".2 .2 .3" density="500"/></body></worldbody></mujoco>"""
  mjm = mjw.load_mjcf_string(xml)
  np.testing.assert_allclose(mjm.body_mass[1:], [500 * 0.096, 500 * 0.096], rtol=1e-9)
  np.testing.assert_allclose(mjm.body_ipos[1], [0.5, 0.1, 0.3], atol=1e-9)  # hull centre (0.1, 0, 0.3) rotated 90 deg about z
  np.testing.assert_allclose(np.sort(mjm.body_inertia[1]), np.sort(mjm.body_inertia[2]), rtol=1e-9)
  tet = mjw.load_mjcf_string("""<mujoco><asset><mesh name="t" vertex="0 0 0 1 0 0 0 1 0 0 0 1"/></asset><worldbody>
    <body><freejoint/><geom type="mesh" mesh="t" mass="3"/></body></worldbody></mujoco>""")
  np.testing.assert_allclose(tet.body_mass[1], 3.0)
  np.testing.assert_allclose(tet.body_ipos[1], [0.25, 0.25, 0.25], atol=1e-12)
  # unit right tetrahedron about its centroid: I = m * [[3/40 ..]] with eigenvalues m * (3/80 * 2 ... ) -> check the trace and positivity
  assert abs(tet.body_inertia[1].sum() - 3.0 * 2 * 3 * (1 / 10 - 1 / 16)) < 1e-9 and (tet.body_inertia[1] > 0).all()


def test_mesh_files_stl_and_obj_load_as_convex_hull_vertices(tmp_path):
  """Binary STL, ASCII STL and OBJ assets: vertices deduplicated to their convex hull, `scale` applied, bounds from the hull."""
  import struct

  cube = np.array([[x, y, z] for x in (-1, 1) for y in (-1, 1) for z in (-1, 1)], dtype=np.float64)
  faces = [(0, 1, 3), (0, 3, 2), (4, 6, 7), (4, 7, 5), (0, 4, 5), (0, 5, 1), (2, 3, 7), (2, 7, 6), (0, 2, 6), (0, 6, 4), (1, 5, 7), (1, 7, 3)]
  with open(tmp_path / "cube.stl", "wb") as f:
    f.write(b"\0" * 80 + struct.pack("<I", len(faces)))
    for tri in faces:
      f.write(struct.pack("<12fH", 0, 0, 0, *cube[list(tri)].ravel(), 0))
  with open(tmp_path / "cube_ascii.stl", "w") as f:
    f.write("solid cube\n")
    for tri in faces:
      f.write(" facet normal 0 0 0\n  outer loop\n" + "".join(f"   vertex {v[0]} {v[1]} {v[2]}\n" for v in cube[list(tri)]) + "  endloop\n endfacet\n")
    f.write("endsolid cube\n")
  with open(tmp_path / "cube.obj", "w") as f:
    f.write("# cube with an interior point\n" + "".join(f"v {v[0]} {v[1]} {v[2]}\n" for v in np.vstack([cube, [[0, 0, 0]]])))
    f.write("".join(f"f {a + 1} {b + 1} {c + 1}\n" for a, b, c in faces))
  for fname in ("cube.stl", "cube_ascii.stl", "cube.obj"):
    xml = f"""<mujoco><compiler meshdir="{tmp_path}"/><asset><mesh name="c" file="{fname}" scale=".1 .2 .3"/></asset><worldbody>
      <body><inertial pos="0 0 0" mass="1" diaginertia=".1 .1 .1"/><freejoint/><geom type="mesh" mesh="c"/></body></worldbody></mujoco>"""
    mjm = mjw.load_mjcf_string(xml)
    assert mjm.mesh_vertnum.tolist() == [8], fname
    got = mjm.mesh_vert[np.lexsort(mjm.mesh_vert.T[::-1])]
    want = cube * [0.1, 0.2, 0.3]
    np.testing.assert_allclose(got, want[np.lexsort(want.T[::-1])], atol=1e-7, err_msg=fname)
    np.testing.assert_allclose(mjm.geom_rbound[0], np.linalg.norm([0.1, 0.2, 0.3]), rtol=1e-6)


def test_device_is_mandatory_for_compute(humanoid):
  if torch.cuda.is_available():
    pytest.skip("a CUDA device is present")
  with pytest.raises(RuntimeError, match="CUDA"):
    mjw.put_model(humanoid)
  m = mjw.put_model(humanoid, device="cpu")
  d = mjw.make_data(humanoid, nworld=1, device="cpu")
  with pytest.raises(RuntimeError, match="no CPU fallback"):
    mjw.step(m, d)


def test_put_data_tiles_host_state_and_get_data_into_reads_it_back(humanoid):
  hd = mjw.make_host_data(humanoid)
  hd.qpos[:] = np.asarray(humanoid.key_qpos[0])
  hd.qvel[:] = np.linspace(-1, 1, humanoid.nv)
  hd.ctrl[:] = 0.25
  hd.time = 1.5
  d = mjw.put_data(humanoid, hd, nworld=4, nconmax=24, njmax=64, device="cpu")
  for w in range(4):
    np.testing.assert_allclose(d.qpos[w].numpy(), hd.qpos, rtol=1e-6)
    np.testing.assert_allclose(d.qvel[w].numpy(), hd.qvel, rtol=1e-6)
  assert float(d.time[2]) == 1.5
  d.qvel[3] += 1.0
  out = mjw.make_host_data(humanoid)
  mjw.get_data_into(out, humanoid, d, world_id=3)
  np.testing.assert_allclose(out.qvel, hd.qvel + 1.0, rtol=1e-6)
  assert out.qvel.dtype == np.float64 and out.nefc == 0 and out.ncon == 0


def test_reset_data_needs_the_device(humanoid):
  """reset_data is a CUDA kernel behind the C ABI (mjb_reset_data); like `step` it has no CPU path (tests/test_gpu_support.py)."""
  m = mjw.put_model(humanoid, device="cpu")
  d = mjw.make_data(humanoid, nworld=4, nconmax=24, njmax=64, device="cpu")
  with pytest.raises(RuntimeError, match="no CPU fallback"):
    mjw.reset_data(m, d, reset=np.array([True, False, True, False]))


def test_override_model_accepts_reference_style_keys(humanoid):
  m = mjw.put_model(humanoid, device="cpu")
  mjw.override_model(m, {"opt.solver": "cg", "opt.iterations": 7, "opt.ls_iterations": 3})
  assert int(m.opt.solver) == int(C.SolverType.CG) and int(m.opt.iterations) == 7 and int(m.opt.ls_iterations) == 3


def test_field_tables_cover_every_dataclass_field():
  for name in types.DATA_REAL_FIELDS:
    assert name not in types.DATA_INT_FIELDS
  assert "qpos" in types.DATA_REAL_FIELDS and "nefc" in types.DATA_INT_FIELDS
  # Contact / Constraint mirrors of the reference dataclasses (types.py:1618, :1659)
  import dataclasses

  assert {"dist", "pos", "frame", "includemargin", "friction", "solref", "solreffriction", "solimp", "dim", "geom", "efc_address",
          "worldid"} <= {f.name for f in dataclasses.fields(types.Contact)}  # fmt: skip
  assert {"J", "pos", "margin", "D", "vel", "aref", "frictionloss", "force", "state", "type", "id"} <= {
    f.name for f in dataclasses.fields(types.Constraint)
  }


def test_connect_and_weld_equalities_compile_their_qpos0_data():
  """MJCF-lite fills the qpos0-dependent half of eq_data like mj_setConst: connect stores the anchor in both body frames, weld the
  anchor in both frames plus the relative orientation q1^-1 q2; site-based equalities reference site ids (mjOBJ_SITE = 6)."""
  from tests import models as TM

  mjm = mjw.load_mjcf_string(TM.SMALL_MODELS["connect_weld"])
  assert mjm.eq_type.tolist() == [0, 1, 0, 1, 1, 2] and mjm.eq_objtype.tolist() == [1, 1, 1, 1, 1, 3]
  assert mjm.eq_active0.tolist() == [True, True, True, True, False, True]
  k = mjw.mjcf.host_kinematics(mjm, mjm.qpos0)
  for i in range(mjm.neq):
    b1, b2, data = int(mjm.eq_obj1id[i]), int(mjm.eq_obj2id[i]), mjm.eq_data[i]
    if mjm.eq_type[i] == 0:  # the two anchors are the same world point at qpos0
      np.testing.assert_allclose(k.xpos[b1] + k.xmat[b1] @ data[:3], k.xpos[b2] + k.xmat[b2] @ data[3:6], atol=1e-12)
    elif mjm.eq_type[i] == 1:
      np.testing.assert_allclose(k.xpos[b1] + k.xmat[b1] @ data[3:6], k.xpos[b2] + k.xmat[b2] @ data[:3], atol=1e-12)
      np.testing.assert_allclose(mjw.mjcf.quat_mul(k.xquat[b1], data[6:10]), k.xquat[b2], atol=1e-12)
  assert mjm.eq_data[3, 10] == 0.5 and mjm.eq_data[1, 10] == 1.0
  m = mjw.put_model(mjm, device="cpu")
  assert m.sizes["neqconnect"] == 2 and m.sizes["neqweld"] == 3 and m.sizes["neqjnt"] == 1
  assert m.eq_connect_adr.tolist() == [0, 2] and m.eq_wld_adr.tolist() == [1, 3, 4]
  sites = mjw.load_mjcf_string(TM.SMALL_MODELS["site_equalities"])
  assert sites.eq_objtype.tolist() == [6, 6, 6, 6] and sites.eq_obj1id.tolist() == [sites.site_names.index(n) for n in ("a2", "c1", "b2", "b2")]
  with pytest.raises(NotImplementedError, match="relpose"):
    mjw.load_mjcf_string('<mujoco><worldbody><body name="a"><freejoint/><geom size=".1"/></body></worldbody>'
                         '<equality><weld body1="a" relpose="0 0 0 1 0 0 0"/></equality></mujoco>')


def test_blob_fields_cannot_be_rebound_and_batched_options_are_rejected(humanoid):
  """The kernels read the packed blob and one option value for all worlds: the reference's per-world randomisation idiom
  (`m.geom_friction = <(nworld, ...)>`, `m.opt.timestep = <(nworld,)>`, indexed there as [worldid % shape[0]]) must fail loudly."""
  from mujoco_warp_b200 import _lib

  m = mjw.put_model(humanoid, device="cpu")
  with pytest.raises(NotImplementedError, match="in place"):
    m.geom_friction = torch.zeros(8, humanoid.ngeom, 3)
  with pytest.raises(NotImplementedError, match="in place"):
    m.dof_damping = m.dof_damping.clone()
  m.geom_friction[0, 1, 0] = 0.5  # in-place edits stay legal and land in the blob
  off = m._offsets["geom_friction"]
  assert float(m._blob[off + 12 : off + 16].view(torch.float32)[0]) == 0.5
  assert _lib.make_option(m).timestep == pytest.approx(float(m.opt.timestep[0]))
  m.opt.timestep = torch.full((8,), 0.002)
  with pytest.raises(NotImplementedError, match="per-world Option.timestep"):
    _lib.make_option(m)


def test_data_field_numel_matches_make_data(humanoid):
  d = mjw.make_data(humanoid, nworld=3, nconmax=24, njmax=64, device="cpu")
  m = mjw.put_model(humanoid, device="cpu")
  sizes = dict(m.sizes)
  sizes.update(d._sizes)
  for n in list(types.DATA_INT_FIELDS) + list(types.DATA_REAL_FIELDS):
    assert types.data_field_numel(n, sizes) == d.field(n).numel(), n


def test_model_cache_round_trip(tmp_path):
  """save_model / load_model: the .npz model cache restores a host model whose every kernel table is identical."""
  import os
  import time

  from mujoco_warp_b200 import io as _io

  path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "benchmarks", "aloha_pot", "scene.xml")
  t0 = time.time()
  mjm = mjw.load_mjcf(path)
  t_compile = time.time() - t0
  f = str(tmp_path / "aloha.npz")
  mjw.save_model(mjm, f)
  t0 = time.time()
  back = mjw.load_model(f)
  t_load = time.time() - t0
  a, b = _io.host_model(mjm), _io.host_model(back)
  assert a.sizes == b.sizes
  for k in a.ints:
    np.testing.assert_array_equal(a.ints[k], b.ints[k], err_msg=k)
  for k in a.reals:
    np.testing.assert_array_equal(a.reals[k], b.reals[k], err_msg=k)
  assert back.key_names == mjm.key_names and float(back.opt.timestep) == float(mjm.opt.timestep)
  np.testing.assert_array_equal(back.keyframe("lift_pot0").qpos, mjm.keyframe("lift_pot0").qpos)
  assert t_load < t_compile
