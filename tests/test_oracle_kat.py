"""Known-answer tests that pin the oracle against the literal expectations held by the reference's own
tests (the only fixtures that exist without MuJoCo C): math_test.py:27-126, broadphase_test.py:49-236,
collision_driver_test.py:882-972, constraint_test.py:270-312.  The scenes are restated here as MJCF."""

import ctypes

import numpy as np
import pytest

import mujoco_warp_b200 as mjw
from mujoco_warp_b200 import constants as C
from oracle import oracle as O

P = lambda x: x.ctypes.data_as(ctypes.POINTER(ctypes.c_double))  # noqa: E731


def seg_seg(a0, a1, b0, b1):
  lib = O.load()
  arr = [np.array(x, dtype=np.float64) for x in (a0, a1, b0, b1)]
  oa, ob = np.zeros(3), np.zeros(3)
  lib.mjo_closest_segment_to_segment_points(P(arr[0]), P(arr[1]), P(arr[2]), P(arr[3]), P(oa), P(ob))
  return oa, ob


# (a0, a1, b0, b1, best_a, best_b, decimals): reference math_test.py:30-95
SEGMENT_CASES = [
  ([0.73432405, 0.12372768, 0.20272314], [1.10600128, 0.88555209, 0.65209485], [0.85599262, 0.61736299, 0.9843583],
   [1.84270939, 0.92891793, 1.36343326], [1.09063, 0.85404, 0.63351], [0.99596, 0.66156, 1.03813], 5),
  ([0, 0, -1], [0, 0, 1], [-1, 0, 0], [1, 0, 0], [0, 0, 0], [0, 0, 0], 5),
  ([0.2, 0.2, 0], [1, 1, 0], [0.2, 0.4, 0], [1, 2, 0], [0.3, 0.3, 0], [0.2, 0.4, 0], 2),
  ([0, 0, -1], [0, 0, 1], [1, 0, -1], [1, 0, 1], [0, 0, 0], [1, 0, 0], 5),
  ([0, 0, -1], [0, 0, 1], [1, 0, 1], [1, 0, 3], [0, 0, 1], [1, 0, 1], 5),
  ([0, 0, -1], [0, 0, -1], [1, 0, 0.1], [1, 0, 0.1], [0, 0, -1], [1, 0, 0.1], 5),
  ([0, 0, -1], [0, 0, 1], [0, 0, -1], [0, 0, 1], [0, 0, 0], [0, 0, 0], 5),
]  # fmt: skip


@pytest.mark.parametrize("case", SEGMENT_CASES)
def test_closest_segment_to_segment_points(case):
  a0, a1, b0, b1, ea, eb, dec = case
  oa, ob = seg_seg(a0, a1, b0, b1)
  assert not np.isnan(oa).any() and not np.isnan(ob).any()
  np.testing.assert_allclose(oa, ea, atol=1.5 * 10.0**-dec)
  np.testing.assert_allclose(ob, eb, atol=1.5 * 10.0**-dec)


def test_upper_triangular_indices():
  lib = O.load()
  assert [lib.mjo_upper_tri_index(2, i, j) for i in range(2) for j in range(i + 1, 2)] == list(range(1))
  assert [lib.mjo_upper_tri_index(10, i, j) for i in range(10) for j in range(i + 1, 10)] == list(range(45))
  assert [lib.mjo_upper_trid_index(1, i, j) for i in range(1) for j in range(i, 1)] == list(range(1))
  assert [lib.mjo_upper_trid_index(10, i, j) for i in range(10) for j in range(i, 10)] == list(range(55))
  assert lib.mjo_upper_trid_index(10, 1, 5) == lib.mjo_upper_trid_index(10, 5, 1)


# ---- broadphase counts (reference broadphase_test.py:55-187) -----------------------------------
BROADPHASE_XML = """
<mujoco>
  <worldbody>
    <body><freejoint/><geom type="sphere" size="0.1"/></body>
    <body><freejoint/><geom type="sphere" size="0.1"/></body>
    <body><freejoint/><geom type="capsule" size="0.1 0.1"/></body>
    <body><freejoint/><geom type="sphere" size="0.1"/></body>
    <body>
      <freejoint/>
      <geom type="sphere" size="0.1"/>
      <geom type="sphere" size="0.1"/>
      <body><geom type="sphere" size="0.1"/><joint type="hinge"/></body>
    </body>
  </worldbody>
  <keyframe>
    <key qpos='0 0 0 1 0 0 0  1 0 0 1 0 0 0  2 0 0 1 0 0 0  3 0 0 1 0 0 0  4 0 0 1 0 0 0  0'/>
    <key qpos='0 0 0 1 0 0 0  .05 0 0 1 0 0 0  2 0 0 1 0 0 0  3 0 0 1 0 0 0  4 0 0 1 0 0 0  0'/>
    <key qpos='0 0 0 1 0 0 0  .01 0 0 1 0 0 0  .02 0 0 1 0 0 0  3 0 0 1 0 0 0  4 0 0 1 0 0 0  0'/>
    <key qpos='0 0 0 1 0 0 0  1 0 0 1 0 0 0  2 0 0 1 0 0 0  2 0 0 1 0 0 0  4 0 0 1 0 0 0  0'/>
  </keyframe>
</mujoco>
"""
BF = C.BroadphaseFilter
FILTERS = [BF.PLANE | BF.SPHERE, BF.PLANE | BF.AABB, BF.PLANE | BF.OBB, BF.PLANE | BF.SPHERE | BF.AABB, BF.PLANE | BF.SPHERE | BF.OBB,
           BF.PLANE | BF.SPHERE | BF.AABB | BF.OBB]  # fmt: skip


def run_collision(mjm, qpos, opt=None, nconmax=64, njmax=64, stages=("kinematics", "com_pos", "collision")):
  om = O.OracleModel(mjm)
  qpos = np.atleast_2d(np.asarray(qpos, dtype=np.float64))
  d = O.make_data(om, nworld=qpos.shape[0], nconmax=nconmax, njmax=njmax)
  d.qpos[:] = qpos
  bits = 0
  for s in stages:
    bits |= O.STAGE[s]
  O.run(om, d, bits, opt=opt)
  return om, d


@pytest.mark.parametrize("broadphase", [C.BroadphaseType.NXN, C.BroadphaseType.SAP_TILE, C.BroadphaseType.SAP_SEGMENTED])
@pytest.mark.parametrize("bfilter", FILTERS)
def test_broadphase_counts(bfilter, broadphase):
  mjm = mjw.load_mjcf_string(BROADPHASE_XML)
  opt = {"broadphase_filter": int(bfilter), "broadphase": int(broadphase)}
  for key, expected in [(0, 0), (1, 1), (2, 3), (3, 1)]:
    _, d = run_collision(mjm, mjm.key_qpos[key], opt)
    assert int(d.ncollision[0]) == expected, (key, int(bfilter))
  # two worlds, four collisions in total
  _, d = run_collision(mjm, np.stack([mjm.key_qpos[1], mjm.key_qpos[2]]), opt)
  assert int(d.ncollision[0]) == 4


def test_broadphase_contype_conaffinity():
  mjm = mjw.load_mjcf_string(BROADPHASE_XML)
  mjm.geom_contype[:3] = 0
  _, d = run_collision(mjm, mjm.key_qpos[1])
  assert int(d.ncollision[0]) == 0


@pytest.mark.parametrize("m1,m2,expected", [(0, 0, 0), (0, 0.011, 1), (0.011, 0, 1), (0.00999, 0, 0), (0, 0.00999, 0), (0.00999, 0.00999, 1)])
def test_broadphase_margin(m1, m2, expected):
  xml = f"""
  <mujoco><worldbody>
    <body><geom type="sphere" size=".1" margin="{m1}"/><joint type="slide" axis="1 0 0"/></body>
    <body><geom type="sphere" size=".1" margin="{m2}"/><joint type="slide" axis="1 0 0"/></body>
  </worldbody><keyframe><key qpos="0 .21"/></keyframe></mujoco>"""
  mjm = mjw.load_mjcf_string(xml)
  _, d = run_collision(mjm, mjm.key_qpos[0])
  assert int(d.ncollision[0]) == expected


@pytest.mark.parametrize("disable,expected", [(0, 0), (int(C.DisableBit.FILTERPARENT), 1)])
def test_broadphase_filterparent(disable, expected):
  xml = """
  <mujoco><worldbody>
    <body><geom type="sphere" size=".1"/><joint type="slide"/>
      <body><geom type="sphere" size=".1"/><joint type="slide"/></body>
    </body>
  </worldbody><keyframe><key qpos="0 0"/></keyframe></mujoco>"""
  mjm = mjw.load_mjcf_string(xml)
  mjm.opt.disableflags = disable
  _, d = run_collision(mjm, mjm.key_qpos[0])
  assert int(d.ncollision[0]) == expected


# ---- contact parameters (reference collision_driver_test.py:882-972) ---------------------------
TWO_SPHERES = """
<mujoco><worldbody>
  <geom type="plane" size="10 10 .01"/>
  <body pos="0 0 0.05"><freejoint/><geom type="sphere" size=".1" {a}/></body>
  <body pos="0.1 0 0.05"><freejoint/><geom type="sphere" size=".1" {b}/></body>
</worldbody></mujoco>"""


@pytest.mark.parametrize("bit", [C.DisableBit.CONSTRAINT, C.DisableBit.CONTACT])
def test_collision_disableflags_give_no_contacts(bit):
  mjm = mjw.load_mjcf_string(TWO_SPHERES.format(a="", b=""))
  _, d = run_collision(mjm, mjm.qpos0)
  assert int(d.nacon[0]) > 0
  mjm.opt.disableflags = int(bit)
  _, d = run_collision(mjm, mjm.qpos0)
  assert int(d.nacon[0]) == 0


def test_min_friction_is_clamped():
  """collision_driver_test.py:918-942: zero friction on both geoms is clamped to MJ_MINMU."""
  xml = """
  <mujoco><worldbody>
    <body><geom type="sphere" size=".1" friction="0 0 0"/><joint type="slide"/></body>
    <body><geom type="sphere" size=".1" friction="0 0 0"/><joint type="slide"/></body>
  </worldbody><keyframe><key qpos="0 .1"/></keyframe></mujoco>"""
  mjm = mjw.load_mjcf_string(xml)
  _, d = run_collision(mjm, mjm.key_qpos[0])
  assert int(d.nacon[0]) == 1
  np.testing.assert_allclose(d.contact_friction[0], 1e-5)  # MJ_MINMU


@pytest.mark.parametrize("p1,p2", [(0, 0), (1, 0), (0, 1)])
def test_contact_parameter_mixing(p1, p2):
  a = f'priority="{p1}" friction="0.3 0.01 0.001" solmix="2" condim="1" solref="0.03 1.1"'
  b = f'priority="{p2}" friction="0.7 0.02 0.003" solmix="1" condim="3" solref="0.01 0.9"'
  mjm = mjw.load_mjcf_string(TWO_SPHERES.format(a=a, b=b))
  _, d = run_collision(mjm, mjm.qpos0)
  n = int(d.nacon[0])
  k = [i for i in range(n) if tuple(d.contact_geom[i]) == (1, 2)]
  assert len(k) == 1
  k = k[0]
  if p1 > p2:
    fr, dim, sr = [0.3, 0.3, 0.01, 0.001, 0.001], 1, [0.03, 1.1]
  elif p2 > p1:
    fr, dim, sr = [0.7, 0.7, 0.02, 0.003, 0.003], 3, [0.01, 0.9]
  else:  # max friction, max condim, solmix-weighted solref (mix = 2 / 3)
    fr, dim = [0.7, 0.7, 0.02, 0.003, 0.003], 3
    sr = [2 / 3 * 0.03 + 1 / 3 * 0.01, 2 / 3 * 1.1 + 1 / 3 * 0.9]
  np.testing.assert_allclose(d.contact_friction[k], fr, rtol=1e-12)
  assert int(d.contact_dim[k]) == dim
  np.testing.assert_allclose(d.contact_solref[k], sr, rtol=1e-12)


def test_efc_address_of_contacts_beyond_njmax_is_minus_one():
  """constraint_test.py:270-312: rows that do not fit report efc_address -1."""
  mjm = mjw.load_mjcf_string(TWO_SPHERES.format(a="", b=""))
  _, d = run_collision(mjm, mjm.qpos0, njmax=5, stages=("kinematics", "com_pos", "crb", "collision", "make_constraint"))
  n = int(d.nacon[0])
  assert n == 3
  adr = d.contact_efc_address[:n]
  assert (adr >= -1).all() and (adr < 5).all()
  assert (adr == -1).sum() > 0
  assert int(d.nefc[0]) >= 5
