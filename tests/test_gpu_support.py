"""CUDA get_state / set_state / contact_force / reset_data (mjb_support.cu through the C ABI) vs the outputs of the REFERENCE's
own functions (tests/golden/support_golden.npz, made by tests/golden/make_support_golden.py) and vs the oracle.

The reference's Data after two steps is loaded into the device Data; the same call sequence then runs on the device.  These
are copies (bit-exact in float32: compared to the float32 rounding of the golden values) plus, for contact_force, a few
float32 sums and a 3x3 rotation (relative 1e-6)."""

import numpy as np
import pytest
import torch

import mujoco_warp_b200 as mjw
from tests import helpers as H
from tests.golden import reference_cases as RC

pytestmark = pytest.mark.gpu


def _device_case(name):
  rec = H.load_support_case(name)
  c = RC.CASES[H.SUPPORT_CASES[name]]
  mjm = RC.load_model(c["model"])
  m = mjw.put_model(mjm)
  m.opt.cone = int(rec["cone"])
  d = mjw.make_data(mjm, nworld=c["nworld"], nconmax=c.get("nconmax", 24), njmax=c.get("njmax", 64))

  def assign(f, a):
    dst = d.field(f)
    dst.copy_(torch.from_numpy(np.ascontiguousarray(a)).to(dst.dtype).reshape(dst.shape).cuda())

  H.support_fill(rec, "stepped_", H.SUPPORT_RESET_FIELDS + H.SUPPORT_CONTACT_FIELDS + ["efc_force"], assign)
  return rec, mjm, m, d


def _check_exact(d, rec, prefix, fields):
  for f in fields:
    if prefix + f not in rec:
      continue
    got = H.dev_np(d, f).reshape(-1)
    want = np.asarray(rec[prefix + f]).reshape(-1)
    want = want.astype(np.float32) if got.dtype == np.float32 else want.astype(got.dtype)
    assert np.array_equal(got, want), (prefix, f, np.abs(got.astype(np.float64) - want).max())


@pytest.mark.parametrize("name", list(H.SUPPORT_CASES))
def test_device_support_functions_match_reference_execution(name):
  rec, mjm, m, d = _device_case(name)
  sig, active = int(rec["sig"]), torch.from_numpy(rec["active"]).cuda()
  nstate = mjw.state_size(m, d, sig)
  assert nstate == rec["state_all"].shape[1]
  state = torch.zeros((d.nworld, nstate), dtype=torch.float32, device="cuda")
  mjw.get_state(m, d, state, sig)
  assert np.array_equal(state.cpu().numpy(), rec["state_all"].astype(np.float32))
  masked = torch.full((d.nworld, nstate + 3), -7.0, dtype=torch.float32, device="cuda")  # a wider row: the stride is honoured
  mjw.get_state(m, d, masked, sig, active=active)
  assert np.array_equal(masked[:, :nstate].cpu().numpy(), rec["state_masked"].astype(np.float32))
  assert float(masked[:, nstate:].max()) == -7.0
  ids = torch.from_numpy(rec["contact_ids"]).cuda()
  for key, to_world in (("force_contact", False), ("force_world", True)):
    force = torch.full((len(ids), 6), -3.0, dtype=torch.float32, device="cuda")
    mjw.contact_force(m, d, ids, to_world, force)
    got, want = force.cpu().numpy().astype(np.float64), rec[key]
    assert np.abs(got - want).max(initial=0.0) <= 2e-6 * (1.0 + np.abs(want).max(initial=0.0)), (key, np.abs(got - want).max())
  mjw.set_state(m, d, torch.from_numpy(rec["new_state"].astype(np.float32)).cuda(), sig, active=active)
  _check_exact(d, rec, "set_", H.SUPPORT_STATE_FIELDS)
  mjw.reset_data(m, d, torch.from_numpy(rec["reset"]).cuda())
  _check_exact(d, rec, "reset_", H.SUPPORT_RESET_FIELDS + H.SUPPORT_CONTACT_FIELDS)
  mjw.reset_data(m, d)
  _check_exact(d, rec, "resetall_", H.SUPPORT_RESET_FIELDS + H.SUPPORT_CONTACT_FIELDS)


def test_state_round_trip_and_mocap_layout_at_full_size():
  """Size-independent properties at the benchmark size (nworld = 8192): set_state(get_state(d)) is the identity on another
  Data, PHYSICS is the qpos | qvel | act prefix-free concatenation, and a model with mocap bodies and two equalities packs
  eq_active / mocap_pos / mocap_quat in MuJoCo's order (the reference's cursor slips are not reproduced, support.py docstring)."""
  mjm = mjw.load_mjcf(H.HUMANOID)
  nworld = 8192
  qpos, qvel, ctrl = H.make_states(mjm, nworld, seed=5)
  m, d = H.device_data(mjw, mjm, qpos, qvel, ctrl, 24, 64)
  for _ in range(2):
    mjw.step(m, d)
  sig = int(mjw.State.INTEGRATION)
  state = torch.empty((nworld, mjw.state_size(m, d, sig)), dtype=torch.float32, device="cuda")
  mjw.get_state(m, d, state, sig)
  d2 = mjw.make_data(mjm, nworld=nworld, nconmax=24, njmax=64)
  mjw.set_state(m, d2, state, sig)
  for f in H.SUPPORT_STATE_FIELDS:
    assert torch.equal(d.field(f), d2.field(f)), f
  phys = torch.empty((nworld, mjm.nq + mjm.nv), dtype=torch.float32, device="cuda")
  mjw.get_state(m, d, phys, int(mjw.State.PHYSICS))
  assert torch.equal(phys, torch.cat([d.qpos, d.qvel], dim=1))
  # both continue identically from the restored state
  mjw.step(m, d)
  mjw.step(m, d2)
  assert torch.equal(d.qpos, d2.qpos) and torch.equal(d.qvel, d2.qvel)

  xml = """<mujoco><worldbody>
    <body name="mc1" mocap="true" pos="1 2 3" quat="0 1 0 0"><geom type="sphere" size=".1" contype="0" conaffinity="0"/></body>
    <body name="mc2" mocap="true" pos="4 5 6"><geom type="sphere" size=".1" contype="0" conaffinity="0"/></body>
    <body><joint name="a" type="hinge"/><geom type="capsule" size=".05 .2"/>
      <body pos="0 0 .5"><joint name="b" type="hinge"/><geom type="capsule" size=".05 .2"/>
        <body pos="0 0 .5"><joint name="c" type="hinge"/><geom type="capsule" size=".05 .2"/></body></body></body>
  </worldbody><equality><joint joint1="a" joint2="b"/><joint joint1="b" joint2="c" active="false"/></equality></mujoco>"""
  mjm = mjw.load_mjcf_string(xml)
  m = mjw.put_model(mjm)
  d = mjw.make_data(mjm, nworld=3)
  sig = int(mjw.State.EQ_ACTIVE | mjw.State.MOCAP_POS | mjw.State.MOCAP_QUAT)
  state = torch.zeros((3, mjw.state_size(m, d, sig)), dtype=torch.float32, device="cuda")
  mjw.get_state(m, d, state, sig)
  want = [1, 0, 1, 2, 3, 4, 5, 6, 0, 1, 0, 0, 1, 0, 0, 0]
  assert state.shape[1] == 16 and state.cpu().numpy().tolist() == [want] * 3
  state[1] = torch.arange(16, dtype=torch.float32, device="cuda")
  mjw.set_state(m, d, state, sig, active=torch.tensor([False, True, False], device="cuda"))
  assert d.eq_active[1].tolist() == [0, 1] and d.eq_active[0].tolist() == [1, 0]
  assert d.mocap_pos[1].reshape(-1).tolist() == [2, 3, 4, 5, 6, 7] and d.mocap_quat[1].reshape(-1).tolist() == list(range(8, 16))
  d.mocap_pos.add_(1.0)
  mjw.reset_data(m, d, torch.tensor([True, False, True], device="cuda"))
  assert d.mocap_pos[0].reshape(-1).tolist() == [1, 2, 3, 4, 5, 6] and d.mocap_quat[2].reshape(-1).tolist() == [0, 1, 0, 0, 1, 0, 0, 0]
  assert d.mocap_pos[1].reshape(-1).tolist() == [3, 4, 5, 6, 7, 8] and d.eq_active[2].tolist() == [1, 0]


def test_reset_data_is_graph_capturable_and_masked_reset_continues_like_fresh_worlds():
  mjm = mjw.load_mjcf(H.HUMANOID)
  qpos, qvel, ctrl = H.make_states(mjm, 64, seed=6)
  m, d = H.device_data(mjw, mjm, qpos, qvel, ctrl, 24, 64)
  mask = torch.zeros(64, dtype=torch.bool, device="cuda")
  mjw.step(m, d)
  mjw.reset_data(m, d, mask)  # binds, warms up
  torch.cuda.synchronize()
  g = torch.cuda.CUDAGraph()
  s = torch.cuda.Stream()
  with torch.cuda.stream(s):
    with torch.cuda.graph(g, stream=s):
      mjw.step(m, d)
      mjw.reset_data(m, d, mask)
  mask[::2] = True  # the captured reset reads the mask at replay time
  g.replay()
  torch.cuda.synchronize()
  fresh = mjw.make_data(mjm, nworld=64, nconmax=24, njmax=64)
  assert torch.equal(d.qpos[::2], fresh.qpos[::2]) and float(d.qvel[::2].abs().max()) == 0.0 and float(d.time[::2].abs().max()) == 0.0
  assert float(d.qvel[1::2].abs().max()) > 0 and float(d.time[1].item()) > 0
  mask.zero_()
  d.ctrl.zero_()
  g.replay()
  mjw.step(m, fresh)
  torch.cuda.synchronize()
  assert torch.equal(d.qpos[::2], fresh.qpos[::2]) and torch.equal(d.qvel[::2], fresh.qvel[::2])
