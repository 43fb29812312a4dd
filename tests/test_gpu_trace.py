"""BASELINE.json config 1: humanoid, nworld=1, Euler + Newton, 1000 steps -- the parity trace.

The float64 oracle integrates ONE world for 1000 steps under the benchmark's OU/Halton control noise
(reference benchmark.py:42-81) and records the state entering every step.  The device then takes all 1000
recorded states as a batch of 1000 worlds and steps each ONCE (teacher forcing), so every step of the
trace is compared without the chaotic divergence a free-running float32 trajectory accumulates:

  * nefc / ne / nf / nl / per-world contact count: bit-exact on every step,
  * qpos 2e-6, qvel 2e-3, qacc 5e-2 (relative-absolute measure of helpers.max_err; qacc is the solver
    output whose float32 noise floor is the solver tolerance), efc_force 2e-2 of the largest force.

A second test lets the device run free for the first 100 steps of the same trace and checks it stays
within 2e-3 / 5e-2 of the oracle (before the divergence of the falling humanoid dominates).
"""

import numpy as np
import pytest
import torch

import mujoco_warp_b200 as mjw
from tests import helpers as H

pytestmark = pytest.mark.gpu

NSTEP, NCONMAX, NJMAX = 1000, 24, 64


def _halton(index, base):
  n0, b, hn, f = index, float(base), 0.0, 1.0 / base
  while n0 > 0:
    n1 = n0 // base
    hn += f * (n0 - n1 * base)
    f /= b
    n0 = n1
  return hn


def _ctrl_noise(ctrl, step, nu, std=0.01, rate=0.1, dt=0.005):
  """benchmark.py:59-81 for world 0 with centre 0 (actuators of the humanoid have no ctrl range midpoint)."""
  r = np.exp(-dt / max(rate, 1e-15))
  scale = std * np.sqrt(1 - r * r)
  out = ctrl * r
  for a in range(nu):
    out[a] += scale * (2 * _halton((step + 1) * 1, a + 2) - 1)
  return out


@pytest.fixture(scope="module")
def trace(humanoid):
  from oracle import oracle as O

  om = O.OracleModel(humanoid)
  od = O.make_data(om, nworld=1, nconmax=NCONMAX, njmax=NJMAX)
  od.qpos[0] = np.asarray(humanoid.key_qpos[0])
  od.qvel[:] = 0
  O.run(om, od, "forward")
  od.qacc_warmstart[:] = od.qacc
  rec = {k: [] for k in "qpos qvel ctrl warm qpos1 qvel1 qacc nefc ne nf nl ncon force".split()}
  ctrl = np.zeros(humanoid.nu)
  dt = float(humanoid.opt.timestep)
  for s in range(NSTEP):
    ctrl = _ctrl_noise(ctrl, s, humanoid.nu, dt=dt)
    od.ctrl[0] = ctrl
    rec["qpos"].append(od.qpos[0].copy())
    rec["qvel"].append(od.qvel[0].copy())
    rec["ctrl"].append(ctrl.copy())
    rec["warm"].append(od.qacc_warmstart[0].copy())
    O.run(om, od, "step")
    rec["qpos1"].append(od.qpos[0].copy())
    rec["qvel1"].append(od.qvel[0].copy())
    rec["qacc"].append(od.qacc[0].copy())
    for k, f in (("nefc", "nefc"), ("ne", "ne"), ("nf", "nf"), ("nl", "nl"), ("ncon", "wcon_n")):
      rec[k].append(int(getattr(od, f)[0]))
    rec["force"].append(od.efc_force[0].copy())
  return {k: np.asarray(v) for k, v in rec.items()}


def test_every_step_of_the_trace_teacher_forced(humanoid, trace):
  t = trace
  m, d = H.device_data(mjw, humanoid, t["qpos"], t["qvel"], t["ctrl"], NCONMAX, NJMAX, warmstart=t["warm"])
  mjw.step(m, d)
  torch.cuda.synchronize()
  assert int(d.status.max()) == 0
  for k, f in (("nefc", "nefc"), ("ne", "ne"), ("nf", "nf"), ("nl", "nl"), ("ncon", "wcon_n")):
    np.testing.assert_array_equal(H.dev_np(d, f), t[k], err_msg=f)
  assert t["nefc"].max() > 8, "trace never made contact: not a useful parity case"
  assert H.max_err(H.dev_np(d, "qpos"), t["qpos1"]) < 2e-6
  assert H.max_err(H.dev_np(d, "qvel"), t["qvel1"]) < 2e-3
  assert H.max_err(H.dev_np(d, "qacc"), t["qacc"]) < 5e-2
  f_dev, f_ref = H.dev_np(d, "efc_force"), t["force"]
  for w in range(NSTEP):
    n = int(t["nefc"][w])
    if n:
      scale = max(1.0, float(np.abs(f_ref[w, :n]).max()))
      assert float(np.abs(f_dev[w, :n] - f_ref[w, :n]).max()) / scale < 2e-2, w


def test_free_running_device_tracks_the_first_100_steps(humanoid, trace):
  t = trace
  m, d = H.device_data(mjw, humanoid, t["qpos"][:1], t["qvel"][:1], t["ctrl"][:1], NCONMAX, NJMAX, warmstart=t["warm"][:1])
  for s in range(100):
    d.ctrl.copy_(torch.from_numpy(t["ctrl"][s : s + 1].astype(np.float32)).cuda())
    mjw.step(m, d)
  assert H.max_err(H.dev_np(d, "qpos"), t["qpos1"][99:100]) < 2e-3
  assert H.max_err(H.dev_np(d, "qvel"), t["qvel1"][99:100]) < 5e-2
  assert abs(float(d.time[0]) - 100 * float(humanoid.opt.timestep)) < 1e-5
