"""Generates tests/golden/*.npz from the CPU oracle (float64) on seeded inputs.

The reference itself cannot run in this container (it needs warp-lang and mujoco), so
these fixtures are outputs of the oracle restatement, NOT of the reference: they pin the oracle and
the CUDA path against regressions and travel to the GPU box; parity with MuJoCo C stays "unpinned".

  python tests/golden/make_golden.py
"""

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import mujoco_warp_b200 as mjw  # noqa: E402
from oracle import oracle as O  # noqa: E402
from tests import helpers as H  # noqa: E402

NWORLD, NCONMAX, NJMAX, NSTEP, SEED = 8, 24, 64, 10, 2024
FORWARD_FIELDS = H.SMOOTH_FIELDS + H.EFC_FIELDS + ["efc_force", "qfrc_constraint", "qacc", "efc_Ma"]
INT_FIELDS = ["ne", "nf", "nl", "nefc", "nacon", "ncollision", "wcon_n", "efc_type", "efc_id"]


def main():
  mjm = mjw.load_mjcf(H.HUMANOID)
  qpos, qvel, ctrl = H.make_states(mjm, NWORLD, seed=SEED)
  om, d = H.oracle_data(mjm, qpos, qvel, ctrl, NCONMAX, NJMAX)
  O.run(om, d, "forward")
  out = {"qpos0": qpos, "qvel0": qvel, "ctrl": ctrl, "seed": SEED, "nconmax": NCONMAX, "njmax": NJMAX}
  for f in FORWARD_FIELDS + INT_FIELDS:
    out["fwd_" + f] = np.array(getattr(d, f))
  # contacts in per-world order
  for f in H.CONTACT_FIELDS + H.CONTACT_INT_FIELDS:
    out["fwd_" + f] = np.concatenate([H.world_contacts_orc(d, w, f) for w in range(NWORLD)], axis=0)
  # trajectory
  om, d = H.oracle_data(mjm, qpos, qvel, ctrl, NCONMAX, NJMAX)
  traj_q, traj_v, traj_nefc = [], [], []
  for _ in range(NSTEP):
    O.run(om, d, "step")
    traj_q.append(d.qpos.copy())
    traj_v.append(d.qvel.copy())
    traj_nefc.append(d.nefc.copy())
  out["traj_qpos"], out["traj_qvel"], out["traj_nefc"] = np.array(traj_q), np.array(traj_v), np.array(traj_nefc)
  path = os.path.join(ROOT, "tests", "golden", "humanoid_forward_step.npz")
  np.savez_compressed(path, **out)
  print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
  main()
