"""Golden vectors of the whole step path produced by EXECUTING the reference's own code.

`ref_runner.py` imports /root/reference/mujoco_warp (unmodified source) under the numpy Warp interpreter and the
mujoco stub, and this script drives `put_model -> make_data -> forward / step` on seeded states for every case below,
storing inputs and outputs in `reference_step.npz`.  `tests/test_oracle_reference_golden.py` checks the CPU oracle
against the file (float64 both sides: agreement ~1e-9), `tests/test_gpu_reference_golden.py` checks the CUDA path
against it directly (float32 tolerances).  Run here (the reference tree does not exist on the GPU box):

  python tests/golden/make_reference_golden.py [--check] [case ...]

`--check` also runs the oracle on the same inputs and prints the worst error per case without writing the file.
"""

from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_runner as R  # noqa: E402

import reference_cases as RC  # noqa: E402
from reference_cases import CASES, COUNTS, CON_INT, CON_REAL, EFC_INT, EFC_REAL, REAL_FIELDS, case_inputs, compare, run_oracle  # noqa: E402

OUT = RC.NPZ

def run_reference(name):
  """Executes the reference on a case; returns {field: array} (per step for step cases: the state after the LAST step,
  plus the full forward fields of the last `forward` inside it)."""
  wp, ref = R.load_reference()
  types_mod = sys.modules["mujoco_warp._src.types"]
  c = CASES[name]
  mjm, qpos, qvel, ctrl = case_inputs(name)
  m, d = R.make(mjm, c["nworld"], c.get("nconmax", 24), c.get("njmax", 64))
  for k, v in c.get("opt", {}).items():
    if k == "broadphase":
      v = types_mod.BroadphaseType(v)
    setattr(m.opt, k, v)
  for flag in c.get("disable", []):
    m.opt.disableflags |= int(getattr(types_mod.DisableBit, flag))
  for flag in c.get("enable", []):
    m.opt.enableflags |= int(getattr(types_mod.EnableBit, flag))
  R.set_state(d, qpos, qvel, ctrl)
  out = {"in_qpos": qpos, "in_qvel": qvel, "in_ctrl": ctrl}
  fn = ref.forward if c["mode"] == "forward" else ref.step
  traj_qpos, traj_qvel = [], []
  ctrl_traj = RC.case_ctrl_traj(name, mjm)
  for t in range(c.get("nstep", 1)):
    if ctrl_traj is not None:
      d.ctrl.numpy()[...] = ctrl_traj[t]
    fn(m, d)
    traj_qpos.append(np.array(d.qpos.numpy()))
    traj_qvel.append(np.array(d.qvel.numpy()))
  out["traj_qpos"], out["traj_qvel"] = np.stack(traj_qpos), np.stack(traj_qvel)
  for f in REAL_FIELDS + COUNTS:
    out[f] = np.array(getattr(d, f).numpy())
  for f in EFC_REAL + EFC_INT:
    out["efc_" + f] = np.array(getattr(d.efc, f).numpy())
  for f in CON_REAL + CON_INT:
    out["contact_" + f] = np.array(getattr(d.contact, f).numpy())
  if m.is_sparse:
    _densify(m, d, mjm, out)
  return out


def _densify(m, d, mjm, out):
  """A model the reference stores sparse: qM (legacy format: dof_Madr + the walk up dof_parentid, smooth.py crb) and efc.J (rows as
  rownnz / rowadr / colind) are stored DENSE in the fixture, which is the layout this library keeps for every model; qLD (sparse
  L^T D L factor) has no dense-Cholesky counterpart and is left out."""
  nv, nworld, njmax = int(mjm.nv), out["qM"].shape[0], int(d.njmax)
  madr, parent = np.asarray(mjm.dof_Madr), np.asarray(mjm.dof_parentid)
  qM = np.zeros((nworld, nv, nv))
  src = out["qM"].reshape(nworld, -1)
  for i in range(nv):
    adr, j = int(madr[i]), i
    while j >= 0:
      qM[:, i, j] = qM[:, j, i] = src[:, adr]
      adr, j = adr + 1, int(parent[j])
  out["qM"] = qM
  J = np.zeros((nworld, int(d.njmax_pad), int(m.nv_pad)))  # the dense path's padded shape (io.py:650)
  val = out["efc_J"].reshape(nworld, -1)
  rownnz, rowadr = d.efc.J_rownnz.numpy(), d.efc.J_rowadr.numpy()
  colind = d.efc.J_colind.numpy().reshape(nworld, -1)
  for w in range(nworld):
    for r in range(int(out["nefc"][w])):
      a, n = int(rowadr[w, r]), int(rownnz[w, r])
      J[w, r, colind[w, a : a + n]] = val[w, a : a + n]
  out["efc_J"] = J
  out["efc_J_rownnz"] = np.array(rownnz)
  del out["qLD"]
  # solver.py:2189 zeroes qfrc_constraint of ALL worlds on every solver iteration of the sparse path, while the kernel that refills it
  # (:1955) skips worlds that already converged: a world that finishes before the slowest one ends up with qfrc_constraint = 0.  An
  # artefact of which worlds share a batch, not a property of the world: left out (J^T efc_force is checked through efc_J and efc_force).
  del out["qfrc_constraint"]


def main(argv):
  check = "--check" in argv
  names = [a for a in argv if not a.startswith("--")] or list(CASES)
  store = {}
  if os.path.exists(OUT) and names != list(CASES):
    with np.load(OUT) as z:
      store = {k: z[k] for k in z.files}
  for name in names:
    t = time.time()
    out = run_reference(name)
    dt = time.time() - t
    od = run_oracle(name)
    worst = compare(name, out, od)
    bad = {k: v for k, v in worst.items() if not v < 1e-7}
    print(f"{name:40s} ref {dt:6.1f}s  nefc {out['nefc'].tolist()} nacon {out['nacon'].tolist()}  worst {max(worst.values()):.2e}  {bad if bad else ''}")
    for k, v in out.items():
      store[f"{name}/{k}"] = v
  if not check:
    np.savez_compressed(OUT, **store)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
  main(sys.argv[1:])
