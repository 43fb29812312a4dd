"""Prints per-field max errors between the CUDA library and the CPU oracle (debug aid; GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mujoco_warp_b200 as mjw
from tests import helpers as H
from oracle import oracle as O

nworld = int(sys.argv[1]) if len(sys.argv) > 1 else 64
mjm = mjw.load_mjcf(H.HUMANOID)
qpos, qvel, ctrl = H.make_states(mjm, nworld, seed=1)
om, od = H.oracle_data(mjm, qpos, qvel, ctrl, 24, 64)
m, d = H.device_data(mjw, mjm, qpos, qvel, ctrl, 24, 64)
O.run(om, od, "forward")
mjw.forward(m, d)
torch.cuda.synchronize()
print("status", d.status.unique().tolist(), "nacon", int(d.nacon[0]), od.nacon[0], "ncollision", int(d.ncollision[0]), od.ncollision[0])
for name in ("nefc", "ne", "nf", "nl", "wcon_n", "solver_niter"):
  a, b = H.dev_np(d, name), getattr(od, name)
  print(f"{name:14s} equal={np.array_equal(a, b)} dev[:8]={a[:8]} orc[:8]={b[:8]}")
for name in H.SMOOTH_FIELDS + H.EFC_FIELDS + ["efc_force", "qfrc_constraint", "efc_Ma", "qacc"]:
  a, b = H.dev_np(d, name), getattr(od, name)
  if name.startswith("efc_") and name not in ("efc_Ma",):
    errs = []
    for w in range(nworld):
      n = min(int(od.nefc[w]), 64)
      errs.append(H.max_err(a[w, :n], b[w, :n]))
    e = max(errs)
  else:
    e = H.max_err(a, b)
  print(f"{name:18s} {e:.3e}")
bad = 0
for w in range(nworld):
  if int(d.wcon_n[w]) != int(od.wcon_n[w]):
    bad += 1
    continue
  for name in H.CONTACT_FIELDS:
    e = H.max_err(H.world_contacts_dev(d, w, name), H.world_contacts_orc(od, w, name))
    if e > 1e-4:
      print("contact", w, name, e)
  for name in H.CONTACT_INT_FIELDS:
    if not np.array_equal(H.world_contacts_dev(d, w, name), H.world_contacts_orc(od, w, name)):
      print("contact int mismatch", w, name)
print("worlds with contact count mismatch:", bad)
# multi-step
for i in range(20):
  O.run(om, od, "step")
  mjw.step(m, d)
torch.cuda.synchronize()
for name in ("qpos", "qvel", "qacc"):
  print(f"after 20 steps {name:6s} {H.max_err(H.dev_np(d, name), getattr(od, name)):.3e}")
print("niter dev", H.dev_np(d, "solver_niter")[:16], "orc", od.solver_niter[:16])
print("nan", bool(torch.isnan(d.qpos).any()))
