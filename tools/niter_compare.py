"""Solver iterations of the float32 device vs the float64 oracle from the SAME steady-state benchmark states."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mujoco_warp_b200 as mjw
from tests import helpers as H
from oracle import oracle as O

nworld = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
mjm = mjw.load_mjcf(H.HUMANOID)
m = mjw.put_model(mjm)
hd = mjw.make_host_data(mjm); hd.qpos[:] = mjm.key_qpos[0]
d = mjw.put_data(mjm, hd, nworld=nworld, nconmax=24, njmax=64, m=m)
for i in range(260):
  mjw.ctrl_noise(m, d, i); mjw.step(m, d)
torch.cuda.synchronize()
mjw.ctrl_noise(m, d, 260)
qpos, qvel, ctrl, warm = (H.dev_np(d, n).astype(np.float64) for n in ("qpos", "qvel", "ctrl", "qacc_warmstart"))
om, od = H.oracle_data(mjm, qpos, qvel, ctrl, 24, 64, warmstart=warm)
O.run(om, od, "forward", nthread=16)
mjw.forward(m, d); torch.cuda.synchronize()
nd, no = H.dev_np(d, "solver_niter"), od.solver_niter
print("nefc mean", od.nefc.mean(), "device niter mean %.3f  oracle niter mean %.3f" % (nd.mean(), no.mean()))
print("hist device", np.bincount(nd, minlength=8)[:10], "oracle", np.bincount(no, minlength=8)[:10])
print("device - oracle hist", {int(k): int(v) for k, v in zip(*np.unique(nd.astype(int) - no.astype(int), return_counts=True))})
