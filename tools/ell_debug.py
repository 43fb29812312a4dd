import numpy as np, torch, sys
sys.path.insert(0, ".")
import mujoco_warp_b200 as mjw
from mujoco_warp_b200 import constants as C
from tests import helpers as H
from oracle import oracle as O
mjm = mjw.load_mjcf(H.HUMANOID)
qpos, qvel, ctrl = H.make_states(mjm, 64, seed=31)
qpos[:, 2] -= 0.1
for solver in (2, 1):
  opt = {"cone": 1, "solver": solver, "iterations": 100}
  om, od = H.oracle_data(mjm, qpos, qvel, ctrl, 24, 64)
  O.run(om, od, "forward", opt=opt, nthread=4)
  m, d = H.device_data(mjw, mjm, qpos, qvel, ctrl, 24, 64)
  for k, v in opt.items(): setattr(m.opt, k, v)
  mjw.forward(m, d); torch.cuda.synchronize()
  qa = H.dev_np(d, "qacc"); err = np.abs(qa - od.qacc).max(axis=1) / (1 + np.abs(od.qacc).max(axis=1))
  nit_d = H.dev_np(d, "solver_niter").reshape(-1); nit_o = od.solver_niter.reshape(-1)
  print("solver", solver, "max err", err.max(), "worlds>1e-2:", np.nonzero(err > 1e-2)[0])
  print(" niter dev", nit_d[:32]); print(" niter ora", nit_o[:32])
  # cost of the device solution evaluated in float64: KKT residual
  nv = mjm.nv
  for w in np.argsort(-err)[:4]:
    n = int(od.nefc[w]); J = od.efc_J[w].reshape(64, -1)[:n, :nv]
    f_d = H.dev_np(d, "efc_force")[w][:n]; res_d = H.dev_np(d, "efc_Ma")[w] - od.qfrc_smooth[w] - J.T @ f_d
    res_o = od.efc_Ma[w] - od.qfrc_smooth[w] - J.T @ od.efc_force[w][:n]
    sd = H.dev_np(d, "efc_state")[w][:n]; so = od.efc_state[w][:n]
    print(" world", w, "err", err[w], "kkt dev", np.abs(res_d).max(), "kkt ora", np.abs(res_o).max(), "niter", nit_d[w], nit_o[w], "state diff", (sd != so).sum(), "ncone", (so == 4).sum())
