"""Times forward / prefix-fused stage sets for several nworld (CUDA events) to separate compute from memory effects."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib
import torch
import mujoco_warp_b200 as mjw
F = importlib.import_module("mujoco_warp_b200.forward")
from tests import helpers as H

mjm = mjw.load_mjcf(H.HUMANOID)
m = mjw.put_model(mjm)


def timeit(d, bits, n=10):
  for _ in range(3):
    F._run(m, d, bits)
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  torch.cuda.synchronize()
  e0.record()
  for _ in range(n):
    F._run(m, d, bits)
  e1.record()
  e1.synchronize()
  return 1e3 * e0.elapsed_time(e1) / n


order = [F.STAGE_KINEMATICS, F.STAGE_COM_POS, F.STAGE_CRB, F.STAGE_COLLISION, F.STAGE_MAKE_CONSTRAINT, F.STAGE_TRANSMISSION,
         F.STAGE_ZERO_SENSOR | F.STAGE_FWD_VELOCITY, F.STAGE_FWD_ACTUATION, F.STAGE_FWD_ACC_FACTOR, F.STAGE_SOLVE]
names = ["kin", "com", "crb", "coll", "mkc", "trn", "vel", "act", "acc", "solve"]
for nworld in [int(a) for a in sys.argv[1:]] or [1024, 2048, 4096, 8192, 16384]:
  hd = mjw.make_host_data(mjm)
  hd.qpos[:] = mjm.key_qpos[0]
  d = mjw.put_data(mjm, hd, nworld=nworld, nconmax=24, njmax=64, m=m)
  for i in range(30):
    mjw.ctrl_noise(m, d, i)
    mjw.step(m, d)
  torch.cuda.synchronize()
  line = f"nworld {nworld:6d} nefc {float(d.nefc.float().mean()):.1f} niter {float(d.solver_niter.float().mean()):.2f} | prefix us:"
  bits = 0
  for nm, b in zip(names, order):
    bits |= b
    line += f" {nm} {timeit(d, bits):.0f}"
  line += f" | solve-only {timeit(d, F.STAGE_SOLVE):.0f} us/world-fwd {timeit(d, F.STAGES_FORWARD) / nworld * 1e3:.1f} ns"
  print(line, flush=True)
