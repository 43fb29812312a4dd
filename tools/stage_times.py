"""Times each pipeline stage separately (CUDA events), humanoid nworld=8192 after a few steps."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mujoco_warp_b200 as mjw
import importlib
F = importlib.import_module("mujoco_warp_b200.forward")
from tests import helpers as H

nworld = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
mjm = mjw.load_mjcf(H.HUMANOID)
m = mjw.put_model(mjm)
hd = mjw.make_host_data(mjm)
hd.qpos[:] = mjm.key_qpos[0]
d = mjw.put_data(mjm, hd, nworld=nworld, nconmax=24, njmax=64, m=m)
for i in range(30):
  mjw.ctrl_noise(m, d, i)
  mjw.step(m, d)
torch.cuda.synchronize()
print("nefc mean", float(d.nefc.float().mean()), "niter mean", float(d.solver_niter.float().mean()), "nacon", int(d.nacon[0]))
stages = [
  ("kinematics", F.STAGE_KINEMATICS), ("com_pos", F.STAGE_COM_POS), ("crb", F.STAGE_CRB), ("collision", F.STAGE_COLLISION),
  ("make_constraint", F.STAGE_MAKE_CONSTRAINT), ("transmission", F.STAGE_TRANSMISSION), ("fwd_velocity", F.STAGE_FWD_VELOCITY),
  ("fwd_actuation", F.STAGE_FWD_ACTUATION), ("fwd_acc_factor", F.STAGE_FWD_ACC_FACTOR), ("solve", F.STAGE_SOLVE),
  ("forward", F.STAGES_FORWARD), ("fwd_position", F.STAGES_FWD_POSITION),
]
ws = d.qacc_warmstart.clone()
for name, bits in stages:
  for _ in range(3):
    F._run(m, d, bits)
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  n = 10
  for _ in range(n):
    F._run(m, d, bits)
  e1.record()
  e1.synchronize()
  print(f"{name:16s} {1e3 * e0.elapsed_time(e1) / n:9.1f} us")
