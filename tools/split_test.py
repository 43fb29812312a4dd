"""Compares one fused launch per step with several launches per step (stage groups), both graph-captured."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib
import torch
import mujoco_warp_b200 as mjw
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
S = F
groups = {
  "fused1": [S.STAGES_FORWARD | S.STAGE_EULER],
  "split3": [S.STAGES_FORWARD & ~S.STAGE_SOLVE, S.STAGE_SOLVE, S.STAGE_EULER],
  "split6": [S.STAGE_KINEMATICS | S.STAGE_COM_POS | S.STAGE_CRB | S.STAGE_TRANSMISSION, S.STAGE_COLLISION, S.STAGE_MAKE_CONSTRAINT,
             S.STAGE_ZERO_SENSOR | S.STAGE_FWD_VELOCITY | S.STAGE_FWD_ACTUATION | S.STAGE_FWD_ACC_FACTOR, S.STAGE_SOLVE, S.STAGE_EULER],
  "split4": [S.STAGES_FWD_POSITION, S.STAGE_ZERO_SENSOR | S.STAGE_FWD_VELOCITY | S.STAGE_FWD_ACTUATION | S.STAGE_FWD_ACC_FACTOR, S.STAGE_SOLVE,
             S.STAGE_EULER],
}
state = (d.qpos.clone(), d.qvel.clone(), d.qacc_warmstart.clone(), d.act.clone() if d.act.numel() else None)
for name, gl in groups.items():
  d.qpos.copy_(state[0]); d.qvel.copy_(state[1]); d.qacc_warmstart.copy_(state[2])
  st = torch.cuda.Stream()
  with torch.cuda.stream(st):
    for _ in range(3):
      for b in gl:
        F._run(m, d, b)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
      for b in gl:
        F._run(m, d, b)
    for _ in range(5):
      g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    n = 50
    for _ in range(n):
      g.replay()
    e1.record()
    e1.synchronize()
    print(f"{name:8s} {1e3 * e0.elapsed_time(e1) / n:9.1f} us/step  niter {float(d.solver_niter.float().mean()):.2f} nefc {float(d.nefc.float().mean()):.1f}", flush=True)
