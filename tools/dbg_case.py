"""Per-world counts of a reference golden case on the device vs the golden: python tools/dbg_case.py <case>"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mujoco_warp_b200 as mjw
from tests import helpers as H
from tests.golden import reference_cases as RC

name = sys.argv[1]
c = RC.CASES[name]; g = RC.load_golden(name); mjm = RC.load_case_model(name)
m, d = H.device_data(mjw, mjm, g["in_qpos"], g["in_qvel"], g["in_ctrl"], c.get("nconmax", 24), c.get("njmax", 64))
for k, v in RC.case_opt(name).items():
  setattr(m.opt, k, v)
(mjw.forward if c["mode"] == "forward" else mjw.step)(m, d)
torch.cuda.synchronize()
print("nefc dev", H.dev_np(d, "nefc").tolist()); print("nefc ref", g["nefc"].tolist())
n = int(d.nacon[0]); nr = int(g["nacon"][0])
wd, wr = H.dev_np(d, "contact_worldid")[:n], g["contact_worldid"][:nr]
gd, gr = H.dev_np(d, "contact_geom")[:n], g["contact_geom"][:nr]
np.set_printoptions(precision=5, suppress=True, linewidth=180)
for w in range(c["nworld"]):
  a = sorted(map(tuple, gd[wd == w].tolist())); b = sorted(map(tuple, gr[wr == w].tolist()))
  if a != b:
    print("world", w, "dev", a, "ref", b)
    for tag, pos, dist, frame, sel in (("dev", H.dev_np(d, "contact_pos")[:n], H.dev_np(d, "contact_dist")[:n], H.dev_np(d, "contact_frame")[:n], wd == w),
                                       ("ref", g["contact_pos"][:nr], g["contact_dist"][:nr], g["contact_frame"][:nr], wr == w)):
      for p, ds, fr in zip(pos[sel], dist[sel], frame[sel].reshape(-1, 9)):
        print("   ", tag, p, "%.5f" % ds, fr[:3])
