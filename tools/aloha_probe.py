"""aloha_pot collision statistics in the replay's steady state: pairs passing the broadphase per world, contacts per world, pair types."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench

nworld = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
w = bench.Workload("aloha_pot", nworld)
with torch.cuda.stream(w.stream):
  w.run(5); w.capture()
  for k in range(300):
    w.step()
    if k % 100 == 99:
      torch.cuda.synchronize()
      d = w.d
      print(k + 1, "ncollision/world %.2f" % (float(d.ncollision[0]) / nworld), "nacon/world %.2f" % (float(d.nacon[0]) / nworld), "nefc %.1f" % d.nefc.float().mean())
  g, _ = w.group_times(20)
  print({k: round(v * 1e3, 1) for k, v in g.items()})
# which geom types do the colliding pairs have (world 0)?
d = w.d
n = int(d.nacon[0]); geom = d.contact.geom[:n].cpu().numpy(); wid = d.contact.worldid[:n].cpu().numpy()
gt = np.asarray(w.mjm.geom_type)
from collections import Counter
print("contact pair types", Counter((int(gt[a]), int(gt[b])) for a, b in geom[wid == 0]))
