"""Solver / constraint statistics of the humanoid benchmark in steady state (after N steps from the keyframe): nefc, ncon, niter histograms."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
w = bench.Workload("humanoid", 8192)
with torch.cuda.stream(w.stream):
  w.run(10)
  w.capture()
  for k in range(n):
    w.step()
    if k % 50 == 49 or k == n - 1:
      torch.cuda.synchronize()
      d = w.d
      ni = d.solver_niter.float()
      print(k + 1, "nefc mean %.1f max %d" % (d.nefc.float().mean(), d.nefc.max()), "ne/nf/nl %.1f %.1f %.1f" % (d.ne.float().mean(), d.nf.float().mean(), d.nl.float().mean()),
            "ncon/world %.2f" % (float(d.nacon[0]) / 8192), "niter mean %.2f" % ni.mean(), "hist", torch.bincount(d.solver_niter.clamp(0, 10)).tolist(),
            "quad rows %.1f" % ((d.efc.state == 1).sum().item() / 8192))
