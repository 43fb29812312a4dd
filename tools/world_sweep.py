"""Steady-state step time of a bench config at several world counts:  python tools/world_sweep.py humanoid 1024 2048 4096"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

name = sys.argv[1]
for nworld in [int(a) for a in sys.argv[2:]]:
  w = bench.Workload(name, nworld, use_graph=True)
  with torch.cuda.stream(w.stream):
    w.run(bench.SETTLE_STEPS)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(w.stream)
    w.run(200)
    e1.record(w.stream)
  e1.synchronize()
  ms = e0.elapsed_time(e1) / 200
  print(f"{name} nworld {nworld:6d}  {ms * 1e3:8.1f} us/step  {nworld / ms / 1e3:8.3f} M env-steps/s", flush=True)
