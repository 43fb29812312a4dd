"""Steps a benchmark config eagerly (no CUDA graph) for profiling under ncu:  python tools/aloha_run.py [config] [nworld] [nstep]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "aloha_pot"
nworld = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
nstep = int(sys.argv[3]) if len(sys.argv) > 3 else 320
w = bench.Workload(name, nworld, use_graph=False)
with torch.cuda.stream(w.stream):
  w.run(nstep)
torch.cuda.synchronize()
print("done", float(w.d.nacon[0]) / nworld)
