"""Where the end-to-end loop of bench.py loses time against the device-resident replay: variants of the loop on one GPU.
Finding (round 2): none of the transfer variants matters (a 0.69 MB upload is 30 us and overlaps); the 25 % "gap" of the round-1
loop was its WORKLOAD -- it uploaded uncorrelated +-0.1 controls, which agitate the humanoids and cost solver iterations, while the
device-resident loop applies the benchmark's small Ornstein-Uhlenbeck noise.  This probe still uploads the agitating controls, so
every variant that uploads shows the slower step; bench.py now uploads frames of the OU process.
  python tools/e2e_probe.py [nsteps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import mujoco_warp_b200 as mjw

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
w = bench.Workload("humanoid", 8192)
m, d, stream = w.m, w.d, w.stream
nu, nq, nv, nworld = m.nu, m.nq, m.nv, 8192
with torch.cuda.stream(stream):
  w.run(10); w.capture(); w.run(250); stream.synchronize()
  h_ctrl = [torch.empty((nworld, nu)).uniform_(-0.1, 0.1).pin_memory() for _ in range(2)]
  h_obs = [torch.empty((nworld, nq + nv)).pin_memory() for _ in range(2)]
  stage = [torch.empty_like(d.ctrl) for _ in range(2)]
  d_obs = [torch.empty((nworld, nq + nv), device="cuda") for _ in range(2)]
  sig = int(mjw.State.QPOS) | int(mjw.State.QVEL)
  graphs = []
  for b in range(2):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
      mjw.set_state(m, d, stage[b], int(mjw.State.CTRL)); mjw.step(m, d); mjw.get_state(m, d, d_obs[b], sig)
    graphs.append(g)
  copy = torch.cuda.Stream()
  evc = [torch.cuda.Event() for _ in range(2)]; evs = [torch.cuda.Event() for _ in range(2)]; evo = [torch.cuda.Event() for _ in range(2)]

  def loop(h2d, d2h, hostsync, same_stream=False):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for k in range(n):
      b = k & 1
      if h2d:
        if same_stream:
          stage[b].copy_(h_ctrl[b], non_blocking=True)
        else:
          with torch.cuda.stream(copy):
            stage[b].copy_(h_ctrl[b], non_blocking=True); evc[b].record(copy)
          stream.wait_event(evc[b])
      graphs[b].replay()
      if d2h:
        if same_stream:
          h_obs[b].copy_(d_obs[b], non_blocking=True); evo[b].record(stream)
        else:
          evs[b].record(stream)
          with torch.cuda.stream(copy):
            copy.wait_event(evs[b]); h_obs[b].copy_(d_obs[b], non_blocking=True); evo[b].record(copy)
        if hostsync and k > 0:
          evo[(k - 1) & 1].synchronize()
    torch.cuda.synchronize()
    return nworld * n / (time.perf_counter() - t0) / 1e6

  # zero-copy: the graph's set_state / get_state kernels address the pinned host buffers directly
  zg = []
  for b in range(2):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
      mjw.set_state(m, d, h_ctrl[b], int(mjw.State.CTRL)); mjw.step(m, d); mjw.get_state(m, d, h_obs[b], sig)
    zg.append(g)

  def zloop(hostsync):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for k in range(n):
      b = k & 1
      zg[b].replay(); evo[b].record(stream)
      if hostsync and k > 0:
        evo[(k - 1) & 1].synchronize()
    torch.cuda.synchronize()
    return nworld * n / (time.perf_counter() - t0) / 1e6

  for hs in (0, 1):
    zloop(hs); print("zero-copy hostsync=%d   %.2f M env-steps/s" % (hs, zloop(hs)))
  h_obs[0].zero_(); zg[0].replay(); torch.cuda.synchronize()
  print("obs check", float((h_obs[0][:, :nq] - d.qpos.cpu()).abs().max()), float(h_obs[0].abs().max()))
  for name, args in [("graph only", (0, 0, 0)), ("h2d", (1, 0, 0)), ("d2h nosync", (0, 1, 0)), ("h2d+d2h nosync", (1, 1, 0)), ("full (bench)", (1, 1, 1)),
                     ("full same-stream", (1, 1, 1, True))]:
    loop(*args); print("%-20s %.2f M env-steps/s" % (name, loop(*args)))
