"""Variants of the end-to-end loop (host ctrl -> H2D -> step -> D2H obs) timed against each other on one GPU."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mujoco_warp_b200 as mjw
from tests import helpers as H

nworld, Ke = 8192, 200
mjm = mjw.load_mjcf(H.HUMANOID)
m = mjw.put_model(mjm)
hd = mjw.make_host_data(mjm); hd.qpos[:] = mjm.key_qpos[0]
d = mjw.put_data(mjm, hd, nworld=nworld, nconmax=24, njmax=64, m=m)
nu, nq, nv = m.nu, m.nq, m.nv
stream = torch.cuda.Stream(); copy_stream = torch.cuda.Stream()
with torch.cuda.stream(stream):
  for i in range(30):
    mjw.ctrl_noise(m, d, i); mjw.step(m, d)
stream.synchronize()


def run(depth, fused_graph):
  NB = depth
  h_ctrl = [torch.empty((nworld, nu), dtype=torch.float32).pin_memory().uniform_(-0.1, 0.1) for _ in range(NB)]
  h_obs = [torch.empty((nworld, nq + nv), dtype=torch.float32).pin_memory() for _ in range(NB)]
  d_stage = [torch.empty_like(d.ctrl) for _ in range(NB)]
  d_obs = [torch.empty((nworld, nq + nv), dtype=torch.float32, device="cuda") for _ in range(NB)]
  ev_ctrl = [torch.cuda.Event() for _ in range(NB)]; ev_step = [torch.cuda.Event() for _ in range(NB)]; ev_obs = [torch.cuda.Event() for _ in range(NB)]
  graphs = []
  with torch.cuda.stream(stream):
    if fused_graph:
      for b in range(NB):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
          d.ctrl.copy_(d_stage[b]); mjw.step(m, d); d_obs[b][:, :nq].copy_(d.qpos); d_obs[b][:, nq:].copy_(d.qvel)
        graphs.append(g)
    else:
      g = torch.cuda.CUDAGraph()
      with torch.cuda.graph(g, stream=stream):
        mjw.step(m, d)
      graphs = [g] * NB
  torch.cuda.synchronize()

  def upload(k):
    b = k % NB
    with torch.cuda.stream(copy_stream):
      d_stage[b].copy_(h_ctrl[b], non_blocking=True); ev_ctrl[b].record(copy_stream)

  def step(k):
    b = k % NB
    upload(k + 1)
    with torch.cuda.stream(stream):
      stream.wait_event(ev_ctrl[b])
      if fused_graph:
        graphs[b].replay()
      else:
        d.ctrl.copy_(d_stage[b], non_blocking=True); graphs[b].replay()
        d_obs[b][:, :nq].copy_(d.qpos, non_blocking=True); d_obs[b][:, nq:].copy_(d.qvel, non_blocking=True)
      ev_step[b].record(stream)
    with torch.cuda.stream(copy_stream):
      copy_stream.wait_event(ev_step[b]); h_obs[b].copy_(d_obs[b], non_blocking=True); ev_obs[b].record(copy_stream)
    lag = depth - 1
    if k >= lag:
      ev_obs[(k - lag) % NB].synchronize()

  def loop(n):
    upload(0)
    for k in range(n):
      step(k)
    torch.cuda.synchronize()

  loop(6)
  torch.cuda.synchronize(); t0 = time.perf_counter(); loop(Ke); dt = time.perf_counter() - t0
  return nworld * Ke / dt / 1e6, dt / Ke * 1e3


for depth in (2, 3):
  for fused in (False, True):
    v, ms = run(depth, fused)
    print(f"depth {depth} fused_graph {fused}: {v:.2f} M env-steps/s  {ms:.3f} ms/step", flush=True)
# device-only reference
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(stream):
  with torch.cuda.graph(g, stream=stream):
    mjw.step(m, d)
  torch.cuda.synchronize(); t0 = time.perf_counter()
  for _ in range(Ke): g.replay()
  torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"device-only graph replay: {nworld * Ke / dt / 1e6:.2f} M  {dt / Ke * 1e3:.3f} ms/step")
