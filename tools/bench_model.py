"""Device-resident env-steps/s of `mjw.step` on a secondary benchmark config (not the bench.py headline line).

  python tools/bench_model.py panda [nworld] [nconmax] [njmax] [nstep]

panda: benchmarks/franka_emika_panda/scene.xml (reference benchmarks/config.txt:22: nworld 32768, nconmax 1, njmax 5;
BASELINE.json config 3 quotes nworld 8192).  Initial state qpos0, controls = OU noise around 0 like the benchmark loop
(reference benchmark.py:42), CUDA-graph step, CUDA-event timing after warm-up.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import mujoco_warp_b200 as mjw

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODELS = {"panda": ("benchmarks/franka_emika_panda/scene.xml", 32768, 1, 5), "humanoid": ("benchmarks/humanoid/humanoid.xml", 8192, 24, 64),
          "apollo": ("benchmarks/apptronik_apollo/scene_flat.xml", 8192, 16, 64),  # reference benchmarks/config.txt:17
          "aloha": ("benchmarks/aloha_pot/scene.xml", 8192, 24, 128)}  # config.txt:15, replay = keyframes lift_pot*
REPLAY = {"aloha": "lift_pot"}


make_trajectory = mjw.make_trajectory  # reference io.py:2603


def main():
  name = sys.argv[1] if len(sys.argv) > 1 else "panda"
  path, nworld, nconmax, njmax = MODELS[name]
  args = [int(a) for a in sys.argv[2:]]
  nworld, nconmax, njmax = (args + [nworld, nconmax, njmax][len(args):])[:3] if len(args) < 3 else args[:3]
  nstep = args[3] if len(args) > 3 else 200
  mjm = mjw.load_mjcf(os.path.join(ROOT, path))
  if name == "apollo":  # box-box margins: the reference accepts the scene only with NATIVECCD disabled (io.py:386-395)
    mjm.opt.disableflags |= int(mjw.DisableBit.NATIVECCD)
  m = mjw.put_model(mjm)
  traj = None
  if name in REPLAY:  # testspeed.py:277-282: state of the first matching keyframe, controls replayed along the keyframe sequence
    keys = [i for i, n in enumerate(mjm.key_names) if n.startswith(REPLAY[name])]
    hd = mjw.make_host_data(mjm)
    hd.qpos[:] = mjm.key_qpos[keys[0]]
    hd.ctrl[:] = mjm.key_ctrl[keys[0]]
    d = mjw.put_data(mjm, hd, nworld=nworld, nconmax=nconmax, njmax=njmax)
    traj = torch.tensor(make_trajectory(mjm, keys), dtype=torch.float32, device="cuda")
  else:
    d = mjw.make_data(mjm, nworld=nworld, nconmax=nconmax, njmax=njmax)
    if mjm.nkey:
      d.qpos.copy_(torch.tensor(mjm.key_qpos[0], dtype=torch.float32, device="cuda").expand(nworld, -1))
  center = None
  if traj is None and mjm.nkey and getattr(mjm, "key_ctrl", None) is not None and mjm.nu:  # noise centred on the keyframe controls (testspeed.py:286)
    center = torch.tensor(mjm.key_ctrl[0], dtype=torch.float32, device="cuda").expand(nworld, -1).contiguous()
    d.ctrl.copy_(center)
  for i in range(20):
    if traj is None:
      mjw.ctrl_noise(m, d, i, ctrl_center=center)
    mjw.step(m, d)
  torch.cuda.synchronize()
  g = torch.cuda.CUDAGraph()
  s = torch.cuda.Stream()
  with torch.cuda.stream(s):
    mjw.step(m, d)
    with torch.cuda.graph(g, stream=s):
      mjw.step(m, d)
  torch.cuda.synchronize()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  for _ in range(10):
    g.replay()
  if traj is None:
    e0.record()
    for i in range(nstep):
      g.replay()
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1) / nstep
  else:  # controls of step i set before it, outside the timed region (benchmark.py:139-169 times each step on its own)
    ms = 0.0
    with torch.cuda.stream(s):
      for i in range(nstep):
        d.ctrl.copy_(traj[min(20 + i, traj.shape[0] - 1)].expand(nworld, -1))
        e0.record(s)
        g.replay()
        e1.record(s)
        e1.synchronize()
        ms += e0.elapsed_time(e1)
    ms /= nstep
  # per stage-group kernel time (CUDA events around every launch of 20 eager steps)
  from mujoco_warp_b200 import _lib
  lib = _lib.load()
  lib.mjb_timing_reset()
  lib.mjb_timing_enable(1)
  for _ in range(20):
    mjw.step(m, d)
  torch.cuda.synchronize()
  groups = {lib.mjb_group_name(g).decode(): round(lib.mjb_timing_group_ms(g) / 20, 4) for g in range(lib.mjb_group_count())}
  lib.mjb_timing_enable(0)
  lib.mjb_timing_reset()
  out = {"groups_ms": groups, "model": path, "nworld": nworld, "nconmax": nconmax, "njmax": njmax, "steps": nstep, "ms_per_step": ms,
         "env_steps_per_s": nworld / ms * 1e3, "mean_nefc": float(d.nefc.float().mean()), "mean_niter": float(d.solver_niter.float().mean()), "mean_ncon": float(d.nacon[0]) / nworld,
         "nan_worlds": int(torch.isnan(d.qpos).any(dim=1).sum()), "status_bits": int(d.status.max())}  # fmt: skip
  print(json.dumps(out))


if __name__ == "__main__":
  main()
