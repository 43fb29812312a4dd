"""bench.py -- env-steps/s of the batched physics step on N B200s (see BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--nworld 8192] [--configs humanoid,panda,...]

Headline line (metric / value / e2e / roofline / cpu_baseline): benchmarks/humanoid/humanoid.xml, 8192 worlds PER GPU (weak scaling:
each rank owns its own worlds, no per-step collective), one batched `mjw.step` per step preceded by the benchmark's control-noise
kernel (reference benchmark.py:42), timed in steady state (>= 250 steps after the keyframe start).  The same JSON line also carries
  "strong":     8192 worlds TOTAL split over the N ranks, with its own ms_per_step,
  "randomised": the humanoid from randomised initial states (SURVEY 8d recipe, seed 1234 + world),
  "configs":    BASELINE.json configs 3-5 (panda 8192, apollo_flat 4096, aloha_pot 4096 with the lift_pot replay), each with
                value, per-group kernel times, its own byte-model roofline and a CPU baseline from the oracle.
One JSON line is printed by rank 0.
"""

from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HUMANOID = os.path.join(ROOT, "benchmarks", "humanoid", "humanoid.xml")
NCONMAX, NJMAX = 24, 64  # reference benchmarks/config.txt:23
ALG_BYTES_PER_ENV_STEP = 32000  # contract traffic of one humanoid env-step
SETTLE_STEPS = 250  # untimed steps between the keyframe start and every timed window (steady state, whatever --steps is)
# BASELINE.json configs (sizes: reference benchmarks/config.txt:15-24 at BASELINE.json's nworld)
CONFIGS = {
  "humanoid": dict(path="benchmarks/humanoid/humanoid.xml", nworld=8192, nconmax=24, njmax=64, replay=None),
  "panda": dict(path="benchmarks/franka_emika_panda/scene.xml", nworld=8192, nconmax=1, njmax=5, replay=None),
  "apollo_flat": dict(path="benchmarks/apptronik_apollo/scene_flat.xml", nworld=4096, nconmax=16, njmax=64, replay=None, disable=("NATIVECCD",),
                      note="box-box pairs carry a margin: the reference's put_model only accepts this scene with DisableBit.NATIVECCD "
                           "(io.py:386-395), i.e. box-box through the box_box primitive, which is the route timed here"),
  "aloha_pot": dict(path="benchmarks/aloha_pot/scene.xml", nworld=4096, nconmax=24, njmax=128, replay="lift_pot"),
  # not one of BASELINE.json's five configs (reference benchmarks/config.txt:18): the same robot on a 1121 x 588 height field; measured
  # on request (--configs apollo_hfield), not part of the default line
  "apollo_hfield": dict(path="benchmarks/apptronik_apollo/scene_hfield.xml", nworld=4096, nconmax=32, njmax=128, replay=None, disable=("NATIVECCD",)),
  # reference benchmarks/config.txt:22 (nv = 81: the reference's sparse path; stored dense here), on request (--configs three_humanoids)
  "three_humanoids": dict(path="benchmarks/humanoid/three_humanoids.xml", nworld=8192, nconmax=100, njmax=192, replay=None),
}
# Algorithmic traffic of each stage-group kernel: Data fields it must read (produced by an earlier
# kernel or the caller) and fields it defines, write-once/read-once, capacity-sized for the pooled /
# padded arrays (DESIGN.md "Roofline accounting").  bytes/world = 4 x sum of per-world element counts.
GROUP_IO = {
  "smooth_position": (["qpos"],
                      ["xpos", "xquat", "xmat", "xipos", "ximat", "xanchor", "xaxis", "geom_xpos", "geom_xmat", "site_xpos", "site_xmat",
                       "subtree_com", "cdof", "cinert", "crb", "qM", "actuator_length", "actuator_moment"]),
  "collision": (["geom_xpos", "geom_xmat"],
                ["contact_dist", "contact_pos", "contact_frame", "contact_includemargin", "contact_friction", "contact_solref",
                 "contact_solreffriction", "contact_solimp", "contact_dim", "contact_geom", "contact_worldid", "contact_type",
                 "contact_geomcollisionid"]),
  "make_constraint": (["qpos", "qvel", "cdof", "subtree_com", "contact_dist", "contact_pos", "contact_frame", "contact_includemargin",
                       "contact_friction", "contact_solref", "contact_solimp", "contact_dim", "contact_geom"],
                      ["efc_J", "efc_pos", "efc_margin", "efc_D", "efc_vel", "efc_aref", "efc_frictionloss", "efc_type", "efc_id",
                       "contact_efc_address"]),
  "smooth_velocity": (["qpos", "qvel", "ctrl", "act", "cdof", "cinert", "qM", "actuator_length", "actuator_moment", "qfrc_applied",
                       "xfrc_applied", "xipos", "subtree_com"],
                      ["actuator_velocity", "cvel", "cdof_dot", "qfrc_spring", "qfrc_damper", "qfrc_passive", "qfrc_bias", "cacc",
                       "cfrc_int", "actuator_force", "qfrc_actuator", "act_dot", "qfrc_smooth", "qacc_smooth", "qLD"]),
  "solve": (["efc_J", "efc_D", "efc_aref", "efc_frictionloss", "qM", "qacc_warmstart", "qfrc_smooth", "qacc_smooth"],
            ["efc_force", "efc_state", "qfrc_constraint", "qacc", "efc_Ma", "solver_niter"]),
  "euler": (["qM", "efc_Ma", "qacc", "qpos", "qvel", "act", "act_dot"], ["qpos", "qvel", "act", "qacc_warmstart", "time"]),
}


def group_alg_bytes(d, nworld):
  """Algorithmic bytes per world of every stage-group kernel (see GROUP_IO)."""
  out = {}
  for g, (rd, wr) in GROUP_IO.items():
    n = 0
    for name in rd + wr:
      n += d.field(name).numel() * d.field(name).element_size()
    out[g] = n / nworld
  return out


METRIC = "humanoid env-steps/sec at nworld=8192, 1/2/4/8 B200; % of HBM roofline"


def workload_name(nworld):
  return (f"benchmarks/humanoid/humanoid.xml nworld={nworld} per GPU, nconmax=24 njmax=64, Newton solver, dense Jacobian, Euler, "
          "keyframe 'squat' + OU/Halton ctrl noise")  # fmt: skip


def _ncu_traffic(kernel):
  """DRAM bytes per launch of `kernel` from the newest committed `ncu --set full` summary (profiles/rN_ncu_full_groups.json)."""
  import glob

  files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_groups.json")))
  if not files:
    return None, None
  with open(files[-1]) as f:
    e = json.load(f).get(kernel)
  return (e["dram_bytes_per_launch"], os.path.relpath(files[-1], ROOT)) if e else (None, None)


def _peaks():
  p = os.path.join(ROOT, "MEASURED_PEAKS.json")
  if os.path.exists(p):
    with open(p) as f:
      return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
  return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
  """Samples nvidia-smi SM clocks / throttle reasons during the timed region."""

  def __init__(self, index=0):
    self.index, self.samples, self.reasons, self.maxmhz = index, [], set(), None
    self._stop = threading.Event()
    self._t = threading.Thread(target=self._run, daemon=True)

  def _run(self):
    q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")  # fmt: skip
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    while not self._stop.is_set():
      try:
        out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                             capture_output=True, text=True, timeout=5).stdout.strip().split("\n")[0]  # fmt: skip
        f = [x.strip() for x in out.split(",")]
        self.samples.append(float(f[0]))
        self.maxmhz = float(f[1])
        for n, v in zip(names, f[2:]):
          if v.lower().startswith("active"):
            self.reasons.add(n)
      except Exception:
        pass
      self._stop.wait(0.1)

  def __enter__(self):
    self._t.start()
    return self

  def __exit__(self, *a):
    self._stop.set()
    self._t.join(timeout=6)

  def summary(self):
    return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.maxmhz, "reasons": sorted(self.reasons)}


def _ncu_binding(kernel):
  """Binding-resource view of `kernel` from the newest committed `ncu --set full` summary: the kernels are issue / latency bound, not
  bandwidth bound, so the HBM fraction alone invites the wrong optimisation."""
  import glob

  files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_groups.json")))
  if not files:
    return None
  with open(files[-1]) as f:
    e = json.load(f).get(kernel)
  if not e:
    return None
  pick = {
    "issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "inst_per_cycle": "sm__inst_executed.avg.per_cycle_elapsed",
    "fp32_pipe_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "registers": "launch__registers_per_thread",
    "stall_long_scoreboard": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "stall_no_instruction": "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
  }
  out = {k: e[v] for k, v in pick.items() if v in e}
  out["source"] = os.path.relpath(files[-1], ROOT)
  return out


def load_config(name):
  import mujoco_warp_b200 as mjw

  cfg = CONFIGS[name]
  mjm = mjw.load_mjcf(os.path.join(ROOT, cfg["path"]))
  for flag in cfg.get("disable", ()):
    mjm.opt.disableflags |= int(getattr(mjw.DisableBit, flag))
  keys = mjw.find_keys(mjm, cfg["replay"]) if cfg["replay"] else ([0] if mjm.nkey else [])
  return cfg, mjm, keys


def randomised_states(mjm, nworld, first_world=0):
  """SURVEY 8(d) randomised-state variant: qpos = key_qpos + U(-0.05, 0.05) on hinge / slide coordinates, root z += U(0, 0.3),
  root quat = normalised(key_quat + N(0, 0.05)), qvel ~ U(-0.1, 0.1), numpy default_rng(seed = 1234 + worldid)."""
  base = np.asarray(mjm.key_qpos[0] if mjm.nkey else mjm.qpos0, dtype=np.float64)
  qpos = np.repeat(base[None], nworld, axis=0)
  qvel = np.zeros((nworld, mjm.nv))
  jt, qa = np.asarray(mjm.jnt_type), np.asarray(mjm.jnt_qposadr)
  for w in range(nworld):
    rng = np.random.default_rng(1234 + first_world + w)
    for j in range(mjm.njnt):
      a = int(qa[j])
      if jt[j] == 0:
        qpos[w, a + 2] += rng.uniform(0.0, 0.3)
        q = qpos[w, a + 3 : a + 7] + rng.normal(0.0, 0.05, 4)
        qpos[w, a + 3 : a + 7] = q / np.linalg.norm(q)
      elif jt[j] in (2, 3):
        qpos[w, a] += rng.uniform(-0.05, 0.05)
    qvel[w] = rng.uniform(-0.1, 0.1, mjm.nv)
  return qpos, qvel


def _cpu_sample(name, nthread):
  """Times the CPU oracle (float64 C restatement of the reference path) on a bounded sample of a config's workload: `nw` worlds
  from the config's start keyframe stepped with OpenMP over worlds.  Returns a stepping closure and the measured step rate."""
  from oracle import oracle as O

  cfg, mjm, keys = load_config(name)
  om = O.OracleModel(mjm)
  nw = max(64, 16 * nthread) if name != "aloha_pot" else max(16, 2 * nthread)  # aloha: ~9000 candidate pairs per world on the CPU
  d = O.make_data(om, nworld=nw, nconmax=cfg["nconmax"], njmax=cfg["njmax"])
  if keys:
    d.qpos[:] = mjm.key_qpos[keys[0]]
    d.ctrl[:] = mjm.key_ctrl[keys[0]]
  if not cfg["replay"]:
    rng = np.random.default_rng(0)
    d.ctrl[:] += rng.uniform(-0.1, 0.1, d.ctrl.shape)
  stage = "step_implicitfast" if int(mjm.opt.integrator) == 3 else "step"

  def step():
    O.run(om, d, stage, nthread=nthread)

  for _ in range(2):
    step()
  t0 = time.perf_counter()
  n = 0
  while time.perf_counter() - t0 < 1.0 or n < 2:
    step()
    n += 1
  return step, nw, n / (time.perf_counter() - t0)


def cpu_baseline(name, nthread, target_s=12.0):
  step, nw, rate = _cpu_sample(name, nthread)
  ns = int(max(2, min(100000, rate * target_s)))
  t0 = time.perf_counter()
  for _ in range(ns):
    step()
  dt = time.perf_counter() - t0
  return {
    "value": nw * ns / dt, "unit": "env-steps/s", "cores": nthread, "kind": "port",
    "sample": f"{nw} {name} worlds x {ns} steps from the start keyframe with fixed ctrl, float64 C oracle, OpenMP over worlds ({dt:.1f} s)",
  }  # fmt: skip


def run_reference(args, rank):
  """--impl reference: the CPU restatement of the reference path (MuJoCo C / Warp are not
  installable offline, so kind == 'port'), all host threads, bounded sample per step."""
  if rank != 0:
    return
  nthread = os.cpu_count() or 1
  step, nw, rate = _cpu_sample("humanoid", nthread)
  # bounded sample per "step": nw worlds; keep the whole run within a couple of minutes
  steps, warmup = args.steps, args.warmup
  budget = 120.0
  if (steps + warmup) / max(rate, 1e-9) > budget:
    steps = max(1, int(budget * rate) - warmup)
  for _ in range(warmup):
    step()
  t0 = time.perf_counter()
  for _ in range(steps):
    step()
  dt = time.perf_counter() - t0
  value = nw * steps / dt
  cb = {"value": value, "unit": "env-steps/s", "cores": nthread, "kind": "port",
        "sample": f"{nw} humanoid worlds per step (bounded sample of the nworld=8192 workload), float64 C oracle, OpenMP over worlds"}  # fmt: skip
  print(json.dumps({
    "impl": "reference", "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": steps,
    "warmup": warmup, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
    "dtype": "f64", "data": "synthetic", "config": {"workload": workload_name(args.nworld), "sample": f"each step advances a bounded sample of {nw} of the {args.nworld} worlds on the host cores"},
    "cpu_baseline": cb, "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
  }))  # fmt: skip


class Workload:
  """One benchmark config resident on the current GPU: model, nworld-batched data, the captured `step` graph and the per-step control
  update (OU / Halton noise around the keyframe controls, or the keyframe-replay trajectory)."""

  def __init__(self, name, nworld, rank=0, first_world=0, randomise=False, use_graph=True):
    import torch

    import mujoco_warp_b200 as mjw
    from mujoco_warp_b200 import _lib

    self.torch, self.mjw, self.lib = torch, mjw, _lib.load()
    self.name, self.nworld, self.rank = name, nworld, rank
    self.cfg, self.mjm, keys = load_config(name)
    mjm = self.mjm
    self.m = mjw.put_model(mjm)
    hd = mjw.make_host_data(mjm)
    if keys:  # keyframe 0 (or the first replay keyframe) replicated, reference testspeed.py:277-287
      hd.qpos[:] = mjm.key_qpos[keys[0]]
      hd.ctrl[:] = mjm.key_ctrl[keys[0]]
    self.d = mjw.put_data(mjm, hd, nworld=nworld, nconmax=self.cfg["nconmax"], njmax=self.cfg["njmax"], m=self.m)
    if randomise:
      qpos, qvel = randomised_states(mjm, nworld, first_world)
      self.d.qpos.copy_(torch.from_numpy(qpos.astype(np.float32)))
      self.d.qvel.copy_(torch.from_numpy(qvel.astype(np.float32)))
    self.center = torch.from_numpy(np.asarray(hd.ctrl, dtype=np.float32)).cuda() if mjm.nu else None
    self.traj = None
    if self.cfg["replay"]:
      self.traj = torch.from_numpy(mjw.make_trajectory(mjm, keys).astype(np.float32)).cuda()
    mjw.forward(self.m, self.d)
    self.d.qacc_warmstart.copy_(self.d.qacc)
    torch.cuda.synchronize()
    self.stream = torch.cuda.Stream()
    self.step_no = 0
    self.graph = None
    self.use_graph = use_graph
    self.launches_per_step = None

  def ctrl_update(self):
    """Per-step controls, outside the captured graph (the step index changes): benchmark.py:42 ctrl_noise / testspeed replay."""
    if self.traj is not None:
      self.d.ctrl.copy_(self.traj[self.step_no % self.traj.shape[0]].expand_as(self.d.ctrl))
    elif self.mjm.nu:
      self.mjw.ctrl_noise(self.m, self.d, self.step_no + self.rank * 100003, ctrl_center=self.center)
    self.step_no += 1

  def step(self):
    self.ctrl_update()
    if self.graph is not None:
      self.graph.replay()
    else:
      self.mjw.step(self.m, self.d)

  def run(self, n):
    for _ in range(n):
      self.step()

  def capture(self):
    """Captures `mjw.step` once (reference benchmark.py:124-137) and counts its kernel launches."""
    n0 = self.lib.mjb_launch_count()
    if self.use_graph:
      self.graph = self.torch.cuda.CUDAGraph()
      with self.torch.cuda.graph(self.graph, stream=self.stream):
        self.mjw.step(self.m, self.d)
      self.launches_per_step = self.lib.mjb_launch_count() - n0
    else:
      self.mjw.step(self.m, self.d)
      self.launches_per_step = self.lib.mjb_launch_count() - n0
      self.step_no += 1

  def group_times(self, nk):
    """Average kernel milliseconds per step of every stage group over `nk` eager steps (CUDA events around each launch)."""
    lib = self.lib
    graph, self.graph = self.graph, None
    lib.mjb_timing_reset()
    lib.mjb_timing_enable(1)
    self.run(nk)
    self.stream.synchronize()
    group_ms = {lib.mjb_group_name(g).decode(): lib.mjb_timing_group_ms(g) / nk for g in range(lib.mjb_group_count())}
    kernel_ms = lib.mjb_timing_ms() / nk
    lib.mjb_timing_enable(0)
    lib.mjb_timing_reset()
    self.graph = graph
    return group_ms, kernel_ms

  def timed(self, K, dist=None, clock_index=None):
    """K steps bracketed by barrier + synchronize, CUDA events on the launching stream; returns (ms, clocks summary or None)."""
    torch = self.torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if dist is not None:
      dist.barrier()
    torch.cuda.synchronize()
    clk = ClockSampler(clock_index) if clock_index is not None else None
    if clk:
      clk.__enter__()
    e0.record(self.stream)
    self.run(K)
    e1.record(self.stream)
    e1.synchronize()
    if clk:
      clk.__exit__()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), (clk.summary() if clk else None)

  def alg_bytes_per_env_step(self):
    """SURVEY 8(d) recipe for any config: caller inputs read once + every Data field `step` defines written once, per world."""
    d = self.d
    names = {"qpos", "qvel", "ctrl", "qacc_warmstart", "qfrc_applied", "xfrc_applied", "time"}
    for _, wr in GROUP_IO.values():
      names.update(wr)
    return sum(d.field(n).numel() * d.field(n).element_size() for n in names) / self.nworld

  def health(self):
    torch = self.torch
    return int(torch.isnan(self.d.qpos).any(dim=1).sum()), int(self.d.status.max()) if self.nworld else 0


def bench_secondary(name, nworld, K, W, rank, peak, with_cpu, nthread):
  """One BASELINE.json config beyond the headline: device-resident value, per-group kernel times, byte-model roofline, CPU baseline."""
  w = Workload(name, nworld, rank)
  torch = w.torch
  with torch.cuda.stream(w.stream):
    w.run(W)
    w.capture()
    w.run(SETTLE_STEPS)
    w.stream.synchronize()
    group_ms, kernel_ms = w.group_times(min(K, 20))
    ms, _ = w.timed(K)
  nan_worlds, status = w.health()
  value = nworld * K / (ms * 1e-3)
  balg = w.alg_bytes_per_env_step()
  gbytes = group_alg_bytes(w.d, nworld)
  dom = max(group_ms, key=lambda g: group_ms[g])
  out = {
    "workload": f"{CONFIGS[name]['path']} nworld={nworld} nconmax={CONFIGS[name]['nconmax']} njmax={CONFIGS[name]['njmax']}"
                + (f" replay={CONFIGS[name]['replay']}" if CONFIGS[name]["replay"] else ""),
    "value": value, "unit": "env-steps/s", "ms_per_step": ms / K, "steps": K, "settle_steps": SETTLE_STEPS, "nan_worlds": nan_worlds,
    "status_bits": status, "gpu_launches_per_step": int(w.launches_per_step),
    "mean_nefc": float(w.d.nefc.float().mean()), "mean_ncon": float(w.d.nacon[0]) / nworld,
    "mean_solver_niter": float(w.d.solver_niter.float().mean()),
    "roofline": {"bound": "hbm", "alg_bytes_per_env_step": balg, "achieved": balg * value / 1e9, "peak": peak, "unit": "GB/s",
                 "frac": balg * value / 1e9 / peak, "kernel": f"mjb_group_kernel<{dom}>", "kernel_ms": group_ms[dom],
                 "kernel_frac": gbytes[dom] * nworld / (group_ms[dom] * 1e-3) / 1e9 / peak,
                 "groups_ms": group_ms, "kernels_ms": kernel_ms},
  }  # fmt: skip
  if "note" in CONFIGS[name]:
    out["note"] = CONFIGS[name]["note"]
  if with_cpu:
    out["cpu_baseline"] = cpu_baseline(name, nthread, target_s=5.0)
  del w
  torch.cuda.empty_cache()
  return out


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=200)
  ap.add_argument("--warmup", type=int, default=10)
  ap.add_argument("--impl", default="b200")
  ap.add_argument("--nworld", type=int, default=8192)
  ap.add_argument("--configs", default="humanoid,randomised,strong,panda,apollo_flat,aloha_pot",
                  help="comma list of the blocks to measure besides the headline (headline = humanoid, always)")
  ap.add_argument("--no-cpu-baseline", action="store_true")
  ap.add_argument("--no-graph", action="store_true")
  args = ap.parse_args()
  args.warmup = max(args.warmup, 3)
  blocks = set(args.configs.split(","))

  rank = int(os.environ.get("RANK", "0"))
  world = int(os.environ.get("WORLD_SIZE", "1"))
  local_rank = int(os.environ.get("LOCAL_RANK", "0"))
  if args.impl == "reference":
    run_reference(args, rank)
    return

  import torch

  if not torch.cuda.is_available():
    raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
  torch.cuda.set_device(local_rank)
  dist = None
  if world > 1:
    import torch.distributed as dist

    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
  nworld, K, W = args.nworld, args.steps, args.warmup
  nthread = os.cpu_count() or 1
  peak, peak_src = _peaks()

  # ---- headline: humanoid, nworld per GPU (weak scaling), steady state
  w = Workload("humanoid", nworld, rank, first_world=rank * nworld, use_graph=not args.no_graph)
  m, d, stream, lib = w.m, w.d, w.stream, w.lib
  with torch.cuda.stream(stream):
    w.run(W)
    w.capture()
    w.run(max(SETTLE_STEPS - W, 0))  # untimed: leave the all-worlds-identical keyframe transient behind
    stream.synchronize()
    nk = min(K, 50)
    group_ms, kernel_ms = w.group_times(nk)
    ms, clocks = w.timed(K, dist, local_rank)
    gpu_launches = K * w.launches_per_step

    # ---- end to end through the public API with HOST buffers: per step H2D of the controls from
    # pinned memory, mjw.step, D2H of the observation (qpos | qvel) into pinned memory
    import mujoco_warp_b200 as mjw

    graph = w.graph
    nu, nq, nv = m.nu, m.nq, m.nv
    # Host-side controls: NRING frames of the SAME Ornstein-Uhlenbeck / Halton process the device-resident loop applies (continued
    # from the current step index), generated once and parked in pinned host memory -- the workload of `e2e` must be the workload of
    # `value`.  (Uncorrelated +-0.1 controls, as an earlier version uploaded, shake the humanoids enough to cost the solver 25 %.)
    NRING = 32
    h_ctrl = [torch.empty((nworld, nu), dtype=torch.float32).pin_memory() for _ in range(NRING)]
    ctrl_keep = d.ctrl.clone()
    for h in h_ctrl:
      w.ctrl_update()
      h.copy_(d.ctrl)
    d.ctrl.copy_(ctrl_keep)
    stream.synchronize()
    h_obs = [torch.empty((nworld, nq + nv), dtype=torch.float32).pin_memory() for _ in range(2)]
    d_ctrl_stage = [torch.empty_like(d.ctrl) for _ in range(2)]
    d_obs = [torch.empty((nworld, nq + nv), dtype=torch.float32, device="cuda") for _ in range(2)]
    Ke = min(K, 100)
    copy_stream = torch.cuda.Stream()
    ev_ctrl = [torch.cuda.Event() for _ in range(2)]   # ctrl k uploaded
    ev_step = [torch.cuda.Event() for _ in range(2)]   # step k finished, observation k snapshotted
    ev_obs = [torch.cuda.Event() for _ in range(2)]    # observation k on the host

    # Software pipeline over two buffers, every transfer inside the timed region:
    #   copy stream:    H2D ctrl[k+1] into a staging buffer | D2H obs[k] (after step k)
    #   compute stream: ctrl <- staged ctrl[k] (mjw.set_state), mjw.step, snapshot qpos|qvel (mjw.get_state): one captured graph
    #   host:           waits for obs[k-1] while step k runs (one step of observation latency)
    def upload(k):
      with torch.cuda.stream(copy_stream):
        d_ctrl_stage[k & 1].copy_(h_ctrl[k % NRING], non_blocking=True)
        ev_ctrl[k & 1].record(copy_stream)

    # the per-step device work of the loop is ONE graph: ctrl <- staged ctrl (set_state), mjw.step, observation snapshot (get_state
    # packs qpos | qvel into one row per world); the reference's benchmark loop captures its step the same way (benchmark.py:124-137)
    sig_obs = int(mjw.State.QPOS) | int(mjw.State.QVEL)
    e2e_graphs = [None, None]
    if graph is not None:
      for b in range(2):
        e2e_graphs[b] = torch.cuda.CUDAGraph()
        with torch.cuda.graph(e2e_graphs[b], stream=stream):
          mjw.set_state(m, d, d_ctrl_stage[b], int(mjw.State.CTRL))
          mjw.step(m, d)
          mjw.get_state(m, d, d_obs[b], sig_obs)

    def e2e_step(k):
      upload(k + 1)
      stream.wait_event(ev_ctrl[k & 1])
      if e2e_graphs[k & 1] is not None:
        e2e_graphs[k & 1].replay()
      else:
        mjw.set_state(m, d, d_ctrl_stage[k & 1], int(mjw.State.CTRL))
        mjw.step(m, d)
        mjw.get_state(m, d, d_obs[k & 1], sig_obs)
      ev_step[k & 1].record(stream)
      with torch.cuda.stream(copy_stream):
        copy_stream.wait_event(ev_step[k & 1])
        h_obs[k & 1].copy_(d_obs[k & 1], non_blocking=True)
        ev_obs[k & 1].record(copy_stream)
      if k > 0:
        ev_obs[(k - 1) & 1].synchronize()  # the caller consumes observation k-1 here

    def e2e_run(n):
      upload(0)
      for k in range(n):
        e2e_step(k)
      ev_obs[(n - 1) & 1].synchronize()
      torch.cuda.synchronize()

    e2e_run(4)
    if dist is not None:
      dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_run(Ke)
    e2e_s = time.perf_counter() - t0

  nan_worlds, status = w.health()

  # ---- strong scaling: --nworld worlds TOTAL, contiguous shards over the ranks (sharding.world_shard); same steady-state protocol
  strong_ms = None
  if "strong" in blocks:
    from mujoco_warp_b200 import sharding

    lo, hi = sharding.world_shard(nworld, rank, world)
    if world == 1:
      strong_ms = ms  # one GPU: the strong and the weak workload coincide
    else:
      ws = Workload("humanoid", hi - lo, rank, first_world=lo)
      with torch.cuda.stream(ws.stream):
        ws.run(W)
        ws.capture()
        ws.run(max(SETTLE_STEPS - W, 0))
        strong_ms, _ = ws.timed(K, dist)
      del ws

  # ---- randomised initial states, same sizes, 1 GPU's worth per rank
  rand = None
  if "randomised" in blocks:
    wr = Workload("humanoid", nworld, rank, first_world=rank * nworld, randomise=True)
    with torch.cuda.stream(wr.stream):
      wr.run(W)
      wr.capture()
      wr.run(max(SETTLE_STEPS - W, 0))
      rand_ms, _ = wr.timed(K, dist)
    rand = (rand_ms, *wr.health(), float(wr.d.nefc.float().mean()))
    del wr
  torch.cuda.empty_cache()

  gnames = list(group_ms)
  vals = [ms, e2e_s * 1e3, kernel_ms, strong_ms or 0.0, rand[0] if rand else 0.0] + [group_ms[g] for g in gnames]
  t = torch.tensor(vals, dtype=torch.float64, device="cuda")
  if dist is not None:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # optional observation gather over NVLink (the only collective of the path; outside timing)
    obs = torch.cat([d.qpos, d.qvel], dim=1)
    gathered = torch.empty((world * nworld, obs.shape[1]), dtype=obs.dtype, device="cuda")
    dist.all_gather_into_tensor(gathered, obs)
  ms, e2e_ms, kernel_ms, strong_ms, rand_ms = [float(x) for x in t.tolist()[:5]]
  group_ms = {g: float(x) for g, x in zip(gnames, t.tolist()[5:])}

  # ---- BASELINE.json configs 3-5 (rank 0 only, one GPU: they are reported per GPU; the worlds shard like the humanoid's)
  secondary = {}
  if rank == 0:
    for name in [k for k in CONFIGS if k != "humanoid"]:
      if name in blocks:
        secondary[name] = bench_secondary(name, CONFIGS[name]["nworld"], max(20, min(K, 100)), W, rank, peak,
                                          with_cpu=(not args.no_cpu_baseline and world == 1), nthread=nthread)

  if rank == 0:
    value = world * nworld * K / (ms * 1e-3)
    e2e_value = world * nworld * Ke / (e2e_ms * 1e-3)
    gbytes = group_alg_bytes(d, nworld)
    dom = max(group_ms, key=lambda g: group_ms[g])  # the dominant kernel of the step
    achieved = gbytes[dom] * nworld / (group_ms[dom] * 1e-3) / 1e9
    step_achieved = ALG_BYTES_PER_ENV_STEP * nworld / (kernel_ms * 1e-3) / 1e9
    traffic, traffic_src = _ncu_traffic(f"mjb_group_kernel<{dom}>")
    out = {
      "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
      "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
      "data": "synthetic",
      "config": {
        "workload": workload_name(nworld),
        "parallelism": f"worlds sharded over {world} GPU(s), no per-step collective",
        "l2": "per-GPU working set (~40 kB/world x nworld = ~330 MB) exceeds the 126 MB L2; no explicit flush",
        "steady_state": f"{SETTLE_STEPS} untimed steps (warm-up included) run between the keyframe start and the timed window",
        "cuda_graph": w.graph is not None, "nan_worlds": nan_worlds, "status_bits": status,
      },
      "clocks": clocks,
      "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": nworld * nu * 4, "d2h_bytes_per_step": nworld * (nq + nv) * 4,
              "steps": Ke},
      "gpu_launches": int(gpu_launches),
      "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                   "kernel": f"mjb_group_kernel<{dom}>", "kernel_ms": group_ms[dom], "share_of_step": group_ms[dom] / kernel_ms,
                   "alg_bytes_per_launch": gbytes[dom] * nworld, "peak_source": peak_src,
                   "binding_resource": _ncu_binding(f"mjb_group_kernel<{dom}>"),
                   "step": {"kernels_ms": kernel_ms, "alg_bytes_per_env_step": ALG_BYTES_PER_ENV_STEP, "achieved": step_achieved,
                            "frac": step_achieved / peak},
                   "groups": {g: {"ms": group_ms[g], "alg_bytes_per_world": gbytes[g],
                                  "frac": gbytes[g] * nworld / (group_ms[g] * 1e-3) / 1e9 / peak if group_ms[g] > 0 else None}
                              for g in gnames}},
    }  # fmt: skip
    if "strong" in blocks:
      out["strong"] = {"nworld_total": nworld, "nworld_per_gpu": nworld // world, "value": nworld * K / (strong_ms * 1e-3), "unit": "env-steps/s",
                       "ms_per_step": strong_ms / K, "scaling": "strong",
                       "limiter": "per-step latency of the six dependent stage-group launches with few worlds per SM (no collective on the path)"}  # fmt: skip
    if rand is not None:
      out["randomised"] = {"value": world * nworld * K / (rand_ms * 1e-3), "unit": "env-steps/s", "ms_per_step": rand_ms / K,
                           "nan_worlds": rand[1], "status_bits": rand[2], "mean_nefc": rand[3],
                           "states": "key_qpos + U(-0.05,0.05) hinges, root z + U(0,0.3), root quat + N(0,0.05) normalised, qvel U(-0.1,0.1), default_rng(1234 + world)"}  # fmt: skip
    if secondary:
      out["configs"] = secondary
    if not args.no_cpu_baseline and world == 1:
      out["cpu_baseline"] = cpu_baseline("humanoid", nthread)
    print(json.dumps(out))
  if dist is not None:
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
  main()
