"""bench.py -- humanoid env-steps/s of the batched physics step on N B200s (see BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--nworld 8192]

A "step" is one batched `mjw.step` over nworld=8192 humanoid worlds per GPU (weak scaling: each
rank owns its own 8192 worlds, no per-step collective), preceded by the benchmark's control-noise
kernel (reference benchmark.py:42).  One JSON line is printed by rank 0.
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


def _cpu_sample(nthread, target_s=12.0):
  """Times the CPU oracle (float64 C restatement of the reference path) on a bounded sample of the
  same workload: `nw` humanoid worlds from keyframe 0 stepped `ns` times with OpenMP over worlds."""
  import mujoco_warp_b200 as mjw
  from oracle import oracle as O

  mjm = mjw.load_mjcf(HUMANOID)
  om = O.OracleModel(mjm)
  nw = max(64, 16 * nthread)
  d = O.make_data(om, nworld=nw, nconmax=NCONMAX, njmax=NJMAX)
  d.qpos[:] = mjm.key_qpos[0]
  rng = np.random.default_rng(0)
  d.ctrl[:] = rng.uniform(-0.1, 0.1, d.ctrl.shape)
  for _ in range(5):
    O.run(om, d, "step", nthread=nthread)
  t0 = time.perf_counter()
  n = 0
  while time.perf_counter() - t0 < 1.0:
    O.run(om, d, "step", nthread=nthread)
    n += 1
  rate = n / (time.perf_counter() - t0)
  return om, d, nw, rate


def cpu_baseline(nthread):
  from oracle import oracle as O

  om, d, nw, rate = _cpu_sample(nthread)
  ns = int(max(20, min(100000, rate * 12.0)))
  t0 = time.perf_counter()
  for _ in range(ns):
    O.run(om, d, "step", nthread=nthread)
  dt = time.perf_counter() - t0
  return {
    "value": nw * ns / dt, "unit": "env-steps/s", "cores": nthread, "kind": "port",
    "sample": f"{nw} humanoid worlds x {ns} steps from keyframe 'squat' with fixed random ctrl, float64 C oracle, OpenMP over worlds ({dt:.1f} s)",
  }  # fmt: skip


def run_reference(args, rank):
  """--impl reference: the CPU restatement of the reference path (MuJoCo C / Warp are not
  installable offline, so kind == 'port'), all host threads, bounded sample per step."""
  if rank != 0:
    return
  from oracle import oracle as O

  nthread = os.cpu_count() or 1
  om, d, nw, rate = _cpu_sample(nthread)
  # bounded sample per "step": nw worlds; keep the whole run within a couple of minutes
  steps, warmup = args.steps, args.warmup
  budget = 120.0
  if (steps + warmup) / max(rate, 1e-9) > budget:
    steps = max(1, int(budget * rate) - warmup)
  for _ in range(warmup):
    O.run(om, d, "step", nthread=nthread)
  t0 = time.perf_counter()
  for _ in range(steps):
    O.run(om, d, "step", nthread=nthread)
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


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=200)
  ap.add_argument("--warmup", type=int, default=10)
  ap.add_argument("--impl", default="b200")
  ap.add_argument("--nworld", type=int, default=8192)
  ap.add_argument("--no-cpu-baseline", action="store_true")
  ap.add_argument("--no-graph", action="store_true")
  args = ap.parse_args()
  args.warmup = max(args.warmup, 3)

  rank = int(os.environ.get("RANK", "0"))
  world = int(os.environ.get("WORLD_SIZE", "1"))
  local_rank = int(os.environ.get("LOCAL_RANK", "0"))
  if args.impl == "reference":
    run_reference(args, rank)
    return

  import torch

  import mujoco_warp_b200 as mjw
  from mujoco_warp_b200 import _lib

  if not torch.cuda.is_available():
    raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
  torch.cuda.set_device(local_rank)
  dist = None
  if world > 1:
    import torch.distributed as dist

    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
  lib = _lib.load()
  nworld, K, W = args.nworld, args.steps, args.warmup

  mjm = mjw.load_mjcf(HUMANOID)
  m = mjw.put_model(mjm)
  hd = mjw.make_host_data(mjm)
  hd.qpos[:] = mjm.key_qpos[0]  # keyframe 0 'squat' replicated (reference testspeed.py:284-287)
  d = mjw.put_data(mjm, hd, nworld=nworld, nconmax=NCONMAX, njmax=NJMAX, m=m)
  mjw.forward(m, d)
  d.qacc_warmstart.copy_(d.qacc)
  torch.cuda.synchronize()

  stream = torch.cuda.Stream()
  step_no = [0]

  def one_step():
    mjw.ctrl_noise(m, d, step_no[0] + rank * 100003)
    mjw.step(m, d)
    step_no[0] += 1

  with torch.cuda.stream(stream):
    for _ in range(W):
      one_step()
    stream.synchronize()

    # ---- kernel duration for the roofline (CUDA events around the pipeline kernel, no graph)
    lib.mjb_timing_reset()
    lib.mjb_timing_enable(1)
    nk = min(K, 50)
    for _ in range(nk):
      one_step()
    stream.synchronize()
    group_ms = {lib.mjb_group_name(g).decode(): lib.mjb_timing_group_ms(g) / nk for g in range(lib.mjb_group_count())}
    kernel_ms = lib.mjb_timing_ms() / nk  # all kernels of one step
    lib.mjb_timing_enable(0)
    lib.mjb_timing_reset()

    # ---- timed region: K steps, CUDA events on the launching stream, barrier + sync on both sides
    launches0 = lib.mjb_launch_count()
    graph = None
    if not args.no_graph:
      # capture `step` once (reference benchmark.py:124-137); the noise kernel stays outside the
      # graph because its step index changes
      graph = torch.cuda.CUDAGraph()
      with torch.cuda.graph(graph, stream=stream):
        mjw.step(m, d)
      launches_per_graph = lib.mjb_launch_count() - launches0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if dist is not None:
      dist.barrier()
    torch.cuda.synchronize()
    launches0 = lib.mjb_launch_count()
    with ClockSampler(local_rank) as clk:
      e0.record(stream)
      for _ in range(K):
        if graph is not None:
          mjw.ctrl_noise(m, d, step_no[0] + rank * 100003)
          graph.replay()
          step_no[0] += 1
        else:
          one_step()
      e1.record(stream)
      e1.synchronize()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    gpu_launches = (K * launches_per_graph) if graph is not None else (lib.mjb_launch_count() - launches0)

    # ---- end to end through the public API with HOST buffers: per step H2D of the controls from
    # pinned memory, mjw.step, D2H of the observation (qpos | qvel) into pinned memory
    nu, nq, nv = m.nu, m.nq, m.nv
    h_ctrl = [torch.empty((nworld, nu), dtype=torch.float32).pin_memory() for _ in range(2)]
    for h in h_ctrl:
      h.uniform_(-0.1, 0.1)
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
    #   compute stream: ctrl <- staged ctrl[k], mjw.step (public API), snapshot qpos|qvel
    #   host:           waits for obs[k-1] while step k runs (one step of observation latency)
    def upload(k):
      with torch.cuda.stream(copy_stream):
        d_ctrl_stage[k & 1].copy_(h_ctrl[k & 1], non_blocking=True)
        ev_ctrl[k & 1].record(copy_stream)

    def e2e_step(k):
      upload(k + 1)
      stream.wait_event(ev_ctrl[k & 1])
      d.ctrl.copy_(d_ctrl_stage[k & 1], non_blocking=True)
      if graph is not None:
        graph.replay()  # mjw.step captured once, as the reference's benchmark loop does (benchmark.py:124-137)
      else:
        mjw.step(m, d)
      d_obs[k & 1][:, :nq].copy_(d.qpos, non_blocking=True)
      d_obs[k & 1][:, nq:].copy_(d.qvel, non_blocking=True)
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

  nan_worlds = int(torch.isnan(d.qpos).any(dim=1).sum())
  status = int(torch.bitwise_or(d.status, torch.zeros_like(d.status)).max()) if nworld else 0
  gnames = list(group_ms)
  t = torch.tensor([ms, e2e_s * 1e3, kernel_ms] + [group_ms[g] for g in gnames], dtype=torch.float64, device="cuda")
  if dist is not None:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # optional observation gather over NVLink (the only collective of the path; outside timing)
    obs = torch.cat([d.qpos, d.qvel], dim=1)
    gathered = torch.empty((world * nworld, obs.shape[1]), dtype=obs.dtype, device="cuda")
    dist.all_gather_into_tensor(gathered, obs)
  ms, e2e_ms, kernel_ms = [float(x) for x in t.tolist()[:3]]
  group_ms = {g: float(x) for g, x in zip(gnames, t.tolist()[3:])}

  if rank == 0:
    peak, peak_src = _peaks()
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
        "cuda_graph": graph is not None, "nan_worlds": nan_worlds, "status_bits": status,
      },
      "clocks": clk.summary(),
      "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": nworld * nu * 4, "d2h_bytes_per_step": nworld * (nq + nv) * 4,
              "steps": Ke},
      "gpu_launches": int(gpu_launches),
      "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                   "kernel": f"mjb_group_kernel<{dom}>", "kernel_ms": group_ms[dom], "share_of_step": group_ms[dom] / kernel_ms,
                   "alg_bytes_per_launch": gbytes[dom] * nworld, "peak_source": peak_src,
                   "step": {"kernels_ms": kernel_ms, "alg_bytes_per_env_step": ALG_BYTES_PER_ENV_STEP, "achieved": step_achieved,
                            "frac": step_achieved / peak},
                   "groups": {g: {"ms": group_ms[g], "alg_bytes_per_world": gbytes[g],
                                  "frac": gbytes[g] * nworld / (group_ms[g] * 1e-3) / 1e9 / peak if group_ms[g] > 0 else None}
                              for g in gnames}},
    }  # fmt: skip
    if not args.no_cpu_baseline and world == 1:
      out["cpu_baseline"] = cpu_baseline(os.cpu_count() or 1)
    print(json.dumps(out))
  if dist is not None:
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
  main()
