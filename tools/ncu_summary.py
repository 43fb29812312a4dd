"""Builds the committed profile summaries under profiles/ from ncu exports kept in gpurun_out/.

  python tools/ncu_summary.py launches gpurun_out/r1_launches.csv profiles/r1_launches_summary.md
  python tools/ncu_summary.py full gpurun_out/prof_b4_raw.csv profiles/r1_ncu_full_groups
"""
import collections
import csv
import json
import re
import sys

GROUPS = ["smooth_position", "collision", "make_constraint", "smooth_velocity", "solve", "euler"]


def kname(name):
  m = re.search(r"mjb_group_kernel<\(?(?:int\))?(\d)", name)
  if m:
    return "mjb_group_kernel<%s>" % GROUPS[int(m.group(1))]
  m = re.search(r"(mjb_\w+|\w+)", name)
  return m.group(1) if m else name


def launches(src, dst):
  rows = [r for r in csv.reader(l for l in open(src) if not l.startswith("==")) if r]
  hdr = rows[0]
  ci = {h: i for i, h in enumerate(hdr)}
  agg = collections.OrderedDict()
  for r in rows[1:]:
    if len(r) < len(hdr) or r[ci["Metric Name"]] != "gpu__time_duration.sum":
      continue
    v = float(r[ci["Metric Value"]].replace(",", ""))
    unit = r[ci["Metric Unit"]]
    v_us = v / 1e3 if unit in ("nsecond", "ns") else (v if unit in ("usecond", "us") else v * 1e3)
    k = kname(r[ci["Kernel Name"]])
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += v_us
  tot = sum(a[1] for a in agg.values())
  with open(dst, "w") as f:
    f.write("# ncu launch list (gpu__time_duration.sum, --clock-control none, first 400 launches of bench.py)\n\n")
    f.write("Per-launch times under ncu are cold-cache and serialised: the SHARE of each kernel is the figure to compare with\nbench.py's `roofline.groups`, not the absolute time.\n\n")
    f.write("| kernel | launches | total us | avg us | share |\n|---|---:|---:|---:|---:|\n")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
      f.write(f"| `{k}` | {n} | {t:.1f} | {t / n:.1f} | {100 * t / tot:.1f}% |\n")
  print(open(dst).read())


FULL_METRICS = [
  "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
  "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
  "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
  "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
  "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active",
  "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
  "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
  "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
  "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
  "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]  # fmt: skip


def full(src, dst):
  rows = list(csv.reader(open(src)))
  hdr, units, data = rows[0], rows[1], rows[2:]
  ci = {h: i for i, h in enumerate(hdr)}
  out = {}
  for r in data:
    k = kname(r[ci["Kernel Name"]])
    e = {}
    for mname in FULL_METRICS:
      if mname in ci:
        try:
          e[mname] = float(r[ci[mname]].replace(",", ""))
        except ValueError:
          e[mname] = r[ci[mname]]
        e[mname + ".unit"] = units[ci[mname]]
    # dram bytes per launch in bytes
    tb = 0.0
    for mname in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
      u = units[ci[mname]]
      scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
      tb += float(r[ci[mname]].replace(",", "")) * scale
    e["dram_bytes_per_launch"] = tb
    out[k] = e
  json.dump(out, open(dst + ".json", "w"), indent=1, sort_keys=True)
  with open(dst + ".md", "w") as f:
    f.write("# ncu --set full, one launch of every stage-group kernel (steady state, humanoid nworld=8192)\n\n")
    ks = list(out)
    f.write("| metric | " + " | ".join(k.replace("mjb_group_kernel", "") for k in ks) + " |\n|---|" + "---:|" * len(ks) + "\n")
    for mname in FULL_METRICS + ["dram_bytes_per_launch"]:
      if mname not in out[ks[0]]:
        continue
      f.write(f"| {mname} ({out[ks[0]].get(mname + '.unit', 'byte')}) | " + " | ".join(f"{out[k][mname]:.4g}" if isinstance(out[k][mname], float) else str(out[k][mname]) for k in ks) + " |\n")
  print(open(dst + ".md").read())


if __name__ == "__main__":
  {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
