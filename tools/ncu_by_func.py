"""Aggregates an `ncu --page source --csv --print-source cuda,sass` export per source file:function
(static SASS count, executed warp instructions, stall samples).  usage: ncu_by_func.py export.csv"""
import collections
import csv
import os
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
sections, cur = [], None
for r in rows:
  if r and r[0] == "File Path":
    cur = {"file": r[1], "rows": []}
    sections.append(cur)
  elif cur is not None:
    cur["rows"].append(r)


def func_ranges(fn):
  """line number -> enclosing function name (scans for top-level __device__/__global__ definitions)"""
  out, name = {}, "?"
  try:
    lines = open(fn).read().split("\n")
  except OSError:
    return out
  for i, l in enumerate(lines, 1):
    m = re.match(r"^(?:static\s+)?__(?:device|global|host)__.*?\b(\w+)\s*\(", l)
    if m:
      name = m.group(1)
    out[i] = name
  return out


agg = collections.defaultdict(lambda: [0, 0, 0])
for sec in sections:
  hdr = next((r for r in sec["rows"] if r and r[0] == "Line No"), None)
  if hdr is None:
    continue
  ci = {h: i for i, h in enumerate(hdr)}
  fr = func_ranges(sec["file"])
  base = os.path.basename(sec["file"])
  line = None
  for r in sec["rows"]:
    if not r or r[0] in ("Line No", "Function Name") or len(r) < len(hdr):
      continue
    if r[2] == "-":  # per-source-line aggregate
      line = int(r[0])
      key = (base, fr.get(line, "?"))
      try:
        agg[key][1] += int(r[ci["Instructions Executed"]] or 0)
        agg[key][2] += int(r[ci["# Samples"]] or 0)
      except ValueError:
        pass
    elif line is not None:  # a SASS row of the last source line
      agg[(base, fr.get(line, "?"))][0] += 1
tot = [sum(v[i] for v in agg.values()) for i in range(3)]
print(f"{'function':48s} {'sass':>7s} {'inst_exec':>12s} {'%':>6s} {'samples':>8s} {'%':>6s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][2]):
  print(f"{k[0] + ':' + k[1]:48s} {v[0]:7d} {v[1]:12d} {100 * v[1] / max(tot[1], 1):6.1f} {v[2]:8d} {100 * v[2] / max(tot[2], 1):6.1f}")
print("total", tot)
