"""Hot-code footprint from an `ncu --page source --csv --print-source cuda,sass` export: number of
SASS instructions that executed at all, and per-function breakdown.  usage: ncu_hot_code.py export.csv"""
import collections, csv, os, re, sys
rows = list(csv.reader(open(sys.argv[1])))
sections, cur = [], None
for r in rows:
  if r and r[0] == "File Path":
    cur = {"file": r[1], "rows": []}; sections.append(cur)
  elif cur is not None:
    cur["rows"].append(r)
def func_ranges(fn):
  out, name = {}, "?"
  try: lines = open(fn).read().split("\n")
  except OSError: return out
  for i, l in enumerate(lines, 1):
    m = re.match(r"^(?:static\s+)?__(?:device|global|host)__.*?\b(\w+)\s*\(", l)
    if m: name = m.group(1)
    out[i] = name
  return out
seen = {}
per = collections.defaultdict(lambda: [0, 0])
for sec in sections:
  hdr = next((r for r in sec["rows"] if r and r[0] == "Line No"), None)
  if hdr is None: continue
  ci = {h: i for i, h in enumerate(hdr)}
  fr = func_ranges(sec["file"]); base = os.path.basename(sec["file"]); line = None
  for r in sec["rows"]:
    if not r or r[0] in ("Line No", "Function Name") or len(r) < len(hdr): continue
    if r[2] == "-": line = int(r[0]); continue
    if line is None: continue
    addr = r[2]
    try: ex = int(r[ci["Instructions Executed"]] or 0)
    except ValueError: ex = 0
    if addr in seen: continue
    seen[addr] = ex
    k = base + ":" + fr.get(line, "?")
    per[k][0] += 1
    if ex > 0: per[k][1] += 1
tot = len(seen); hot = sum(1 for v in seen.values() if v > 0)
print(f"SASS instructions {tot} ({tot * 16 / 1024:.1f} kB), executed at least once {hot} ({hot * 16 / 1024:.1f} kB)")
for k, v in sorted(per.items(), key=lambda kv: -kv[1][1])[:25]:
  print(f"{k:48s} static {v[0]:6d} hot {v[1]:6d} ({v[1] * 16 / 1024:.1f} kB)")
