"""Top source lines of one file by stall samples from an `ncu --page source --csv --print-source cuda,sass` export.
usage: ncu_by_line.py export.csv file-substring [ntop]"""
import csv, os, sys

rows = list(csv.reader(open(sys.argv[1])))
want, ntop = sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 25
cur, hdr, out = None, None, []
for r in rows:
  if r and r[0] == "File Path":
    cur, hdr = r[1], None
  elif r and r[0] == "Line No":
    hdr = {h: i for i, h in enumerate(r)}
  elif cur and hdr and want in cur and len(r) >= len(hdr) and r[2] == "-":
    try:
      out.append((int(r[hdr["# Samples"]] or 0), int(r[hdr["Instructions Executed"]] or 0), int(r[0]), cur))
    except ValueError:
      pass
src = {}
for s, n, ln, f in sorted(out, reverse=True)[:ntop]:
  if f not in src:
    try:
      src[f] = open(f).read().split("\n")
    except OSError:
      src[f] = []
  text = src[f][ln - 1].strip()[:110] if ln - 1 < len(src[f]) else ""
  print(f"{s:7d} {n:10d}  {os.path.basename(f)}:{ln}  {text}")
