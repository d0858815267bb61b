#!/usr/bin/env python
"""Prints the hot instructions (by warp-stall samples) of each kernel in an ncu source-page CSV:
  ncu -i X.ncu-rep --page source --csv > src.csv ; python tools/ncu_hot.py src.csv [min_pct]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
min_pct = float(sys.argv[2]) if len(sys.argv) > 2 else 1.2
kern, cur = [], None
for r in rows:
  if r and r[0] == "Kernel Name":
    cur = {"name": r[1], "rows": []}
    kern.append(cur)
  elif cur is not None:
    cur["rows"].append(r)
for k in kern:
  hdr, data = k["rows"][0], k["rows"][1:]
  si, so = hdr.index("# Samples"), hdr.index("Source")
  tot = sum(int(r[si]) for r in data if r[si].isdigit())
  print(k["name"][:70], "samples", tot, "instructions", len(data))
  for idx, r in enumerate(data):
    n = int(r[si]) if r[si].isdigit() else 0
    if n > tot * min_pct / 100:
      reasons = [(hdr[i], int(r[i])) for i in range(len(hdr))
                 if hdr[i].startswith("stall") and r[i].isdigit() and int(r[i]) > 0]
      reasons.sort(key=lambda x: -x[1])
      print("%5d %5.1f%% %-56s %s" % (idx, 100.0 * n / tot, r[so].strip()[:56], reasons[:3]))
