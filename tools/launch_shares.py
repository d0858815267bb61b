#!/usr/bin/env python
"""Aggregates an ncu launch list (`--metrics gpu__time_duration.sum --csv --log-file X.csv`) per
kernel: launches, total time, share.   python tools/launch_shares.py X.csv > profiles/X_shares.md"""
import collections
import csv
import re
import sys


def main():
  rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
  hdr = rows[0]
  kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
  tot = collections.OrderedDict()
  for r in rows[1:]:
    if len(r) <= mv or r[hdr.index("Metric Name")] != "gpu__time_duration.sum":
      continue
    name = re.sub(r"\(.*$", "", r[kn])
    name = re.sub(r"^void ", "", name).replace("(anonymous namespace)::", "")[:60]
    v = float(r[mv].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[mu], 1e-6)
    n, t = tot.get(name, (0, 0.0))
    tot[name] = (n + 1, t + v)
  total = sum(t for _, t in tot.values())
  print("| kernel | launches | total ms | share |\n|---|---|---|---|")
  for name, (n, t) in sorted(tot.items(), key=lambda kv: -kv[1][1])[:24]:
    print("| %s | %d | %.3f | %.1f%% |" % (name, n, t, 100 * t / total))
  print("\ntotal %.1f ms over %d launches" % (total, sum(n for n, _ in tot.values())))


if __name__ == "__main__":
  main()
