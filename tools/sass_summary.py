#!/usr/bin/env python
"""Per-kernel count of the SASS mnemonics that prove (or disprove) a Blackwell-native kernel, from
`cuobjdump -sass` of the built library (no GPU needed):
  UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA tile load/store,
  UBLKCP = bulk copy, HMMA = mma.sync (legacy tensor path), SYNCS = mbarrier ops.

  python tools/sass_summary.py > profiles/rN_sass_summary.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "compression_b200", "libtfc_b200.so")
WANT = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "HMMA", "SYNCS",
        "REDUX", "MUFU"]


def main():
  out = subprocess.run(["cuobjdump", "-sass", SO], check=True, capture_output=True,
                       text=True).stdout
  demangle = {}
  counts = collections.OrderedDict()
  cur = None
  for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
      cur = m.group(1)
      counts[cur] = collections.Counter()
      continue
    if cur is None:
      continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
      op = m.group(1).split(".")[0]
      counts[cur]["_total"] += 1
      for w in WANT:
        if op.startswith(w):
          counts[cur][w] += 1
  names = list(counts)
  dem = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
  for n, d in zip(names, dem):
    d = re.sub(r"\(anonymous namespace\)::", "", d)
    d = re.sub(r"\(.*", "", d).replace("void ", "")
    demangle[n] = d
  print("# SASS instruction summary of compression_b200/libtfc_b200.so (sm_100a)\n")
  print("`cuobjdump -sass` of the in-tree build, instructions per kernel.  UTCHMMA = tcgen05.mma "
        "kind::f16, LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA tile load / store "
        "(cp.async.bulk.tensor), HMMA = mma.sync (legacy tensor path), SYNCS = mbarrier.\n")
  print("| kernel | instructions | " + " | ".join(WANT) + " |")
  print("|---|---|" + "---|" * len(WANT))
  for n in sorted(names, key=lambda k: demangle[k]):
    c = counts[n]
    print("| %s | %d | %s |" % (demangle[n][:70], c["_total"],
                                 " | ".join(str(c[w]) if c[w] else "" for w in WANT)))


if __name__ == "__main__":
  main()
