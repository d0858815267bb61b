#!/usr/bin/env python
"""profiles/rNN_conv_traffic.json (read by bench.py for roofline.traffic) from the raw CSV of
  ncu --set full --clock-control none -k regex:conv_ -c 28 python tools/conv_bench.py --batch 64 --reps 1
(two launches per conv layer in network order; the second one of each is counted, the three
hyper-synthesis layers twice — a bench step runs them in compress and in decompress).

  python tools/conv_traffic.py X_raw.csv "build description" > profiles/rNN_conv_traffic.json"""
import csv
import json
import sys

LAYERS = ["A0", "A1", "A2", "A3", "HA0", "HA1", "HA2", "HS0", "HS1", "HS2", "S0", "S1", "S2", "S3"]
TENSOR = "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"


def num(v, unit):
  f = float(v.replace(",", ""))
  scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6,
           "s": 1e3}
  return f * scale[unit]


def main():
  rows = list(csv.reader(open(sys.argv[1])))
  hdr, units, data = rows[0], rows[1], rows[2:]
  col = {n: hdr.index(n) for n in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum",
                                   "dram__bytes_write.sum", TENSOR)}
  assert len(data) == 2 * len(LAYERS), len(data)
  per, total, ms_total = {}, 0.0, 0.0
  for i, name in enumerate(LAYERS):
    r = data[2 * i + 1]
    ms = num(r[col["gpu__time_duration.sum"]], units[col["gpu__time_duration.sum"]])
    b = (num(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]]) +
         num(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]]))
    kern = r[col["Kernel Name"]].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    per[name] = {"ms": round(ms, 5), "dram_bytes": round(b), "tensor_pipe_pct": r[col[TENSOR]],
                 "kernel": kern}
    times = 2 if name.startswith("HS") else 1
    total += times * b
    ms_total += times * ms
  print(json.dumps({
      "workload": "c2",
      "build": sys.argv[2] if len(sys.argv) > 2 else "",
      "command": "ncu --set full --clock-control none -k regex:conv_ -c 28 python tools/conv_bench.py --batch 64 --reps 1",
      "what": "dram__bytes_read.sum + dram__bytes_write.sum of the second launch of each of the 14 conv layers "
              "(batch 64, 768x512), hyper-synthesis counted twice: the 17 conv launches of one bench step",
      "dram_bytes_per_step": total,
      "ncu_ms_per_step_serialised": round(ms_total, 5),
      "per_layer": per,
      "algorithmic_bytes_per_step_note": "layer inputs + outputs at their dtypes ~ 13.6 GB",
  }, indent=1))


if __name__ == "__main__":
  main()
