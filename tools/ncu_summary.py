#!/usr/bin/env python
"""Per-kernel summary of an `ncu --page raw --csv` export (one row per profiled launch):
duration, DRAM bytes, DRAM / L2 / tensor-pipe utilisation, registers, occupancy.

  ncu -i X.ncu-rep --page raw --csv > X_raw.csv ; python tools/ncu_summary.py X_raw.csv > profiles/X.md
"""
import csv
import sys

WANT = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_read"),
    ("dram__bytes_write.sum", "dram_write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_pct"),
    ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
     "tensor_pipe_pct"),
    ("sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed",
     "bf16_mma_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "dyn_smem"),
]


def main():
  rows = list(csv.reader(open(sys.argv[1])))
  hdr, units, data = rows[0], rows[1], rows[2:]
  idx = {name: hdr.index(name) for name, _ in WANT if name in hdr}
  ki = hdr.index("Kernel Name")
  cols = [(name, short) for name, short in WANT if name in idx]
  print("| # | kernel | " + " | ".join("%s [%s]" % (s, units[idx[n]] or "-") for n, s in cols) + " |")
  print("|---|---|" + "---|" * len(cols))
  for k, r in enumerate(data):
    kern = r[ki].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    vals = []
    for n, _ in cols:
      v = r[idx[n]]
      try:
        f = float(v.replace(",", ""))
        vals.append("%.4g" % f)
      except ValueError:
        vals.append(v)
    print("| %d | %s | %s |" % (k, kern[:48], " | ".join(vals)))


if __name__ == "__main__":
  main()
