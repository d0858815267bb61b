#!/usr/bin/env python
"""Timing experiments on the first analysis layer (conv_thin.cu): sweeps the kernel's debug bits
(1: builders idle, 2: no output stores) in ONE process, 20 launches each after 3 warm-ups.
Results with a non-zero debug value are timing-only (outputs are wrong by construction)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  from compression_b200 import _lib
  from compression_b200.examples import bmshj2018
  dev = torch.device("cuda", 0)
  model = bmshj2018.Bmshj2018Model(192, dev, seed=0)
  layer = model.analysis_transform.layers[0]
  x = torch.rand((64, 512, 768, 3), device=dev)
  for tpc in (0, 8, 32):
    _lib.configure(tiles_per_cta=tpc)
    for dbg in (0, 1, 2, 3, 0):
      _lib.check(_lib.lib().tfc_configure(99, dbg))
      for _ in range(3):
        layer(x, out_dtype=torch.bfloat16)
      torch.cuda.synchronize()
      a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      a.record()
      for _ in range(20):
        layer(x, out_dtype=torch.bfloat16)
      b.record()
      torch.cuda.synchronize()
      print("tiles_per_cta %2d debug %d: %.3f ms" % (tpc, dbg, a.elapsed_time(b) / 20), flush=True)
  _lib.check(_lib.lib().tfc_configure(99, 0))


if __name__ == "__main__":
  main()
