"""Timeline of conv_thin_kernel's tensor-memory GDN epilogue (A0): clock64() stamps of CTA 0's phase
boundaries over tiles 8..15 (the stamps are compiled in by default: TFC_A0_STAMPS in csrc/conv_thin.cu).
Usage: python tools/a0_trace.py [--debug BITS]"""
import argparse
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from compression_b200 import _lib  # noqa: E402
from compression_b200.examples import bmshj2018  # noqa: E402

EPI = ["d1_full", None, "x2 arrived", "d2_full", "D2 loaded", "y done", "y_free", "staged"]


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--debug", type=int, default=0)
  args = ap.parse_args()
  if args.debug:
    _lib.check(_lib.lib().tfc_configure(99, args.debug))
  dev = torch.device("cuda", 0)
  model = bmshj2018.Bmshj2018Model(192, dev, seed=0)
  layer = model.analysis_transform.layers[0]
  x = torch.rand((64, 512, 768, 3), device=dev)
  for _ in range(3):
    layer(x, out_dtype=torch.bfloat16)
  torch.cuda.synchronize()
  buf = (ctypes.c_ulonglong * 512)()
  rc = _lib.lib().tfc_debug_a0_trace(buf, 512)
  assert rc == 0, rc
  t = [list(buf[i * 64:(i + 1) * 64]) for i in range(8)]
  t0 = min(v for v in t[0] if v)
  for i in range(8):
    print("tile %d" % (8 + i))
    for part in range(4):
      row = t[i][part * 12:part * 12 + 8]
      print("  part %d: " % part + "  ".join("%s %6d" % (n, v - t0) for n, v in zip(EPI, row) if n))
    print("  store thread: drained %6d  y_full %6d  store issued %6d" % tuple(v - t0 for v in t[i][60:63]))
    print("  mma: a2_full %6d  gemm2 issued %6d  d1_empty %6d  main issued %6d" %
          tuple(v - t0 for v in t[i][50:54]))


if __name__ == "__main__":
  main()
