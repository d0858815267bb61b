#!/usr/bin/env python
"""Pinned host<->device copy bandwidth of the box (what bounds bench.py's e2e leg from below):
one direction at a time and both at once on two streams.

  python tools/pcie_bw.py [--mb 256] [--reps 10]
"""
import argparse
import time

import torch


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--mb", type=int, default=256)
  ap.add_argument("--reps", type=int, default=10)
  args = ap.parse_args()
  dev = torch.device("cuda", 0)
  n = args.mb << 20
  h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
  h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
  d_in = torch.empty(n, dtype=torch.uint8, device=dev)
  d_out = torch.empty(n, dtype=torch.uint8, device=dev)
  s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

  def timed(fn):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.reps):
      fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / args.reps

  def h2d():
    with torch.cuda.stream(s1):
      d_in.copy_(h_in, non_blocking=True)

  def d2h():
    with torch.cuda.stream(s2):
      h_out.copy_(d_out, non_blocking=True)

  def both():
    h2d()
    d2h()

  gb = n / 1e9
  print("pinned copies of %d MiB: H2D %.1f GB/s, D2H %.1f GB/s, both at once %.1f GB/s each way" % (
      args.mb, gb / timed(h2d), gb / timed(d2h), gb / timed(both)))


if __name__ == "__main__":
  main()
