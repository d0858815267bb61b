#!/usr/bin/env python
"""Entropy-model elementwise kernels against the HBM roofline.

Shapes of BASELINE.json configs[2] (C3, 256 x 1080p over the box; one GPU's micro-batch of 32
by default): z [B,17,30,192], y / sigma [B,68,120,192].  Each kernel is timed alone with CUDA
events; buffers are rotated through a pool larger than L2 so no launch re-reads what the
previous one left in cache.  bytes = ALGORITHMIC bytes per element:
  EntropyBottleneck.__call__ (inference)   4 z + 4 z_hat + 4 likelihood       = 12 B
  EntropyBottleneck symbols                4 z + 4 symbol                     =  8 B
  GaussianConditional.__call__             4 y + 4 sigma + 4 y_hat + 4 lik    = 16 B
  GaussianConditional symbols + indexes    4 y + 4 sigma + 4 symbol + 4 index = 16 B

  python tools/elementwise_bench.py [--batch 32] [--reps 20]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--batch", type=int, default=32)
  ap.add_argument("--reps", type=int, default=20)
  ap.add_argument("--eb-mult", type=int, default=16,
                  help="z is 1/16 of y: the EB kernels are timed on eb-mult x the z shape as well")
  args = ap.parse_args()
  from compression_b200.python.layers import entropy_models
  dev = torch.device("cuda", 0)
  peaks = {}
  try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
  except Exception:
    pass
  hbm = float(peaks.get("hbm_gbs", 6650.0))   # fallback: B200_PROFILING.md
  g = torch.Generator(device=dev).manual_seed(0)
  B = args.batch
  eb = entropy_models.EntropyBottleneck()
  eb.build(192, dev, torch.Generator().manual_seed(0))
  eb.update()
  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))

  def timeit(make_inputs, fn, count, bytes_per_elem, name, shape):
    # pool > 126 MB L2 so consecutive launches do not hit in cache
    per = count * bytes_per_elem
    pool = [make_inputs() for _ in range(max(2, int(400e6 // per) + 1))]
    for p in pool[:2]:
      fn(*p)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for r in range(args.reps):
      fn(*pool[r % len(pool)])
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / args.reps
    gbs = per / ms / 1e6
    print(json.dumps({"kernel": name, "shape": list(shape), "elements": count,
                      "bytes_per_element": bytes_per_elem, "ms": ms, "GB_s": gbs,
                      "hbm_peak_GB_s": hbm, "frac": gbs / hbm}), flush=True)

  for mult in (1, args.eb_mult):
    zs = (B * mult, 17, 30, 192)
    mk = lambda: (torch.randn(zs, generator=g, device=dev) * 4,)
    cnt = int(np.prod(zs))
    timeit(mk, lambda z: eb(z, training=False), cnt, 12, "eb_call (EntropyBottleneck.__call__)", zs)
    timeit(mk, lambda z: eb._quantize_symbols(z), cnt, 8, "eb_symbols", zs)
  ys = (B, 68, 120, 192)
  cnt = int(np.prod(ys))
  mk = lambda: (torch.randn(ys, generator=g, device=dev) * 3,
                torch.rand(ys, generator=g, device=dev) * 4 + 0.05)

  gc0 = entropy_models.GaussianConditional(None, table)
  gc0.build_tables(dev)

  def gc_call(y, s):
    gc = entropy_models.GaussianConditional(s, table)
    gc.adopt_tables(gc0)
    gc.built = True          # likelihood only: the index lookup is timed separately below
    return gc(y, training=False)

  def gc_sym(y, s):
    gc = entropy_models.GaussianConditional(s, table)
    gc.adopt_tables(gc0)
    return gc._symbols_and_indexes(y)

  timeit(mk, gc_call, cnt, 16, "gaussian_call (GaussianConditional.__call__)", ys)
  timeit(mk, gc_sym, cnt, 16, "gaussian_symbols_indexes", ys)


if __name__ == "__main__":
  main()
