#!/usr/bin/env python
"""Where does the multi-batch pipeline spend its time?  Host-side issue time per batch (python +
launch overhead, no synchronisation) against device time per batch, for a few depths."""
import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--steps", type=int, default=40)
  ap.add_argument("--depths", default="6,8,12")
  ap.add_argument("--batch", type=int, default=64)
  ap.add_argument("--mode", default="full", help="full | conv (transforms only) | coder (entropy "
                  "coding only, on the symbols of one analysed batch)")
  ap.add_argument("--tiles-per-cta", type=int, default=8)
  ap.add_argument("--pack", type=int, default=4)
  args = ap.parse_args()
  from compression_b200.examples import bmshj2018
  from compression_b200.python.ops import range_coding_ops as rc
  dev = torch.device("cuda", 0)
  model = bmshj2018.Bmshj2018Model(192, dev, seed=0)
  x = [torch.rand((args.batch, 512, 768, 3), device=dev) for _ in range(4)]
  batches = [x[i % 4] for i in range(args.steps)]
  from compression_b200.python.ops import tensor_ops
  y0, z0, sigma0 = model.analyse(x[0])
  c0 = model.compress(x[0])

  def work(xb):
    if args.mode == "full":
      model.decompress(model.compress(xb))
    elif args.mode == "conv":
      y, z, sigma = model.analyse(xb)
      model.hyper_synthesis_transform(z.to(torch.bfloat16))
      model.synthesis_transform(y.to(torch.bfloat16))
    else:
      model.entropy_bottleneck.compress_device(z0)
      gc = model._conditional(sigma0)
      gc.compress_device(y0)
      model.entropy_bottleneck.decompress_device(c0.side_string, tuple(c0.z_shape) + (192,),
                                                 channels=192, out_dtype=torch.bfloat16)
      gc.decompress_device(c0.string, out_dtype=torch.bfloat16)

  for depth in [int(d) for d in args.depths.split(",")]:
    with model.pipeline(depth, args.tiles_per_cta, args.pack) as pipe:
      for xb in batches[:2]:
        work(xb)
      torch.cuda.synchronize()
      t0 = time.perf_counter()
      cur = pipe._fork()
      with rc.deferred_checks():
        for k, xb in enumerate(batches):
          with torch.cuda.stream(pipe.streams[k % depth]):
            work(xb)
        t1 = time.perf_counter()
        pipe._join(cur)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
    n = len(batches)
    print("depth %2d: host issue %.2f ms/batch, total %.2f ms/batch (%.0f MPix/s)" % (
        depth, (t1 - t0) / n * 1e3, (t2 - t0) / n * 1e3,
        n * args.batch * 512 * 768 / (t2 - t0) / 1e6), flush=True)


if __name__ == "__main__":
  main()
