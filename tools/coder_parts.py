#!/usr/bin/env python
"""Times the four entropy-coder launches of one bmshj2018 batch (z encode/decode through the
EntropyBottleneck tables, y encode/decode through the Gaussian tables) on their own."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--batch", type=int, default=64)
  ap.add_argument("--height", type=int, default=512)
  ap.add_argument("--width", type=int, default=768)
  ap.add_argument("--pack", type=int, default=0)
  ap.add_argument("--reps", type=int, default=5)
  ap.add_argument("--lanes", type=int, default=0, help="0 auto, 1 warps per stream, 2 lane per stream")
  ap.add_argument("--smem-kb", type=int, default=0)
  ap.add_argument("--table-kb", type=int, default=0)
  ap.add_argument("--debug", type=int, default=0, help="library experiment bits (64: packed encoder with one stream per chain warp)")
  ap.add_argument("--repeat", type=int, default=1,
                  help="code the batch's latents this many times over in ONE launch (batch * repeat streams)")
  args = ap.parse_args()
  import compression_b200 as tfc
  from compression_b200.examples import bmshj2018
  dev = torch.device("cuda", 0)
  model = bmshj2018.Bmshj2018Model(192, dev, seed=0)
  x = torch.rand((args.batch, args.height, args.width, 3), device=dev)
  y, z, sigma = model.analyse(x)
  if args.repeat > 1:
    y, z, sigma = (t.repeat(args.repeat, 1, 1, 1) for t in (y, z, sigma))
  string, side_string = model.encode_strings(y, z, sigma)
  c = bmshj2018.Compressed(string, side_string, tuple(x.shape[1:-1]), tuple(y.shape[1:-1]),
                           tuple(z.shape[1:-1]))
  tfc.configure(coder_lanes=args.lanes)
  if args.debug:
    from compression_b200 import _lib
    _lib.check(_lib.lib().tfc_configure(99, args.debug))
  tfc.configure(coder_streams_per_cta=args.pack, coder_smem_kb=args.smem_kb,
                coder_table_kb=args.table_kb)
  eb, gc = model.entropy_bottleneck, model._conditional(sigma)
  nf = 192

  def timed(name, fn, symbols):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.reps):
      fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / args.reps
    print("%-8s %8.3f ms  %6.1f cycles/symbol/stream (1.965 GHz)" % (
        name, ms, ms * 1e-3 * 1.965e9 / symbols), flush=True)

  nz, ny = z[0].numel(), y[0].numel()
  timed("z encode", lambda: eb.compress_device(z), nz)
  timed("z decode", lambda: eb.decompress_device(c.side_string, tuple(c.z_shape) + (nf,), channels=nf,
                                                 out_dtype=torch.bfloat16), nz)
  timed("y encode", lambda: gc.compress_device(y), ny)
  timed("y decode", lambda: gc.decompress_device(c.string, out_dtype=torch.bfloat16), ny)
  tfc.configure(coder_streams_per_cta=0)


if __name__ == "__main__":
  main()
