#!/usr/bin/env python
"""Per-layer SignalConv2D timing at the bmshj2018 C2 shapes (batch B, 768x512):
CUDA-event ms, algorithmic TFLOP/s (SURVEY §8a counting) and output GB/s per layer.

  python tools/conv_bench.py [--batch 64] [--layers A0,A1,S2,S3] [--reps 5] [--check]
"""
import argparse
import json
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
  ap.add_argument("--layers", default="")
  ap.add_argument("--reps", type=int, default=5)
  ap.add_argument("--check", action="store_true", help="compare with the CUDA-core kernel")
  ap.add_argument("--debug", type=int, default=0, help="kernel experiment bits (timing only)")
  ap.add_argument("--tiles-per-cta", type=int, default=0)
  ap.add_argument("--cluster", type=int, default=0, help="0 auto, 1 off, 2, 4")
  ap.add_argument("--stages", type=int, default=0, help="operand-ring stages (0: 4)")
  ap.add_argument("--pair", type=int, default=0, help="cta_group::2 MMAs: 0 auto, 1 off, 2 on")
  args = ap.parse_args()
  from compression_b200.examples import bmshj2018
  from compression_b200.python.layers import gdn as gdn_lib
  from compression_b200 import _lib
  if args.debug:
    _lib.check(_lib.lib().tfc_configure(99, args.debug))
  if args.cluster:
    _lib.configure(conv_cluster=args.cluster)
  if args.stages:
    _lib.configure(conv_stages=args.stages)
  if args.pair:
    _lib.configure(conv_pair=args.pair)
  if args.tiles_per_cta:
    _lib.configure(tiles_per_cta=args.tiles_per_cta)
  dev = torch.device("cuda", 0)
  model = bmshj2018.Bmshj2018Model(192, dev, seed=0)
  B, H, W = args.batch, args.height, args.width
  want = set(filter(None, args.layers.split(",")))
  g = torch.Generator().manual_seed(3)
  plan = [("A", model.analysis_transform, (H, W, 3)),
          ("HA", model.hyper_analysis_transform, (H // 16, W // 16, 192)),
          ("HS", model.hyper_synthesis_transform, (H // 64, W // 64, 192)),
          ("S", model.synthesis_transform, (H // 16, W // 16, 192))]
  for prefix, seq, (h, w, c) in plan:
    for i, layer in enumerate(seq.layers):
      name = "%s%d" % (prefix, i)
      shape = (B, h, w, c)
      _, ho, wo, co = layer.compute_output_shape(shape)
      if not want or name in want:
        x = torch.randn(shape, generator=g).to(dev)
        if c != 3:
          x = x.to(torch.bfloat16)
        out_dtype = torch.float32 if i == len(seq.layers) - 1 else torch.bfloat16
        y = layer(x, out_dtype=out_dtype)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.reps + 1)]
        a.record()
        evs[0].record()
        for i_rep in range(args.reps):
          y = layer(x, out_dtype=out_dtype)
          evs[i_rep + 1].record()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / args.reps
        each = sorted(evs[k].elapsed_time(evs[k + 1]) for k in range(args.reps))
        kk = layer.kernel_support[0] * layer.kernel_support[1]
        if layer.strides_up[0] > 1:
          fl = 2.0 * h * w * c * co * kk
        else:
          fl = 2.0 * ho * wo * c * co * kk
        if isinstance(layer.activation, gdn_lib.GDN):
          fl += 2.0 * ho * wo * co * co
        fl *= B
        rec = {"layer": name, "in": list(shape), "out": [B, ho, wo, co], "ms": round(ms, 4),
               "ms_min": round(each[0], 4), "ms_med": round(each[len(each) // 2], 4),
               "ms_max": round(each[-1], 4),
               "tflops": round(fl / ms / 1e9, 1),
               "io_gbs": round((x.numel() * x.element_size() + y.numel() * y.element_size())
                               / ms / 1e6, 1)}
        if args.check:
          saved = layer.impl
          layer.impl = 1
          ref = layer(x, out_dtype=torch.float32).float()
          layer.impl = saved
          err = (y.float() - ref).abs().max().item()
          rec["max_abs_vs_cuda_core"] = err
          rec["ref_max"] = ref.abs().max().item()
        print(json.dumps(rec), flush=True)
      h, w, c = ho, wo, co


if __name__ == "__main__":
  main()
