#!/usr/bin/env python
"""Where the bls2017 training step's device time goes: torch.profiler over a few steps of
Bls2017Trainer.step at the C5 shape, kernels aggregated by name (CUDA time per step).

  python tools/train_profile.py [--batch 64] [--steps 3] [--top 30]
"""
import argparse
import collections
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--batch", type=int, default=64)
  ap.add_argument("--patch", type=int, default=256)
  ap.add_argument("--steps", type=int, default=3)
  ap.add_argument("--top", type=int, default=30)
  ap.add_argument("--model", default="bls2017", choices=["bls2017", "bmshj2018"])
  args = ap.parse_args()
  dev = torch.device("cuda", 0)
  if args.model == "bls2017":
    from compression_b200.examples import bls2017
    trainer = bls2017.Bls2017Trainer(bls2017.Bls2017Model(128, dev, seed=0), lmbda=0.01)
  else:
    from compression_b200.examples import bmshj2018
    trainer = bmshj2018.Bmshj2018Trainer(bmshj2018.Bmshj2018Model(192, dev, seed=0), lmbda=0.01)
  x = torch.rand((args.batch, args.patch, args.patch, 3), device=dev)
  for _ in range(3):
    trainer.step(x)
  torch.cuda.synchronize()
  from torch.profiler import profile, ProfilerActivity
  with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(args.steps):
      trainer.step(x)
    torch.cuda.synchronize()
  agg = collections.defaultdict(lambda: [0, 0.0])
  wgrad = []
  for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
      a = agg[ev.name[:90]]
      a[0] += 1
      a[1] += ev.device_time
      if "wgrad" in ev.name:
        wgrad.append(ev.device_time / 1e3)
  if wgrad:
    per = len(wgrad) // args.steps
    print("filter-gradient launches of the last step, in backward order (ms): " +
          " ".join("%.3f" % t for t in wgrad[-per:]))
  total = sum(v[1] for v in agg.values())
  print("device time per step: %.3f ms over %d kernel launches" % (
      total / args.steps / 1e3, sum(v[0] for v in agg.values()) // args.steps))
  for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:args.top]:
    print("%8.3f ms %5.1f%% %5d  %s" % (t / args.steps / 1e3, 100 * t / total, n // args.steps, name))


if __name__ == "__main__":
  main()
