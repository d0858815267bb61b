#!/usr/bin/env python
"""bls2017 training-step timing (BASELINE.json configs[4]): 256x256 synthetic crops, batch 64
per GPU, 128 filters, NCCL gradient all-reduce when launched under torchrun.

  python tools/train_bench.py [--batch 64] [--steps 10] [--warmup 3]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/train_bench.py
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--batch", type=int, default=64)
  ap.add_argument("--patch", type=int, default=256)
  ap.add_argument("--steps", type=int, default=10)
  ap.add_argument("--warmup", type=int, default=3)
  ap.add_argument("--no-graph", action="store_true",
                  help="issue the ~600 launches of a step one by one instead of replaying the "
                       "captured CUDA graph of its forward + backward")
  args = ap.parse_args()
  from compression_b200.examples import bls2017
  world = int(os.environ.get("WORLD_SIZE", "1"))
  rank = int(os.environ.get("RANK", "0"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  torch.cuda.set_device(local)
  dev = torch.device("cuda", local)
  if world > 1:
    dist.init_process_group("nccl", device_id=dev)
  model = bls2017.Bls2017Model(128, dev, seed=0)          # same seed: identical replicas
  trainer = bls2017.Bls2017Trainer(model, lmbda=0.01)
  g = torch.Generator(device=dev).manual_seed(100 + rank)  # per-rank crops and noise
  torch.manual_seed(200 + rank)
  x = torch.rand((args.batch, args.patch, args.patch, 3), generator=g, device=dev)
  for _ in range(args.warmup):
    out = trainer.step(x, graph=not args.no_graph)
  if world > 1:
    dist.barrier()
  torch.cuda.synchronize()
  a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  a.record()
  for _ in range(args.steps):
    out = trainer.step(x, graph=not args.no_graph)
  b.record()
  torch.cuda.synchronize()
  ms = torch.tensor([a.elapsed_time(b) / args.steps], device=dev)
  check = trainer.main.param.double().sum().reshape(1)
  if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    lo, hi = check.clone(), check.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    assert float(hi - lo) == 0.0, "replicas diverged"
  if rank == 0:
    flop = 8.83e9 * args.batch * (args.patch / 256.0) ** 2      # fwd + bwd per image
    print(json.dumps({
        "metric": "bls2017 training step", "n_gpus": world, "batch_per_gpu": args.batch,
        "ms_per_step": float(ms), "cuda_graph": not args.no_graph, "images_per_s": world * args.batch / float(ms) * 1e3,
        "tflops_per_gpu": flop / float(ms) / 1e9, "parameters": trainer.num_parameters,
        "loss": out["loss"], "bpp": out["bpp"], "mse": out["mse"], "aux": out["aux"],
        "replicas_identical": True}))
  if world > 1:
    dist.destroy_process_group()


if __name__ == "__main__":
  main()
