#!/usr/bin/env python
"""Per-phase timeline of the multi-batch pipeline (no nsys in this image): a CUDA event at the end
of every phase of every batch on the batch's own stream.  Phase time in the pipeline = end(phase)
- end(previous phase on that stream) = waiting for SMs + running; compared with the same phase
timed alone on the GPU, it shows which phases queue and how many SM-milliseconds the coder holds.

  python tools/pipe_gantt.py [--depth 20] [--pack 4] [--steps 60]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

PHASES = ["analysis+hyper", "z encode", "y encode", "z decode", "hyper synthesis", "y decode",
          "synthesis"]


def one_batch(model, x, mark):
  nf = model.num_filters
  y, z, sigma = model.analyse(x)
  mark(0)
  side = model.entropy_bottleneck.compress_device(z)
  mark(1)
  string = model._conditional(sigma).compress_device(y)
  mark(2)
  z_hat = model.entropy_bottleneck.decompress_device(side, tuple(z.shape[1:3]) + (nf,), nf,
                                                     torch.bfloat16)
  mark(3)
  sigma2 = model.hyper_synthesis_transform(z_hat)[:, :y.shape[1], :y.shape[2], :].contiguous()
  mark(4)
  y_hat = model._conditional(sigma2).decompress_device(string, torch.bfloat16)
  mark(5)
  model.synthesis_transform(y_hat)
  mark(6)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--depth", type=int, default=20)
  ap.add_argument("--pack", type=int, default=4)
  ap.add_argument("--tiles-per-cta", type=int, default=8)
  ap.add_argument("--steps", type=int, default=60)
  ap.add_argument("--batch", type=int, default=64)
  args = ap.parse_args()
  from compression_b200.examples import bmshj2018
  from compression_b200.python.ops import range_coding_ops as rc
  dev = torch.device("cuda", 0)
  model = bmshj2018.Bmshj2018Model(192, dev, seed=0)
  x = torch.rand((args.batch, 512, 768, 3), device=dev)

  def run(depth, steps, pack, tiles):
    ev = {}
    with model.pipeline(depth, tiles, pack) as pipe, rc.deferred_checks() as checks:
      for k in range(depth + 1):                      # warm every stream
        with torch.cuda.stream(pipe.streams[k % depth]):
          one_batch(model, x, lambda p: None)
      torch.cuda.synchronize()
      checks.flush()
      t0 = torch.cuda.Event(enable_timing=True)
      cur = pipe._fork()
      t0.record()
      for k in range(steps):
        st = pipe.streams[k % depth]
        checks.tag = k
        with torch.cuda.stream(st):
          start = torch.cuda.Event(enable_timing=True)
          start.record()
          ev[(k, -1)] = start

          def mark(p, k=k, st=st):
            e = torch.cuda.Event(enable_timing=True)
            e.record(st)
            ev[(k, p)] = e
          one_batch(model, x, mark)
        checks.drain(2 * depth)
      pipe._join(cur)
      torch.cuda.synchronize()
    t = {key: t0.elapsed_time(e) for key, e in ev.items()}
    total = max(t.values())
    # steady state: skip the first and last `depth` batches
    ks = [k for k in range(steps) if depth <= k < steps - depth] or list(range(steps))
    dur = {p: float(np.mean([t[(k, p)] - t[(k, p - 1)] for k in ks])) for p in range(7)}
    # the first phase of a batch also waits for the stream's previous batch: measure from the later
    # of (own issue, previous batch's end on this stream)
    first = []
    for k in ks:
      prev_end = t.get((k - depth, 6), 0.0)
      first.append(t[(k, 0)] - max(prev_end, t[(k, -1)]))
    dur[0] = float(np.mean(first))
    return total / steps, dur

  import ctypes
  from compression_b200 import _lib
  trace = (ctypes.c_uint64 * 4)()

  def coder_trace():
    _lib.check(_lib.lib().tfc_debug_coder_trace(trace, 1))
    v = [int(t) for t in trace]
    # every image has a z stream and a y stream: time per IMAGE (z + y)
    return {"encode_ms_per_image": v[0] / max(v[1] / 2, 1) / 1e6, "encode_streams": v[1],
            "decode_ms_per_image": v[2] / max(v[3] / 2, 1) / 1e6, "decode_streams": v[3]}

  alone_ms, alone = run(1, 6, 0, 0)
  trace_alone = coder_trace()
  ms, dur = run(args.depth, args.steps, args.pack, args.tiles_per_cta)
  trace_pipe = coder_trace()
  out = {"depth": args.depth, "pack": args.pack, "tiles_per_cta": args.tiles_per_cta,
         "ms_per_batch": ms, "ms_per_batch_serial": alone_ms,
         "phases": {PHASES[p]: {"alone_ms": round(alone[p], 3), "in_pipeline_ms": round(dur[p], 3)}
                    for p in range(7)},
         "latency_per_batch_ms": round(sum(dur.values()), 2),
         # y and z streams together (z streams are 1/16 as long): time a coder stream really held
         # its CTA slot, alone and in the pipeline
         "coder_cta_time": {"alone": trace_alone, "in_pipeline": trace_pipe}}
  print(json.dumps(out))


if __name__ == "__main__":
  main()
