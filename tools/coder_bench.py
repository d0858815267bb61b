#!/usr/bin/env python
"""Range-coder throughput sweep (BASELINE.json configs[3]): S streams x n symbols,
precision 16, overflow_width 4, the 64-row Gaussian table of bmshj2018; bit-exact check of a
few streams against the CPU checker, then CUDA-event timing of encode and decode.

  python tools/coder_bench.py [--streams 64] [--n 294912] [--reps 5] [--check 2]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--streams", type=int, default=64)
  ap.add_argument("--n", type=int, default=294912)
  ap.add_argument("--reps", type=int, default=5)
  ap.add_argument("--check", type=int, default=2)
  ap.add_argument("--sigma-scale", type=float, default=1.0)
  ap.add_argument("--rows-max", type=int, default=40, help="indexes drawn from [0, rows-max)")
  ap.add_argument("--cpu", action="store_true", help="also time the CPU checker, all threads")
  ap.add_argument("--pack", type=int, default=0, help="coder streams per CTA (tfc_configure)")
  args = ap.parse_args()
  from compression_b200.python.layers import entropy_models
  from compression_b200.python.ops import range_coding_ops as rc
  dev = torch.device("cuda", 0)
  if args.pack:
    import compression_b200
    compression_b200.configure(coder_streams_per_cta=args.pack)
  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))
  gc = entropy_models.GaussianConditional(None, table)
  gc.build_tables(dev)
  cdf, size, off = gc._quantized_cdf, gc._cdf_length, gc._offset
  g = torch.Generator(device="cpu").manual_seed(1)
  S, n = args.streams, args.n
  # scales as a trained hyperprior produces them: mostly small, log-uniform over the low rows
  idx = torch.randint(0, args.rows_max, (S, n), generator=g, dtype=torch.int32)
  sig = torch.as_tensor(table, dtype=torch.float32)[idx.long()] * args.sigma_scale
  sym = torch.floor(torch.randn((S, n), generator=g) * sig + .5).to(torch.int32)
  idx_d, sym_d = idx.to(dev), sym.to(dev)
  out, nbytes = rc.unbounded_index_range_encode_batch(sym_d, idx_d, cdf, size, off, 16, 4, 0)
  dec = rc.unbounded_index_range_decode_batch(out, nbytes, idx_d, cdf, size, off, 16, 4, 0)
  assert torch.equal(dec, sym_d), "round trip failed"
  if args.check:
    from oracle import oracle_lib
    coder = oracle_lib.load("best")
    strings = rc.strings_to_host(out[:args.check], nbytes[:args.check])
    c, s_, o = cdf.cpu().numpy(), size.cpu().numpy(), off.cpu().numpy()
    for i in range(args.check):
      want = coder.ui_encode(sym[i].numpy(), idx[i].numpy(), c, s_, o, 16, 4)
      assert strings[i] == want, "stream %d differs from the %s coder" % (i, coder.kind)

  def timeit(fn):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.reps):
      fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / args.reps

  cap = out.shape[1]
  enc_ms = timeit(lambda: rc.unbounded_index_range_encode_batch(sym_d, idx_d, cdf, size, off,
                                                                16, 4, 0, cap=cap))
  dec_ms = timeit(lambda: rc.unbounded_index_range_decode_batch(out, nbytes, idx_d, cdf, size,
                                                                off, 16, 4, 0, check=False))
  total = S * n
  bits = float(nbytes.sum().item()) * 8 / total
  clk = 1.965e9
  res = {
      "streams": S, "symbols_per_stream": n, "bits_per_symbol": bits,
      "encode_ms": enc_ms, "decode_ms": dec_ms,
      "encode_msym_s": total / enc_ms / 1e3, "decode_msym_s": total / dec_ms / 1e3,
      "encode_in_MB_s": total * 8 / enc_ms / 1e3, "decode_out_MB_s": total * 4 / dec_ms / 1e3,
      "encode_cycles_per_symbol_per_stream": enc_ms * 1e-3 * clk / n / max(1, -(-S // (148 * 16))),
      "decode_cycles_per_symbol_per_stream": dec_ms * 1e-3 * clk / n / max(1, -(-S // (148 * 16))),
  }
  if args.cpu:
    threads = len(os.sched_getaffinity(0))
    t0 = time.perf_counter()
    o2, n2 = coder.ui_encode_batch(sym.numpy(), idx.numpy(), c, s_, o, threads=threads)
    t1 = time.perf_counter()
    coder.ui_decode_batch(o2, n2, idx.numpy(), n, c, s_, o, threads=threads)
    t2 = time.perf_counter()
    res["cpu"] = {"kind": coder.kind, "threads": threads,
                  "encode_msym_s": total / (t1 - t0) / 1e6, "decode_msym_s": total / (t2 - t1) / 1e6}
  print(json.dumps(res))


if __name__ == "__main__":
  main()
