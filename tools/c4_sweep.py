#!/usr/bin/env python
"""BASELINE.json configs[3] / SURVEY.md §8(d) C4: the coder sweep.

For S in {1K, 4K, 16K, 64K} streams of L = 4096 symbols, precision 16, overflow_width 4:
  * unbounded-index encode + decode with the 64-row Gaussian table of bmshj2018, indexes
    ~ U{0..63}, symbols ~ floor(N(0, table[idx]) + 1/2) (about tail_mass of them escape);
  * bounded RangeEncode / RangeDecode, m = 64, one shared CDF (discretised Gaussian, sigma 6);
  * pmf_to_quantized_cdf on S rows of n in {22, 64, 256, 1480} entries, uniform-random rows
    (range_coding_helper_kernels_test.cc:58-68) and symmetric Gaussians (tie-heavy).
Every configuration is checked bit-exact against the CPU checker (all streams / rows up to
--check-streams, an evenly spaced sample above that) and timed with CUDA events; the checker is
timed on the same sample with all host threads.  One JSON line per configuration.

  python tools/c4_sweep.py [--streams 1024,4096,16384,65536] [--n 4096] [--reps 5]
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


def timeit(fn, reps):
  fn()
  torch.cuda.synchronize()
  a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  a.record()
  for _ in range(reps):
    fn()
  b.record()
  torch.cuda.synchronize()
  return a.elapsed_time(b) / reps


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--streams", default="1024,4096,16384,65536")
  ap.add_argument("--n", type=int, default=4096)
  ap.add_argument("--reps", type=int, default=5)
  ap.add_argument("--check-streams", type=int, default=1024)
  ap.add_argument("--pmf-n", default="22,64,256,1480")
  ap.add_argument("--skip", default="", help="comma list of: unbounded,bounded,pmf")
  ap.add_argument("--coder-lanes", type=int, default=0,
                  help="0: automatic (lane per stream from 2048 streams on), 1: warps per stream "
                       "always, 2: lane per stream always")
  args = ap.parse_args()
  skip = set(args.skip.split(","))
  from compression_b200.python.layers import entropy_models
  from compression_b200.python.ops import range_coding_ops as rc
  from oracle import oracle_lib
  coder = oracle_lib.load("best")
  import compression_b200
  compression_b200.configure(coder_lanes=args.coder_lanes)
  threads = len(os.sched_getaffinity(0))
  dev = torch.device("cuda", 0)
  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))
  gc = entropy_models.GaussianConditional(None, table)
  gc.build_tables(dev)
  cdf, size, off = gc._quantized_cdf, gc._cdf_length, gc._offset
  c_h, s_h, o_h = cdf.cpu().numpy(), size.cpu().numpy(), off.cpu().numpy()
  table_d = torch.as_tensor(table, dtype=torch.float32, device=dev)
  n = args.n
  # bounded coder: m = 64 symbols, one shared CDF
  k = np.arange(64, dtype=np.float64) - 31.5
  pmf64 = np.exp(-0.5 * (k / 6.0) ** 2)
  pmf64 = (pmf64 / pmf64.sum()).astype(np.float32)
  cdf64 = rc.pmf_to_quantized_cdf(torch.as_tensor(pmf64, device=dev), 16)
  cdf64_h = cdf64.cpu().numpy()
  assert np.array_equal(cdf64_h, coder.pmf_to_cdf(pmf64, 16))

  for S in [int(s) for s in args.streams.split(",")]:
    g = torch.Generator(device=dev).manual_seed(1)
    pick = np.arange(S) if S <= args.check_streams else \
        np.linspace(0, S - 1, args.check_streams).astype(np.int64)
    pick_d = torch.as_tensor(pick, device=dev)
    total = S * n
    if "unbounded" not in skip:
      idx = torch.randint(0, 64, (S, n), generator=g, dtype=torch.int32, device=dev)
      sym = torch.floor(torch.randn((S, n), generator=g, device=dev) * table_d[idx.long()] + .5)
      sym = sym.to(torch.int32)
      out, nbytes = rc.unbounded_index_range_encode_batch(sym, idx, cdf, size, off, 16, 4, 0)
      dec = rc.unbounded_index_range_decode_batch(out, nbytes, idx, cdf, size, off, 16, 4, 0)
      assert torch.equal(dec, sym), "round trip failed"
      del dec
      sym_h, idx_h = sym[pick_d].cpu().numpy(), idx[pick_d].cpu().numpy()
      got = rc.strings_to_host(out[pick_d], nbytes[pick_d])
      t0 = time.perf_counter()
      o2, n2 = coder.ui_encode_batch(sym_h, idx_h, c_h, s_h, o_h, threads=threads)
      t1 = time.perf_counter()
      d2 = coder.ui_decode_batch(o2, n2, idx_h, n, c_h, s_h, o_h, threads=threads)
      t2 = time.perf_counter()
      assert np.array_equal(d2, sym_h)
      for i in range(len(pick)):
        assert got[i] == o2[i, :n2[i]].tobytes(), "stream %d differs from the %s coder" % (
            pick[i], coder.kind)
      escapes = float(((sym_h - o_h[idx_h] < 0) | (sym_h - o_h[idx_h] >= s_h[idx_h] - 2)).mean())
      cap = out.shape[1]
      enc_ms = timeit(lambda: rc.unbounded_index_range_encode_batch(
          sym, idx, cdf, size, off, 16, 4, 0, cap=cap), args.reps)
      dec_ms = timeit(lambda: rc.unbounded_index_range_decode_batch(
          out, nbytes, idx, cdf, size, off, 16, 4, 0, check=False), args.reps)
      out_bytes = float(nbytes.sum().item())
      print(json.dumps({
          "op": "unbounded_index_range_encode/decode", "coder_lanes": args.coder_lanes, "streams": S, "symbols_per_stream": n,
          "bits_per_symbol": out_bytes * 8 / total, "escape_fraction": escapes,
          "bit_exact_streams_checked": int(len(pick)), "checker": coder.kind,
          "encode_ms": enc_ms, "decode_ms": dec_ms,
          "encode_msym_s": total / enc_ms / 1e3, "decode_msym_s": total / dec_ms / 1e3,
          "encode_in_MB_s": total * 8 / enc_ms / 1e3, "encode_out_MB_s": out_bytes / enc_ms / 1e3,
          "decode_in_MB_s": (out_bytes + total * 4) / dec_ms / 1e3,
          "decode_out_MB_s": total * 4 / dec_ms / 1e3,
          "cpu": {"threads": threads, "streams": int(len(pick)),
                  "encode_msym_s": len(pick) * n / (t1 - t0) / 1e6,
                  "decode_msym_s": len(pick) * n / (t2 - t1) / 1e6}}), flush=True)
      del idx, sym, out, nbytes
    if "bounded" not in skip:
      u = torch.randn((S, n), generator=g, device=dev) * 6.0 + 31.5
      data = torch.clamp(torch.floor(u + .5), 0, 63).to(torch.int16)
      del u
      out, nbytes = rc.range_encode_batch(data, cdf64.reshape(1, 65), 16, debug_level=0)
      dec = rc.range_decode_batch(out, nbytes, (n,), cdf64.reshape(1, 65), 16, debug_level=0)
      assert torch.equal(dec, data), "bounded round trip failed"
      del dec
      data_h = data[pick_d].cpu().numpy()
      got = rc.strings_to_host(out[pick_d], nbytes[pick_d])
      m = min(len(pick), 256)
      t0 = time.perf_counter()
      want = [coder.rc_encode(data_h[i], cdf64_h.reshape(1, 65), 16) for i in range(m)]
      t1 = time.perf_counter()
      for i in range(m):
        assert got[i] == want[i], "bounded stream %d differs from the %s coder" % (
            pick[i], coder.kind)
      cap = out.shape[1]
      enc_ms = timeit(lambda: rc.range_encode_batch(
          data, cdf64.reshape(1, 65), 16, debug_level=0, cap=cap), args.reps)
      dec_ms = timeit(lambda: rc.range_decode_batch(
          out, nbytes, (n,), cdf64.reshape(1, 65), 16, debug_level=0), args.reps)
      out_bytes = float(nbytes.sum().item())
      print(json.dumps({
          "op": "range_encode/decode (bounded, m=64)", "streams": S, "symbols_per_stream": n,
          "bits_per_symbol": out_bytes * 8 / total, "bit_exact_streams_checked": m,
          "checker": coder.kind, "encode_ms": enc_ms, "decode_ms": dec_ms,
          "encode_msym_s": total / enc_ms / 1e3, "decode_msym_s": total / dec_ms / 1e3,
          "encode_in_MB_s": total * 2 / enc_ms / 1e3, "encode_out_MB_s": out_bytes / enc_ms / 1e3,
          "decode_out_MB_s": total * 2 / dec_ms / 1e3,
          "cpu": {"threads": 1, "streams": m, "encode_msym_s": m * n / (t1 - t0) / 1e6}}),
            flush=True)
      del data, out, nbytes
    if "pmf" not in skip:
      for pn in [int(x) for x in args.pmf_n.split(",")]:
        for kind in ("uniform", "gaussian"):
          if kind == "uniform":
            p = torch.rand((S, pn), generator=g, device=dev) + 1e-6
            p = p / p.sum(-1, keepdim=True)
          else:
            sig = torch.rand((S, 1), generator=g, device=dev) * (pn / 6.0) + 0.2
            kk = torch.arange(pn, device=dev, dtype=torch.float32) - (pn - 1) / 2
            p = torch.exp(-0.5 * (kk / sig) ** 2)
            p = p / p.sum(-1, keepdim=True)
          p = p.contiguous()
          got = rc.pmf_to_quantized_cdf(p, 16)
          rows = pick if len(pick) <= 256 else pick[:: len(pick) // 256]
          p_h = p[torch.as_tensor(rows, device=dev)].cpu().numpy()
          t0 = time.perf_counter()
          want = coder.pmf_to_cdf(p_h, 16, threads=threads)
          t1 = time.perf_counter()
          assert np.array_equal(got[torch.as_tensor(rows, device=dev)].cpu().numpy(), want), \
              "pmf_to_quantized_cdf differs (%s, n=%d)" % (kind, pn)
          ms = timeit(lambda: rc.pmf_to_quantized_cdf(p, 16), max(1, args.reps // 2))
          print(json.dumps({
              "op": "pmf_to_quantized_cdf", "rows": S, "n": pn, "pmf": kind,
              "bit_exact_rows_checked": int(len(rows)), "checker": coder.kind, "ms": ms,
              "rows_per_s": S / ms * 1e3, "entries_per_s": S * pn / ms * 1e3,
              "MB_s": S * pn * 8 / ms / 1e3,
              "cpu": {"threads": threads, "rows": int(len(rows)),
                      "rows_per_s": len(rows) / (t1 - t0)}}), flush=True)
          del p, got


if __name__ == "__main__":
  main()
