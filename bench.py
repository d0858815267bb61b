#!/usr/bin/env python
"""Headline benchmark: bmshj2018-hyperprior encode+decode throughput (MPix/s) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|tiny]

A step = compress + decompress of one batch of synthetic images (BASELINE.json configs[1]:
192 filters, 768x512, batch 64 per GPU; random-init weights, seed 0).  N > 1 is launched by
torchrun, one rank per GPU; every rank codes its own batch (images are independent units, no
collective on the data path), so scaling is weak.  One JSON line is printed by rank 0.

  value   device-resident: images already in HBM, strings stay in HBM, x_hat stays in HBM
  e2e     through the public API with HOST buffers: pinned uint8 images -> H2D -> compress ->
          strings D2H -> strings H2D -> decompress -> uint8 x_hat D2H, all inside the timing
  roofline      the conv kernels (dominant): algorithmic FLOPs / CUDA-event time vs the
                measured bf16 tensor peak (MEASURED_PEAKS.json)
  cpu_baseline  the same pipeline on the host cores (reference range coder from oracle/_ref
                when built, torch-CPU float32 restatement of the transforms), bounded sample
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

# before the CUDA context exists (see compression_b200/__init__.py): one hardware queue per stream
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (height, width, batch per GPU, filters)
    "c2": (512, 768, 64, 192),
    "c3": (1080, 1920, 32, 192),
    "tiny": (128, 192, 4, 192),
}
# DRAM traffic (read + write) of the conv launches of one step as ncu measures it: the sum over
# the 14 layers in profiles/r1e_conv_full_summary.md (batch 64, 768x512); other workloads: null
NCU_TRAFFIC_BYTES_PER_STEP = {"c2": 15.67e9}
FLOP_PER_IMAGE_C2 = 149.31e9   # encode 75.61 + decode 73.70 GFLOP per 768x512 image


def parse_args():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=20)
  ap.add_argument("--warmup", type=int, default=3)
  ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
  ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
  ap.add_argument("--conv-impl", type=int, default=0, help="0 auto, 1 CUDA-core, 2 tcgen05")
  ap.add_argument("--cpu-sample-images", type=int, default=0,
                  help="images per CPU step (0: one per host thread, at most the batch)")
  ap.add_argument("--no-cpu-baseline", action="store_true")
  ap.add_argument("--verbose", action="store_true", help="stage/layer breakdown on stderr")
  ap.add_argument("--depth", type=int, default=0,
                  help="batches in flight on separate CUDA streams (1: strictly serial steps; "
                       "0: 20, or fewer if 20 batches of activations would not fit in HBM)")
  ap.add_argument("--back-lag", type=int, default=0,
                  help="e2e leg: batches queued between a compress and the host's return for its "
                       "strings (0: the batches in flight, at most 12)")
  ap.add_argument("--host-slots", type=int, default=0,
                  help="e2e leg: sets of pinned staging buffers (0: twice the batches in flight, "
                       "at most 8 GB)")
  ap.add_argument("--tiles-per-cta", type=int, default=8,
                  help="transform kernels: tiles per CTA when depth > 1 (0: persistent grid)")
  ap.add_argument("--coder-pack", type=int, default=0,
                  help="coder streams per CTA when depth > 1 (0: by the number of streams in "
                       "flight: 4 at batch 64 x 12, 2 at batch 32 x 8)")
  return ap.parse_args()


def measured_peaks():
  path = os.path.join(ROOT, "MEASURED_PEAKS.json")
  if os.path.exists(path):
    with open(path) as f:
      p = json.load(f)
    return dict(hbm_gbs=p["hbm_gbs"], bf16_tflops=p["bf16_tflops"],
                bf16_tflops_sustained=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                source="measured")
  return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0,
              source="fallback")


# ---- algorithmic work -------------------------------------------------------------------------
def conv_flops(model, h, w):
  """Algorithmic FLOPs per image of every conv (+GDN) layer for an h x w input, following
  SURVEY §8a: 2*MACs of the dense contraction; a transposed conv counts no stuffed zeros."""
  from compression_b200.python.layers import gdn as gdn_lib
  total = {}

  def run(seq, name, hh, ww, cin):
    for i, l in enumerate(seq.layers):
      ho, wo = l.compute_output_shape((1, hh, ww, cin))[1:3]
      kk = l.kernel_support[0] * l.kernel_support[1]
      if l.strides_up[0] > 1:
        f = 2.0 * hh * ww * cin * l.filters * kk
      else:
        f = 2.0 * ho * wo * cin * l.filters * kk
      if isinstance(l.activation, gdn_lib.GDN):
        f += 2.0 * ho * wo * l.filters * l.filters
      total["%s%d" % (name, i)] = f
      hh, ww, cin = ho, wo, l.filters
    return hh, ww

  nf = model.num_filters
  hy, wy = run(model.analysis_transform, "A", h, w, 3)
  hz, wz = run(model.hyper_analysis_transform, "HA", hy, wy, nf)
  hs = run(model.hyper_synthesis_transform, "HS", hz, wz, nf)
  run(model.synthesis_transform, "S", hy, wy, nf)
  enc = sum(v for k, v in total.items() if k[0] in "AH")     # A + HA + HS
  dec = sum(v for k, v in total.items() if k.startswith("HS") or k.startswith("S"))
  return total, enc, dec


class ConvTimer(object):
  """Brackets every SignalConv2D call with CUDA events on the current stream."""

  def __init__(self):
    self.pairs = []
    self.on = False

  def install(self):
    import torch
    from compression_b200.python.layers import signal_conv
    timer = self
    orig = signal_conv.SignalConv2D.__call__

    def timed(layer, inputs, out_dtype=None):
      if not timer.on:
        return orig(layer, inputs, out_dtype)
      a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      a.record()
      out = orig(layer, inputs, out_dtype)
      b.record()
      timer.pairs.append((layer.name, tuple(inputs.shape), a, b))
      return out

    signal_conv.SignalConv2D.__call__ = timed

  def total_ms(self):
    return sum(a.elapsed_time(b) for _, _, a, b in self.pairs)


# ---- clocks -------------------------------------------------------------------------------------
class ClockSampler(object):
  QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
           "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
           "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

  def __init__(self, index):
    self.index = index
    self.proc = None
    self.path = None

  def start(self):
    try:
      fd, self.path = tempfile.mkstemp(suffix=".csv")
      os.close(fd)
      self.proc = subprocess.Popen(
          ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.QUERY,
           "--format=csv,noheader,nounits", "-lms", "100"],
          stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
    except OSError:
      self.proc = None

  def stop(self):
    if self.proc is None:
      return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    self.proc.terminate()
    try:
      self.proc.wait(timeout=5)
    except subprocess.TimeoutExpired:
      self.proc.kill()
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for line in open(self.path):
      f = [t.strip() for t in line.split(",")]
      if len(f) < 9:
        continue
      try:
        sm.append(float(f[1])); mx.append(float(f[2]))
      except ValueError:
        continue
      for n, v in zip(names, f[5:9]):
        if v.lower().startswith("active"):
          reasons.add(n)
    os.unlink(self.path)
    return {"sm_mhz": float(np.median(sm)) if sm else None,
            "sm_max_mhz": float(max(mx)) if mx else None, "reasons": sorted(reasons),
            "samples": len(sm)}


# ---- CPU arm (cpu_baseline and --impl reference) ---------------------------------------------------
def host_threads():
  try:
    return len(os.sched_getaffinity(0))
  except AttributeError:
    return os.cpu_count() or 1


def cpu_arm(h, w, images, steps, warmup, weights=None, seed=0):
  """Times compress+decompress of `images` synthetic images per step on the host cores.
  Returns (MPix/s, ms_per_step, kind, cores)."""
  import torch
  from oracle import cpu_codec, oracle_lib
  coder = oracle_lib.load("best")
  threads = host_threads()
  codec = cpu_codec.CpuBmshj2018(weights or cpu_codec.random_weights(192, seed), coder, threads)
  g = torch.Generator().manual_seed(seed)
  x = torch.rand((images, h, w, 3), generator=g, dtype=torch.float32).numpy()
  times = []
  for it in range(warmup + steps):
    t0 = time.perf_counter()
    c = codec.compress(x)
    x_hat, _ = codec.decompress(c)
    t1 = time.perf_counter()
    if it >= warmup:
      times.append(t1 - t0)
  sec = float(np.mean(times))
  return images * h * w / sec / 1e6, sec * 1e3, coder.kind, threads


def workload_name(args):
  h, w, batch, nf = WORKLOADS[args.workload]
  return "bmshj2018-%d encode+decode %dx%d batch %d per GPU (%s)" % (nf, w, h, batch, args.workload)


def cpu_sample_images(args):
  h, w, batch, nf = WORKLOADS[args.workload]
  n = args.cpu_sample_images if args.cpu_sample_images > 0 else host_threads()
  return max(1, min(n, batch))


def run_reference(args):
  """The reference's CPU implementation of the path on the host cores: its own C++ range coder
  (oracle/_ref, one thread per image) + the torch-CPU float32 restatement of the transforms
  (TensorFlow 1.15 cannot run in this image).  Rank 0 only."""
  rank = int(os.environ.get("RANK", "0"))
  if rank != 0:
    return
  h, w, batch, nf = WORKLOADS[args.workload]
  images = cpu_sample_images(args)
  steps, warmup = max(1, args.steps), max(0, min(args.warmup, 1))
  mpix, ms, kind, cores = cpu_arm(h, w, images, steps, warmup)
  sample = "%d of %d images of %dx%d per step, all %d host threads" % (images, batch, w, h, cores)
  line = {
      "impl": "reference", "metric": "MPix/s encode+decode, bmshj2018 hyperprior",
      "value": mpix, "unit": "MPix/s", "n_gpus": args.gpus, "steps": steps,
      "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
      "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
      "config": {"workload": workload_name(args), "weights": "random-init seed 0",
                 "sample": sample,
                 "note": "CPU: torch-CPU float32 transforms + %s range coder (one thread per "
                         "image); TF 1.15 cannot run in this image" % kind},
      "cpu_baseline": {"value": mpix, "unit": "MPix/s", "cores": cores, "kind": kind,
                       "sample": sample},
      "e2e": {"value": mpix, "unit": "MPix/s", "h2d_bytes_per_step": 0,
              "d2h_bytes_per_step": 0},
      "gpu_launches": 0,
  }
  print(json.dumps(line))


# ---- B200 arm ----------------------------------------------------------------------------------------
def run_b200(args):
  import torch
  import torch.distributed as dist
  import compression_b200 as tfc
  from compression_b200.examples import bmshj2018
  from compression_b200.python.ops import range_coding_ops as rc

  world = int(os.environ.get("WORLD_SIZE", "1"))
  rank = int(os.environ.get("RANK", "0"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  if not torch.cuda.is_available():
    raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference")
  torch.cuda.set_device(local)
  dev = torch.device("cuda", local)
  if world > 1:
    dist.init_process_group("nccl", device_id=dev)
  h, w, batch, nf = WORKLOADS[args.workload]
  model = bmshj2018.Bmshj2018Model(nf, dev, seed=0, impl=args.conv_impl)
  g = torch.Generator().manual_seed(1000 + rank)
  x_u8_host = torch.randint(0, 256, (batch, h, w, 3), generator=g, dtype=torch.uint8).pin_memory()
  x_dev = (x_u8_host.to(dev).float() / 255).contiguous()
  timer = ConvTimer()
  timer.install()

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  pixels_per_step = batch * h * w
  if args.depth <= 0:
    # ~230 bytes of activations per input pixel stay reserved per batch in flight (measured:
    # 15.2 GB for 32 x 1080p, 5.8 GB for 64 x 768x512); keep the pipeline inside 3/4 of HBM
    total_mem = torch.cuda.get_device_properties(dev).total_memory
    # measured (profiles/r1h_pipeline_sweep.txt): C2 device 1632 / e2e 1471 MPix/s at 12 batches in
    # flight, 1738 / 1569 at 16, 1731 / 1640 at 20, 1734 / 1621 at 24
    # (C3 ran at 12 batches in flight, 1299 device / 1196 e2e MPix/s, which 230 B/pixel would put at
    # 183 GB: the figure is an upper bound, 200 B/pixel admits that depth)
    args.depth = max(1, min(20, int(0.85 * total_mem / (200.0 * pixels_per_step))))
  if args.coder_pack <= 0:
    # streams in flight / CTAs wanted: the coder is latency bound per stream, and 4 streams per
    # CTA is the measured optimum once there are more streams than ~4 per SM
    args.coder_pack = max(1, min(4, int(round(batch * args.depth / 150.0))))
  if args.host_slots <= 0:
    # e2e leg: twice as many sets of pinned staging buffers as batches in flight, so that neither
    # host wait (strings of batch k, x_hat of the batch whose buffers are reused) comes before the
    # GPU can have finished that work; at most 8 GB pinned per rank (a set is x_hat
    # plus the string slabs at capacity, ~1.6 bytes per image byte).  Measured
    # (profiles/r1l_e2e_host_waits.txt): C3 e2e 1189 MPix/s at depth + 1 sets and a lag of
    # depth - 1, 1384 at lag 6, 1504 at 24 sets and lag 12; C2 1592 -> 1675.
    args.host_slots = max(args.depth + 1,
                          min(2 * args.depth, int(8e9 / (1.6 * 3 * pixels_per_step))))
  if args.back_lag <= 0:
    args.back_lag = max(1, min(args.depth, 12, args.host_slots // 2))
  pipe = model.pipeline(args.depth, args.tiles_per_cta, args.coder_pack, args.back_lag,
                        args.host_slots)

  with pipe:
    # every stream of the pipeline must have run once before the timed region: a stream's first
    # batch allocates its activation pool (cudaMalloc synchronises the device) and stages its
    # weight copies
    n_warm = max(args.warmup, 3, args.depth + 1)
    pipe.roundtrip_device([x_dev] * n_warm)
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
      clocks.start()
    launches0 = tfc.launch_count()
    timer.on = True
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    pipe.roundtrip_device([x_dev] * args.steps)
    stop.record()
    barrier()
    timer.on = False
    launches = tfc.launch_count() - launches0
    clk = clocks.stop() if rank == 0 else None
    ms = torch.tensor([start.elapsed_time(stop)], device=dev)
    conv_ms = timer.total_ms()
    if world > 1:
      dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms.item()) / args.steps

    # e2e: same metric through host buffers (H2D image, D2H strings, H2D strings, D2H x_hat)
    seen = {"strings": 0, "pixels": 0}

    def on_result(k, compressed, x_hat_u8):
      seen["strings"] = compressed.string.total_bytes + compressed.side_string.total_bytes
      seen["pixels"] += int(x_hat_u8.shape[0]) * int(x_hat_u8.shape[1]) * int(x_hat_u8.shape[2])

    pipe.roundtrip_host([x_u8_host] * (pipe.host_slots), on_result)
    barrier()
    seen["pixels"] = 0
    t0 = time.perf_counter()
    pipe.roundtrip_host([x_u8_host] * args.steps, on_result)
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], device=dev)
    assert seen["pixels"] == args.steps * pixels_per_step, "e2e: not every batch came back"
    string_bytes = seen["strings"]
    if world > 1:
      dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_s.item()) * 1e3 / args.steps
    h2d_bytes, d2h_bytes = int(pipe.h2d_bytes), int(pipe.d2h_bytes)
    if world > 1:
      dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_s.item()) * 1e3 / args.steps

  # the conv kernels alone, one batch, nothing else on the GPU (explains the roofline figure
  # measured under concurrency above)
  model.decompress(model.compress(x_dev))      # stage weights for this stream, warm the caches
  timer.pairs, timed_pairs = [], timer.pairs
  timer.on = True
  model.decompress(model.compress(x_dev))
  torch.cuda.synchronize()
  timer.on = False
  conv_ms_isolated = timer.total_ms()
  timer.pairs, isolated_pairs = timed_pairs, timer.pairs

  # the range coder alone (latency bound: one serial chain per image): CUDA-event time of the
  # y encode / decode launches, one CTA per stream
  y, z, sigma = model.analyse(x_dev)
  gc = model._conditional(sigma)
  ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
  string = gc.compress_device(y)
  gc.decompress_device(string, torch.bfloat16)
  ev[0].record()
  string = gc.compress_device(y)
  ev[1].record()
  gc.decompress_device(string, torch.bfloat16)
  ev[2].record()
  torch.cuda.synchronize()
  y_symbols = int(y[0].numel())
  coder_ms = (ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]))
  del y, z, sigma, string

  if rank != 0:
    if world > 1:
      dist.destroy_process_group()
    return
  if args.verbose:
    timer.pairs = isolated_pairs
    stage_breakdown(model, x_dev, timer, 1)
    timer.pairs = timed_pairs
  pixels = batch * h * w
  peaks = measured_peaks()
  per_layer, enc_f, dec_f = conv_flops(model, h, w)
  flops_step = (enc_f + dec_f) * batch
  conv_ms_step = conv_ms / args.steps
  n_conv = len(timer.pairs) // max(args.steps, 1)
  achieved = flops_step / (conv_ms_step * 1e-3) / 1e12 if conv_ms_step > 0 else 0.0
  peak = peaks["bf16_tflops_sustained"]
  line = {
      "metric": "MPix/s encode+decode, bmshj2018 hyperprior",
      "value": world * pixels / (ms_per_step * 1e-3) / 1e6, "unit": "MPix/s", "n_gpus": world,
      "steps": args.steps, "warmup": n_warm, "ms_per_step": ms_per_step,
      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
      "data": "synthetic",
      "config": {"workload": workload_name(args),
                 "weights": "random-init seed 0", "l2": "inputs_exceed_l2 (%.0f MB images)"
                 % (x_dev.numel() * 4 / 1e6), "parallelism": "batch-sharded x%d" % world,
                 "conv_impl": args.conv_impl,
                 "pipeline": {"batches_in_flight": args.depth, "e2e_back_lag": pipe.back_lag, "e2e_host_slots": pipe.host_slots,
                              "tiles_per_cta": args.tiles_per_cta if args.depth > 1 else 0,
                              "coder_streams_per_cta": args.coder_pack if args.depth > 1 else 0,
                              "hbm_peak_allocated_gb": round(
                                  torch.cuda.max_memory_allocated(dev) / 1e9, 1)}},
      "e2e": {"value": world * pixels / (e2e_ms * 1e-3) / 1e6, "unit": "MPix/s",
              "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
              "string_bytes_per_step": int(string_bytes), "ms_per_step": e2e_ms},
      "gpu_launches": int(launches),
      "clocks": clk,
      # The transforms dominate the GPU's SM-time (the coder holds ~16 CTAs for longer, see
      # "coder").  achieved/frac: the 17 conv launches of one batch timed per launch with CUDA
      # events, the batch alone on the GPU; "in_pipeline": the same brackets inside the timed
      # region, where they include waiting for SMs shared with other batches' kernels.
      "roofline": {"bound": "tensor", "kernel": "SignalConv2D kernels (%d launches/step)" % n_conv,
                   "achieved": flops_step / (conv_ms_isolated * 1e-3) / 1e12, "peak": peak,
                   "unit": "TFLOP/s",
                   "frac": flops_step / (conv_ms_isolated * 1e-3) / 1e12 / peak,
                   "traffic": NCU_TRAFFIC_BYTES_PER_STEP.get(args.workload),
                   "peak_source": peaks["source"] + " bf16 sustained",
                   "ms_per_step": conv_ms_isolated, "flops_per_step": flops_step,
                   "in_pipeline": {"ms_per_step": conv_ms_step, "achieved": achieved,
                                   "frac": achieved / peak,
                                   "share_of_step": conv_ms_step / ms_per_step}},
      "coder": {"bound": "latency", "kernel": "encode/decode_unbounded_v4 (y, one CTA per image)",
                "symbols_per_stream": y_symbols, "streams": batch,
                "encode_ms": coder_ms[0], "decode_ms": coder_ms[1],
                "encode_cycles_per_symbol": coder_ms[0] * 1e-3 * (clk["sm_mhz"] or 1965.0) * 1e6
                                            / y_symbols,
                "decode_cycles_per_symbol": coder_ms[1] * 1e-3 * (clk["sm_mhz"] or 1965.0) * 1e6
                                            / y_symbols,
                "hbm_frac": batch * y_symbols * 12.6 / (sum(coder_ms) * 1e-3) / 1e9
                            / peaks["hbm_gbs"]},
  }
  if not args.no_cpu_baseline and world == 1:
    images = cpu_sample_images(args)
    mpix, cms, kind, cores = cpu_arm(h, w, images, 2, 1, weights=model.export_weights())
    line["cpu_baseline"] = {"value": mpix, "unit": "MPix/s", "cores": cores, "kind": kind,
                            "sample": "%d of %d images of %dx%d, 2 steps after 1 warm-up, all host threads; "
                                      "torch-CPU f32 transforms + %s range coder"
                                      % (images, batch, w, h, kind)}
  print(json.dumps(line))
  if world > 1:
    dist.destroy_process_group()


def stage_breakdown(model, x_dev, timer, steps):
  """One instrumented step (synchronising between stages) + per-layer conv times."""
  import torch

  def timed(name, fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    sys.stderr.write("  %-28s %8.3f ms\n" % (name, (time.perf_counter() - t0) * 1e3))
    return out

  sys.stderr.write("stage breakdown (one synchronised step):\n")
  y, z, sigma = timed("analyse (A, HA, HS)", lambda: model.analyse(x_dev))
  side = timed("EB compress (z)", lambda: model.entropy_bottleneck.compress_device(z))
  gc = model._conditional(sigma)
  string = timed("GC compress (y)", lambda: gc.compress_device(y))
  nf = model.num_filters
  z_hat = timed("EB decompress (z)", lambda: model.entropy_bottleneck.decompress_device(
      side, tuple(z.shape[1:3]) + (nf,), nf, torch.bfloat16))
  sigma2 = timed("hyper synthesis", lambda: model.hyper_synthesis_transform(z_hat)[
      :, :y.shape[1], :y.shape[2], :].contiguous())
  gc2 = model._conditional(sigma2)
  y_hat = timed("GC decompress (y)", lambda: gc2.decompress_device(string, torch.bfloat16))
  timed("synthesis", lambda: model.synthesis_transform(y_hat))
  sys.stderr.write("conv layers (mean over %d timed steps):\n" % steps)
  agg = {}
  for name, shape, a, b in timer.pairs:
    key = (name, shape)
    agg.setdefault(key, []).append(a.elapsed_time(b))
  for (name, shape), v in agg.items():
    sys.stderr.write("  %-10s %-24s %8.3f ms x%d\n" % (name, shape, float(np.mean(v)),
                                                      len(v) // max(steps, 1)))


def main():
  args = parse_args()
  if args.impl == "reference":
    run_reference(args)
  else:
    run_b200(args)


if __name__ == "__main__":
  main()
