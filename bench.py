#!/usr/bin/env python
"""Headline benchmark: bmshj2018-hyperprior encode+decode throughput (MPix/s) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c3|tiny]
                  [--blocks parity,c3,c4,c5,c1 | none]

A step = compress + decompress of one batch of synthetic images (BASELINE.json configs[1]:
192 filters, 768x512, batch 64 per GPU; random-init weights, seed 0).  N > 1 is launched by
torchrun, one rank per GPU; every rank codes its own batch (images are independent units, no
collective on the data path), so scaling is weak.  One JSON line is printed by rank 0.

  value   device-resident: images already in HBM, strings stay in HBM, x_hat stays in HBM
  e2e     through the public API with HOST buffers: pinned uint8 images -> H2D -> compress ->
          strings D2H -> strings H2D -> decompress -> uint8 x_hat D2H, all inside the timing
  roofline      the conv kernels (dominant): algorithmic FLOPs / CUDA-event time.  `frac` =
                `frac_burst`: the 17 conv launches of one batch alone on the GPU against the
                measured BURST bf16 peak; `frac_step`: the same FLOPs over the whole timed step
                (ms_per_step) against the measured SUSTAINED peak (MEASURED_PEAKS.json)
  parity        after the timed region: images 0, 1, B/2, B-1 of the benched batch — the CPU
                checker (oracle/) must encode the GPU's symbols to the GPU's bytes and decode
                the GPU's bytes to the GPU's symbols, and the last image coded alone must give
                the bytes it gave inside the batch; the run fails otherwise
  cpu_baseline  the same pipeline on the host cores (reference range coder from oracle/_ref
                when built, torch-CPU float32 restatement of the transforms), bounded sample
  c1 c3 c4 c5   the other BASELINE.json configs, one block each in the same line: bls2017 on
                the host cores next to the GPU; 1920x1080 batch 256 split over the N ranks
                (strong scaling); the coder / pmf_to_quantized_cdf sweep; the bls2017 training
                step with its NCCL all-reduce at the N the run was launched with
"""
import argparse
import gc as pygc
import glob
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

# before the CUDA context exists (see compression_b200/__init__.py): one hardware queue per stream
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
# BASELINE.json: "gradient allreduce uses NCCL over NVLink only" — no InfiniBand / socket transport
os.environ.setdefault("NCCL_P2P_LEVEL", "NVL")
os.environ.setdefault("NCCL_IB_DISABLE", "1")
os.environ.setdefault("NCCL_NET_DISABLE", "1")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (height, width, batch per GPU, filters)
    "c2": (512, 768, 64, 192),
    "c3": (1080, 1920, 32, 192),
    "tiny": (128, 192, 4, 192),
}
ALL_BLOCKS = ("parity", "c3", "c4", "c5", "c1")
METRIC = "MPix/s encode+decode, bmshj2018 hyperprior"


def parse_args():
  ap = argparse.ArgumentParser()
  ap.add_argument("--gpus", type=int, default=1)
  ap.add_argument("--steps", type=int, default=20)
  ap.add_argument("--warmup", type=int, default=3)
  ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
  ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
  ap.add_argument("--blocks", default="",
                  help="comma list of extra blocks (%s), 'none', or empty for the default: all "
                       "of them on the c2 workload" % ",".join(ALL_BLOCKS))
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
  ap.add_argument("--coder-lanes", type=int, default=0,
                  help="range coder design: 0 automatic, 1 warps per stream, 2 one stream per lane")
  ap.add_argument("--c3", default="", help="c3 block overrides: micro,depth,pack (e.g. 16,16,8)")
  ap.add_argument("--conv-pair", type=int, default=0,
                  help="implicit-GEMM kernel on CTA pairs (cta_group::2): 0 auto, 1 off, 2 on")
  ap.add_argument("--conv-stages", type=int, default=0,
                  help="operand-ring stages of the implicit-GEMM kernel in the pipeline (0: 4)")
  ap.add_argument("--coder-smem-kb", type=int, default=0,
                  help="shared memory per coder CTA in the pipeline (0: 226 KB)")
  ap.add_argument("--coder-table-kb", type=int, default=0,
                  help="shared memory for the coder's staged tables in the pipeline (0: 72 KB)")
  ap.add_argument("--stage-priority", default="1",
                  help="transform stage streams from the high-priority pool (1) or the default one "
                       "(0), or 'a,h,s': stream priority levels (0, -1, -2, -3) of the analysis, "
                       "hyper-synthesis and synthesis stages")
  ap.add_argument("--debug-bits", type=int, default=0,
                  help="library experiment bits (tfc_configure key 99), for A/B runs")
  ap.add_argument("--c3-tail", default="", help="the same for the c3 block's micro-batches")
  ap.add_argument("--tail", default="",
                  help="'batches,pack': coder streams per CTA for the last batches of a run")
  ap.add_argument("--coder-pack", type=int, default=0,
                  help="coder streams per CTA when depth > 1 (0: by the number of streams in "
                       "flight: 4 at batch 64 x 12, 2 at batch 32 x 8)")
  return ap.parse_args()


def wanted_blocks(args):
  if args.blocks == "none":
    return ()
  if args.blocks:
    return tuple(b for b in args.blocks.split(",") if b in ALL_BLOCKS)
  return ALL_BLOCKS if args.workload == "c2" else ("parity",)


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


def ncu_conv_traffic(workload):
  """DRAM bytes (read + write) of the conv launches of one step as ncu measured them on the build
  that is being benched: the newest committed profiles/r*_conv_traffic.json whose `workload`
  matches.  (value, file) or (None, None): bench.py never invents this number."""
  best = (None, None)
  for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_conv_traffic.json"))):
    try:
      with open(path) as f:
        rec = json.load(f)
    except (OSError, ValueError):
      continue
    if rec.get("workload") == workload and rec.get("dram_bytes_per_step"):
      best = (float(rec["dram_bytes_per_step"]), os.path.relpath(path, ROOT))
  return best


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
  run(model.hyper_synthesis_transform, "HS", hz, wz, nf)
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
  """`nvidia-smi -lms 100` in the background.  The process is started BEFORE the warm-up (its
  NVML start-up takes a few hundred ms on a fresh box and holds driver locks that can stall
  kernel launches: inside a 0.27 s timed window that is the whole measurement) and only the
  samples stamped inside [mark(), stop()] are used."""
  QUERY = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
           "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
           "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

  def __init__(self, index):
    self.index = index
    self.proc = None
    self.path = None
    self.t0 = None

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

  def mark(self):
    import datetime
    self.t0 = datetime.datetime.now()

  def stop(self):
    import datetime
    if self.proc is None:
      return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    t1 = datetime.datetime.now()
    self.proc.terminate()
    try:
      self.proc.wait(timeout=5)
    except subprocess.TimeoutExpired:
      self.proc.kill()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    rows = []
    for line in open(self.path):
      f = [t.strip() for t in line.split(",")]
      if len(f) < 10:
        continue
      try:
        ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f")
        try:
          watts = float(f[4])
        except ValueError:
          watts = None
        rows.append((ts, float(f[2]), float(f[3]), f[6:10], watts))
      except ValueError:
        continue
    os.unlink(self.path)
    if self.t0 is not None:
      inside = [r for r in rows if self.t0 <= r[0] <= t1]
      if not inside and rows:            # a window shorter than the sampling period
        inside = [min(rows, key=lambda r: abs((r[0] - self.t0).total_seconds()))]
      rows = inside
    sm, mx, reasons = [r[1] for r in rows], [r[2] for r in rows], set()
    for r in rows:
      for n, v in zip(names, r[3]):
        if v.lower().startswith("active"):
          reasons.add(n)
    watts = [r[4] for r in rows if r[4] is not None]
    return {"sm_mhz": float(np.median(sm)) if sm else None,
            "sm_max_mhz": float(max(mx)) if mx else None, "reasons": sorted(reasons),
            "samples": len(sm), "power_w": float(np.median(watts)) if watts else None}


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
      "impl": "reference", "metric": METRIC,
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


# ---- parity of the benched batch (after the timed region) -----------------------------------------
def parity_block(model, x_dev, tag):
  """Images 0, 1, B/2, B-1 of `x_dev`: the CPU checker encodes the GPU's symbols + indexes to the
  GPU's bytes and decodes the GPU's bytes back to them (y and z), the GPU decoder returns the
  same symbols, and the last image coded alone gives the bytes it gave inside the batch.
  Raises on any mismatch."""
  import torch
  from oracle import oracle_lib
  coder = oracle_lib.load("best")
  t = model.export_tables()
  nf = model.num_filters
  b = int(x_dev.shape[0])
  picks = sorted(set([0, min(1, b - 1), b // 2, b - 1]))
  c = model.compress(x_dev)
  host = model.to_host(c)
  y, z, sigma = model.analyse(x_dev)
  gc = model._conditional(sigma)
  y_sym, y_idx, _ = gc._symbols_and_indexes(y.contiguous())
  z_sym = model.entropy_bottleneck._quantize_symbols(z)
  z_idx = np.broadcast_to(np.arange(nf, dtype=np.int32), tuple(z_sym.shape[1:])).copy()
  # the GPU decoder (which re-derives sigma from the side string) returns the coded symbols
  z_hat = model.entropy_bottleneck.decompress_device(c.side_string, tuple(c.z_shape) + (nf,), nf,
                                                     torch.bfloat16)
  sigma2 = model.hyper_synthesis_transform(z_hat)[:, :c.y_shape[0], :c.y_shape[1], :].contiguous()
  y_hat = model._conditional(sigma2).decompress_device(c.string)
  gpu_roundtrip = bool(torch.equal(y_hat.to(torch.int32), y_sym)) and bool(torch.equal(sigma2, sigma))
  strings_equal = symbols_equal = True
  for i in picks:
    ys, yi = y_sym[i].cpu().numpy(), y_idx[i].cpu().numpy()
    zs = z_sym[i].cpu().numpy()
    strings_equal &= host.string[i] == coder.ui_encode(ys, yi, t["gc_cdf"], t["gc_len"],
                                                       t["gc_off"], 16, 4)
    strings_equal &= host.side_string[i] == coder.ui_encode(zs, z_idx, t["eb_cdf"], t["eb_len"],
                                                            t["eb_off"], 16, 4)
    symbols_equal &= np.array_equal(
        coder.ui_decode(host.string[i], yi, t["gc_cdf"], t["gc_len"], t["gc_off"], 16, 4), ys)
    symbols_equal &= np.array_equal(
        coder.ui_decode(host.side_string[i], z_idx, t["eb_cdf"], t["eb_len"], t["eb_off"], 16, 4),
        zs)
  alone = model.to_host(model.compress(x_dev[b - 1:].contiguous()))
  alone_equal = alone.string[0] == host.string[b - 1] and \
      alone.side_string[0] == host.side_string[b - 1]
  out = {"config": tag, "images_checked": len(picks), "images": picks,
         "strings_equal": bool(strings_equal), "symbols_equal": bool(symbols_equal),
         "gpu_decoder_returns_symbols": gpu_roundtrip,
         "last_image_alone_equals_in_batch": bool(alone_equal),
         "symbols_per_image": int(y_sym[0].numel() + z_sym[0].numel()),
         "checker": coder.kind}
  if not (strings_equal and symbols_equal and gpu_roundtrip and alone_equal):
    raise SystemExit("PARITY FAILURE: %s" % json.dumps(out))
  return out


# ---- BASELINE config 3: 1920x1080, batch 256 split over the ranks ---------------------------------
def c3_block(args, model, world, rank, dev, barrier, allmax):
  import torch
  h, w, total = 1080, 1920, 256
  share = total // world
  # One 1080p stream is 1.57 M symbols = 85 ms of serial decode: the coder is latency bound per
  # image, so throughput needs ~300 images in flight per GPU whatever the share of the batch;
  # every batch in flight keeps its own activation pool (0.46 GB per image), which is what stops
  # at 18 micro-batches of 16 (134 GB reserved).  Measured (profiles/r2d_c3_sweep.txt): micro 16,
  # 14 in flight, 8 streams per coder CTA 1008 MPix/s; 18 in flight 1315; 4 per CTA 1539.
  # after the stream-aliasing fix (profiles/r2l_c3_sweep.txt): 24 in flight 1464 / 1461 e2e
  micro, depth, pack = min(16, share), 24, 4
  if args.c3:
    micro, depth, pack = [int(v) for v in args.c3.split(",")]
    micro = min(micro, share)
  nmicro = share // micro
  passes = 2 * world                        # every rank codes 512 images in the timed region
  g = torch.Generator().manual_seed(3000 + rank)
  xs_u8 = [torch.randint(0, 256, (micro, h, w, 3), generator=g, dtype=torch.uint8).pin_memory()
           for _ in range(2)]
  xs_dev = [(x.to(dev).float() / 255).contiguous() for x in xs_u8]
  c3_tail = None
  if args.c3_tail:
    c3_tail = tuple(int(v) for v in args.c3_tail.split(","))
  pipe = model.pipeline(depth, args.tiles_per_cta, pack, back_lag=max(2, depth // 2),
                        host_slots=depth + 6, tail=c3_tail)
  out = {"workload": "bmshj2018-192 encode+decode 1920x1080, batch 256 split over %d GPU(s)" % world,
         "images_per_gpu": share, "micro_batch": micro, "passes_timed": passes,
         "batches_in_flight": depth, "coder_streams_per_cta": pack, "scaling": "strong"}
  with pipe:
    n = nmicro * passes
    seq = [xs_dev[k % 2] for k in range(n)]
    pipe.roundtrip_device(seq[:min(n, depth + 1)])
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    pipe.roundtrip_device(seq)
    b.record()
    barrier()
    ms = allmax(a.elapsed_time(b))
    pix = world * n * micro * h * w
    out["value"] = pix / (ms * 1e-3) / 1e6
    out["ms_per_256_images"] = ms / passes
    seen = {"pixels": 0}

    def on_result(k, c, xh):
      seen["pixels"] += int(xh.shape[0]) * int(xh.shape[1]) * int(xh.shape[2])

    seq8 = [xs_u8[k % 2] for k in range(n)]
    pipe.roundtrip_host(seq8[:pipe.host_slots], on_result)
    barrier()
    seen["pixels"] = 0
    t0 = time.perf_counter()
    pipe.roundtrip_host(seq8, on_result)
    torch.cuda.synchronize()
    sec = allmax(time.perf_counter() - t0)
    assert seen["pixels"] == n * micro * h * w
    out["e2e"] = {"value": pix / sec / 1e6, "unit": "MPix/s",
                  "h2d_bytes_per_micro_batch": int(pipe.h2d_bytes),
                  "d2h_bytes_per_micro_batch": int(pipe.d2h_bytes)}
  out["unit"] = "MPix/s"
  out["hbm_peak_reserved_gb"] = round(torch.cuda.max_memory_reserved(dev) / 1e9, 1)
  if rank == 0:
    out["parity"] = parity_block(model, xs_dev[0][:4].contiguous(), "c3 (4 of a micro-batch)")
  barrier()
  return out


# ---- BASELINE config 5: bls2017 training step + NCCL all-reduce -----------------------------------
def c5_block(args, world, rank, dev, barrier, allmax):
  import torch
  import torch.distributed as dist
  from compression_b200.examples import bls2017
  batch, patch, steps, warmup = 64, 256, 10, 3
  model = bls2017.Bls2017Model(128, dev, seed=0)          # same seed: identical replicas
  trainer = bls2017.Bls2017Trainer(model, lmbda=0.01)
  g = torch.Generator(device=dev).manual_seed(100 + rank)  # per-rank crops and noise
  torch.manual_seed(200 + rank)
  x = torch.rand((batch, patch, patch, 3), generator=g, device=dev)
  for _ in range(warmup):
    stats = trainer.step(x, graph=True)
  barrier()
  a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  a.record()
  for _ in range(steps):
    stats = trainer.step(x, graph=True)
  b.record()
  torch.cuda.synchronize()
  ms = allmax(a.elapsed_time(b) / steps)
  # the exchange step alone: both flat buckets, as the step issues them
  ar_ms = 0.0
  if world > 1:
    barrier()
    a.record()
    for _ in range(20):
      trainer.main.all_reduce()
      trainer.aux.all_reduce()
    b.record()
    torch.cuda.synchronize()
    ar_ms = allmax(a.elapsed_time(b) / 20)
  check = trainer.main.param.double().sum().reshape(1)
  identical = True
  if world > 1:
    lo, hi = check.clone(), check.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    identical = float(hi - lo) == 0.0
  flop = 8.83e9 * batch                                    # fwd + bwd per image
  out = {"workload": "bls2017-128 training step, 256x256 crops, batch 64 per GPU, NCCL all-reduce",
         "n_gpus": world, "ms_per_step": ms, "allreduce_ms": ar_ms,
         "allreduce_bytes": int(4 * (trainer.main.param.numel() + trainer.aux.param.numel())),
         "images_per_s": world * batch / ms * 1e3, "tflops_per_gpu": flop / ms / 1e9,
         "parameters": trainer.num_parameters, "replicas_identical": bool(identical),
         "loss": stats["loss"], "bpp": stats["bpp"], "mse": stats["mse"], "cuda_graph": True,
         "nccl_env": {k: os.environ.get(k) for k in ("NCCL_P2P_LEVEL", "NCCL_IB_DISABLE",
                                                     "NCCL_NET_DISABLE")}}
  if not identical:
    raise SystemExit("c5: replicas diverged: %s" % json.dumps(out))
  del trainer, model
  torch.cuda.empty_cache()
  return out


# ---- BASELINE config 4: coder + pmf_to_quantized_cdf sweep ----------------------------------------
def c4_block(dev):
  import torch
  from compression_b200.python.layers import entropy_models
  from compression_b200.python.ops import range_coding_ops as rc
  from oracle import oracle_lib
  coder = oracle_lib.load("best")
  threads = host_threads()

  def timeit(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
      fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps

  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))
  gc = entropy_models.GaussianConditional(None, table)
  gc.build_tables(dev)
  cdf, size, off = gc._quantized_cdf, gc._cdf_length, gc._offset
  c_h, s_h, o_h = cdf.cpu().numpy(), size.cpu().numpy(), off.cpu().numpy()
  table_d = torch.as_tensor(table, dtype=torch.float32, device=dev)
  k = np.arange(64, dtype=np.float64) - 31.5
  pmf64 = np.exp(-0.5 * (k / 6.0) ** 2)
  pmf64 = (pmf64 / pmf64.sum()).astype(np.float32)
  cdf64 = rc.pmf_to_quantized_cdf(torch.as_tensor(pmf64, device=dev), 16)
  cdf64_h = cdf64.cpu().numpy()
  n, checked = 4096, 0
  rows_out = {"unbounded": [], "bounded": [], "pmf": []}
  for S in (1024, 4096, 16384, 65536):
    g = torch.Generator(device=dev).manual_seed(1)
    pick = np.linspace(0, S - 1, 16).astype(np.int64)
    pick_d = torch.as_tensor(pick, device=dev)
    total = S * n
    idx = torch.randint(0, 64, (S, n), generator=g, dtype=torch.int32, device=dev)
    sym = torch.floor(torch.randn((S, n), generator=g, device=dev) * table_d[idx.long()] + .5)
    sym = sym.to(torch.int32)
    out, nbytes = rc.unbounded_index_range_encode_batch(sym, idx, cdf, size, off, 16, 4, 0)
    dec = rc.unbounded_index_range_decode_batch(out, nbytes, idx, cdf, size, off, 16, 4, 0)
    assert torch.equal(dec, sym), "c4: unbounded round trip failed at %d streams" % S
    del dec
    sym_h, idx_h = sym[pick_d].cpu().numpy(), idx[pick_d].cpu().numpy()
    got = rc.strings_to_host(out[pick_d], nbytes[pick_d])
    for i in range(len(pick)):
      want = coder.ui_encode(sym_h[i], idx_h[i], c_h, s_h, o_h, 16, 4)
      assert got[i] == want, "c4: stream %d of %d differs from the %s coder" % (pick[i], S, coder.kind)
    checked += len(pick)
    cap = out.shape[1]
    enc_ms = timeit(lambda: rc.unbounded_index_range_encode_batch(
        sym, idx, cdf, size, off, 16, 4, 0, cap=cap))
    dec_ms = timeit(lambda: rc.unbounded_index_range_decode_batch(
        out, nbytes, idx, cdf, size, off, 16, 4, 0, check=False))
    rows_out["unbounded"].append({"streams": S, "encode_gsym_s": total / enc_ms / 1e6,
                                  "decode_gsym_s": total / dec_ms / 1e6,
                                  "encode_ms": enc_ms, "decode_ms": dec_ms})
    del idx, sym, out, nbytes
    u = torch.randn((S, n), generator=g, device=dev) * 6.0 + 31.5
    data = torch.clamp(torch.floor(u + .5), 0, 63).to(torch.int16)
    del u
    out, nbytes = rc.range_encode_batch(data, cdf64.reshape(1, 65), 16, debug_level=0)
    dec = rc.range_decode_batch(out, nbytes, (n,), cdf64.reshape(1, 65), 16, debug_level=0)
    assert torch.equal(dec, data), "c4: bounded round trip failed at %d streams" % S
    del dec
    data_h = data[pick_d].cpu().numpy()
    got = rc.strings_to_host(out[pick_d], nbytes[pick_d])
    for i in range(len(pick)):
      assert got[i] == coder.rc_encode(data_h[i], cdf64_h.reshape(1, 65), 16), \
          "c4: bounded stream %d of %d differs from the %s coder" % (pick[i], S, coder.kind)
    checked += len(pick)
    cap = out.shape[1]
    enc_ms = timeit(lambda: rc.range_encode_batch(data, cdf64.reshape(1, 65), 16, debug_level=0,
                                                  cap=cap))
    dec_ms = timeit(lambda: rc.range_decode_batch(out, nbytes, (n,), cdf64.reshape(1, 65), 16,
                                                  debug_level=0))
    rows_out["bounded"].append({"streams": S, "encode_gsym_s": total / enc_ms / 1e6,
                                "decode_gsym_s": total / dec_ms / 1e6,
                                "encode_ms": enc_ms, "decode_ms": dec_ms})
    del data, out, nbytes
  rows_checked = 0
  S = 65536
  g = torch.Generator(device=dev).manual_seed(2)
  for pn in (22, 64, 256, 1480):
    rec = {"rows": S, "n": pn}
    for kind in ("uniform", "gaussian"):
      if kind == "uniform":
        p = torch.rand((S, pn), generator=g, device=dev) + 1e-6
      else:
        sig = torch.rand((S, 1), generator=g, device=dev) * (pn / 6.0) + 0.2
        kk = torch.arange(pn, device=dev, dtype=torch.float32) - (pn - 1) / 2
        p = torch.exp(-0.5 * (kk / sig) ** 2)
      p = (p / p.sum(-1, keepdim=True)).contiguous()
      got = rc.pmf_to_quantized_cdf(p, 16)
      rows = np.linspace(0, S - 1, 32).astype(np.int64)
      rows_d = torch.as_tensor(rows, device=dev)
      want = coder.pmf_to_cdf(p[rows_d].cpu().numpy(), 16, threads=threads)
      assert np.array_equal(got[rows_d].cpu().numpy(), want), \
          "c4: pmf_to_quantized_cdf differs (%s, n=%d)" % (kind, pn)
      rows_checked += len(rows)
      ms = timeit(lambda: rc.pmf_to_quantized_cdf(p, 16), 2)
      rec["%s_ms" % kind] = ms
      rec["%s_mrows_s" % kind] = S / ms / 1e3
      del p, got
    rows_out["pmf"].append(rec)
  torch.cuda.empty_cache()
  return {"workload": "range coder sweep: S streams x 4096 symbols, precision 16, overflow_width 4, "
                      "64-row Gaussian table (unbounded) / m=64 shared CDF (bounded); "
                      "pmf_to_quantized_cdf on 65536 rows x n",
          "unbounded_index": rows_out["unbounded"], "bounded": rows_out["bounded"],
          "pmf_to_quantized_cdf": rows_out["pmf"], "bit_exact_streams_checked": checked,
          "bit_exact_rows_checked": rows_checked, "checker": coder.kind}


# ---- BASELINE config 1: bls2017 compress/decompress of [8,256,256,3] on the host cores ------------
def c1_block(dev):
  import torch
  from compression_b200.examples import bls2017
  from oracle import cpu_codec, oracle_lib
  coder = oracle_lib.load("best")
  threads = host_threads()
  model = bls2017.Bls2017Model(128, dev, seed=0)
  cpu = cpu_codec.CpuBls2017(model.export_weights(), coder, threads, tables=model.export_tables())
  g = torch.Generator().manual_seed(8)
  x = torch.rand((8, 256, 256, 3), generator=g, dtype=torch.float32)
  xn = x.numpy()
  times = []
  for it in range(4):
    t0 = time.perf_counter()
    c = cpu.compress(xn)
    x_hat_cpu, y_sym_cpu = cpu.decompress(c)
    if it:
      times.append(time.perf_counter() - t0)
  cpu_ms = float(np.mean(times)) * 1e3
  xd = x.to(dev)
  model.decompress(model.compress(xd))
  torch.cuda.synchronize()
  a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  a.record()
  for _ in range(10):
    cg = model.compress(xd)
    x_hat = model.decompress(cg)
  b.record()
  torch.cuda.synchronize()
  gpu_ms = a.elapsed_time(b) / 10
  # the GPU's strings are the CPU coder's strings of the GPU's symbols, and decode on the CPU
  host = model.to_host(cg)
  eb = model.entropy_bottleneck
  sym = eb._quantize_symbols(model.analysis_transform(xd)).cpu().numpy()
  idx = np.broadcast_to(np.arange(128, dtype=np.int32), sym.shape[1:]).copy()
  t = model.export_tables()
  ok = all(host.string[i] == coder.ui_encode(sym[i], idx, t["eb_cdf"], t["eb_len"], t["eb_off"],
                                             16, 4) for i in range(8))
  agree = float(np.mean(sym == c["y_sym"]))
  if not ok:
    raise SystemExit("c1: GPU strings differ from the CPU coder's")
  pix = 8 * 256 * 256
  out = {"workload": "bls2017-128 compress+decompress of [8,256,256,3] (BASELINE configs[0])",
         "cpu": {"ms": cpu_ms, "mpix_s": pix / cpu_ms / 1e3, "cores": threads,
                 "kind": "torch-CPU f32 transforms + %s range coder" % coder.kind},
         "gpu": {"ms": gpu_ms, "mpix_s": pix / gpu_ms / 1e3, "note": "one batch at a time, "
                 "device-resident, launch-latency bound at this size"},
         "strings_equal_cpu_coder": bool(ok),
         "symbols_agree_with_f32_cpu_transforms": agree}
  del model
  torch.cuda.empty_cache()
  return out


def pin_to_cores(local, world):
  """One contiguous slice of the host cores per rank: the ranks' issue threads and CUDA worker
  threads stop migrating over each other's cores."""
  try:
    cores = sorted(os.sched_getaffinity(0))
    per = len(cores) // max(1, world)
    if world > 1 and per >= 2:
      os.sched_setaffinity(0, set(cores[local * per:(local + 1) * per]))
      return per
  except (AttributeError, OSError):
    pass
  return None


# ---- B200 arm ----------------------------------------------------------------------------------------
def run_b200(args):
  world = int(os.environ.get("WORLD_SIZE", "1"))
  rank = int(os.environ.get("RANK", "0"))
  local = int(os.environ.get("LOCAL_RANK", "0"))
  pinned_cores = pin_to_cores(local, world)
  import torch
  import torch.distributed as dist
  import compression_b200 as tfc
  from compression_b200.examples import bmshj2018

  if not torch.cuda.is_available():
    raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference")
  torch.cuda.set_device(local)
  dev = torch.device("cuda", local)
  if world > 1:
    dist.init_process_group("nccl", device_id=dev)
  blocks = wanted_blocks(args)
  h, w, batch, nf = WORKLOADS[args.workload]
  if args.debug_bits:
    from compression_b200 import _lib as _tfc_lib
    _tfc_lib.check(_tfc_lib.lib().tfc_configure(99, args.debug_bits))
  if args.coder_lanes:
    tfc.configure(coder_lanes=args.coder_lanes)
  if args.conv_pair:
    tfc.configure(conv_pair=args.conv_pair)
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

  def allmax(v):
    t = torch.tensor([float(v)], device=dev, dtype=torch.float64)
    if world > 1:
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

  pixels_per_step = batch * h * w
  if args.depth <= 0:
    # ~230 bytes of activations per input pixel stay reserved per batch in flight (measured:
    # 15.2 GB for 32 x 1080p, 5.8 GB for 64 x 768x512); keep the pipeline inside 3/4 of HBM
    total_mem = torch.cuda.get_device_properties(dev).total_memory
    args.depth = max(1, min(20, int(0.85 * total_mem / (200.0 * pixels_per_step))))
  if args.coder_pack <= 0:
    # streams in flight / CTAs wanted: the coder is latency bound per stream, and 4 streams per
    # CTA is the measured optimum once there are more streams than ~4 per SM
    # (measured with the in-kernel occupancy trace, profiles/r2c_pipeline_gantt_coder_trace.jsonl:
    # coder SM-time per batch 4.1 ms of the GPU at 4 streams per CTA, 2.85 ms at 8)
    args.coder_pack = max(1, min(8, int(round(batch * args.depth / 150.0))))
  if args.host_slots <= 0:
    # e2e leg: twice as many sets of pinned staging buffers as batches in flight, so that neither
    # host wait (strings of batch k, x_hat of the batch whose buffers are reused) comes before the
    # GPU can have finished that work; at most 8 GB pinned per rank
    args.host_slots = max(args.depth + 1,
                          min(2 * args.depth, int(8e9 / (1.6 * 3 * pixels_per_step))))
  if args.back_lag <= 0:
    args.back_lag = max(1, min(args.depth, 12, args.host_slots // 2))
  pipe = model.pipeline(args.depth, args.tiles_per_cta, args.coder_pack, args.back_lag,
                        args.host_slots, conv_stages=args.conv_stages,
                        coder_smem_kb=args.coder_smem_kb, coder_table_kb=args.coder_table_kb,
                        tail=(tuple(int(v) for v in args.tail.split(",")) if args.tail else None),
                        stage_priority=(tuple(int(v) for v in args.stage_priority.split(","))
                                        if "," in args.stage_priority else bool(int(args.stage_priority))))

  with pipe:
    # every stream of the pipeline must have run once before the timed region: a stream's first
    # batch allocates its activation pool (cudaMalloc synchronises the device) and stages its
    # weight copies
    n_warm = max(args.warmup, 3, args.depth + 1)
    clocks = ClockSampler(local)
    if rank == 0:
      clocks.start()
    pipe.roundtrip_device([x_dev] * n_warm)
    barrier()
    # no cyclic garbage collection inside the timed windows: a generation-2 pass over the heap of
    # this process takes tens of ms, which is a visible share of a 0.27 s window when it lands
    # while the GPU's queue is still short
    pygc.collect()
    pygc.disable()
    clocks.mark()
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
    conv_ms = timer.total_ms()
    ms_per_step = allmax(start.elapsed_time(stop)) / args.steps

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
    e2e_ms = allmax(time.perf_counter() - t0) * 1e3 / args.steps
    assert seen["pixels"] == args.steps * pixels_per_step, "e2e: not every batch came back"
    string_bytes = seen["strings"]
    h2d_bytes, d2h_bytes = int(pipe.h2d_bytes), int(pipe.d2h_bytes)
    host_busy_ms = pipe.host_busy_s * 1e3 / args.steps
    host_wait_ms = pipe.host_wait_s * 1e3 / args.steps
    pygc.enable()

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
  del y, z, sigma, string, gc

  extra = {}
  if "parity" in blocks:
    extra["parity"] = parity_block(model, x_dev, args.workload) if rank == 0 else None
    barrier()
  pipe_back_lag, pipe_host_slots = pipe.back_lag, pipe.host_slots
  del pipe
  torch.cuda.empty_cache()          # per-stream activation pools of the C2 leg
  if "c3" in blocks:
    extra["c3"] = c3_block(args, model, world, rank, dev, barrier, allmax)
  torch.cuda.empty_cache()
  if "c5" in blocks:
    extra["c5"] = c5_block(args, world, rank, dev, barrier, allmax)
  if rank == 0:
    if "c4" in blocks:
      extra["c4"] = c4_block(dev)
    if "c1" in blocks:
      extra["c1"] = c1_block(dev)
  if rank != 0:
    if world > 1:
      dist.barrier()
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
  tf_isolated = flops_step / (conv_ms_isolated * 1e-3) / 1e12
  tf_step = flops_step / (ms_per_step * 1e-3) / 1e12
  traffic, traffic_file = ncu_conv_traffic(args.workload)
  sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
  line = {
      "metric": METRIC,
      "value": world * pixels / (ms_per_step * 1e-3) / 1e6, "unit": "MPix/s", "n_gpus": world,
      "steps": args.steps, "warmup": n_warm, "ms_per_step": ms_per_step,
      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
      "data": "synthetic",
      "config": {"workload": workload_name(args),
                 "weights": "random-init seed 0", "l2": "inputs_exceed_l2 (%.0f MB images)"
                 % (x_dev.numel() * 4 / 1e6), "parallelism": "batch-sharded x%d" % world,
                 "conv_impl": args.conv_impl, "conv_pair": args.conv_pair,
                 "pinned_cores_per_rank": pinned_cores,
                 "timed_window": "%d steps = %.2f pipeline waves of %d batches in flight"
                 % (args.steps, args.steps / float(args.depth), args.depth),
                 "pipeline": {"batches_in_flight": args.depth, "e2e_back_lag": pipe_back_lag,
                              "e2e_host_slots": pipe_host_slots,
                              "tiles_per_cta": args.tiles_per_cta if args.depth > 1 else 0,
                              "coder_streams_per_cta": args.coder_pack if args.depth > 1 else 0,
                              "conv_stages": args.conv_stages, "coder_smem_kb": args.coder_smem_kb,
                              "coder_table_kb": args.coder_table_kb,
                              "hbm_peak_allocated_gb": round(
                                  torch.cuda.max_memory_allocated(dev) / 1e9, 1)}},
      "e2e": {"value": world * pixels / (e2e_ms * 1e-3) / 1e6, "unit": "MPix/s",
              "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
              "string_bytes_per_step": int(string_bytes), "ms_per_step": e2e_ms,
              "host_busy_ms_per_step": host_busy_ms, "host_wait_ms_per_step": host_wait_ms,
              "image_dtype": "uint8 in, uint8 out (scaling and quantize_image fused into the "
                             "first / last transform kernel)"},
      "gpu_launches": int(launches),
      "clocks": clk,
      # The transforms dominate the GPU's SM-time (the coder holds a few CTAs for longer, see
      # "coder").  achieved / frac (= frac_burst): the conv launches of one batch timed per launch
      # with CUDA events, the batch alone on the GPU, against the BURST peak; frac_step: the same
      # FLOPs over the whole driver-timed step against the SUSTAINED peak; "in_pipeline": the
      # per-launch brackets inside the timed region, where they include waiting for shared SMs.
      "roofline": {"bound": "tensor", "kernel": "SignalConv2D kernels (%d launches/step)" % n_conv,
                   "achieved": tf_isolated, "peak": peaks["bf16_tflops"], "unit": "TFLOP/s",
                   "frac": tf_isolated / peaks["bf16_tflops"],
                   "frac_burst": tf_isolated / peaks["bf16_tflops"],
                   "achieved_step": tf_step, "peak_sustained": peaks["bf16_tflops_sustained"],
                   "frac_step": tf_step / peaks["bf16_tflops_sustained"],
                   "traffic": traffic, "traffic_source": traffic_file,
                   "peak_source": peaks["source"] + " bf16 (burst for the isolated launches, "
                                                    "sustained for the step)",
                   "ms_per_step": conv_ms_isolated, "flops_per_step": flops_step,
                   "in_pipeline": {"ms_per_step": conv_ms_step,
                                   "share_of_step": conv_ms_step / ms_per_step}},
      "coder": {"bound": "latency", "kernel": "encode/decode_unbounded_v4 (y, one CTA per image)",
                "symbols_per_stream": y_symbols, "streams": batch,
                "encode_ms": coder_ms[0], "decode_ms": coder_ms[1],
                "encode_cycles_per_symbol": coder_ms[0] * 1e-3 * sm_mhz * 1e6 / y_symbols,
                "decode_cycles_per_symbol": coder_ms[1] * 1e-3 * sm_mhz * 1e6 / y_symbols,
                "hbm_frac": batch * y_symbols * 12.6 / (sum(coder_ms) * 1e-3) / 1e9
                            / peaks["hbm_gbs"]},
  }
  for k, v in extra.items():
    if v is not None:
      line[k] = v
  if not args.no_cpu_baseline and world == 1:
    images = cpu_sample_images(args)
    mpix, cms, kind, cores = cpu_arm(h, w, images, 2, 1, weights=model.export_weights())
    line["cpu_baseline"] = {"value": mpix, "unit": "MPix/s", "cores": cores, "kind": kind,
                            "sample": "%d of %d images of %dx%d, 2 steps after 1 warm-up, all host threads; "
                                      "torch-CPU f32 transforms + %s range coder"
                                      % (images, batch, w, h, kind)}
  print(json.dumps(line))
  if world > 1:
    dist.barrier()
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
