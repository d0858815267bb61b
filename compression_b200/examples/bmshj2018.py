"""Scale-hyperprior codec — mirror of the model in `/root/reference/examples/bmshj2018.py`
(transforms :73-198, compress :285-362, decompress :365-408) over the B200 kernels.

The reference builds a TF graph per image and runs it in a Session; here `Bmshj2018Model`
holds the four transforms and the two entropy models and exposes
  compress(x)            [B,H,W,3] float32 in [0,1] on the device -> Compressed (strings in HBM)
  decompress(compressed) -> x_hat [B,H,W,3] float32 on the device
  pack / unpack          host-side strings (what PackedTensors would carry)
Activations between layers are bfloat16 (operands of the tensor-core convs); everything the
entropy models consume (y, z, sigma) is float32.
"""
import collections

import numpy as np
import torch

from compression_b200.python.layers import entropy_models
from compression_b200.python.layers import gdn
from compression_b200.python.layers import signal_conv
from compression_b200.python.ops import tensor_ops

SCALES_MIN = 0.11
SCALES_MAX = 256
SCALES_LEVELS = 64

Compressed = collections.namedtuple(
    "Compressed", ["string", "side_string", "x_shape", "y_shape", "z_shape"])


quantize_image = signal_conv.quantize_image     # bmshj2018.py:60-63


class _Sequential(object):

  def __init__(self, layers):
    self._layers = layers

  def build(self, channels, device, generator=None):
    for l in self._layers:
      l.build(channels, device, generator)
      channels = l.filters
    return self

  def __call__(self, tensor, out_dtype=torch.float32):
    n = len(self._layers)
    for i, layer in enumerate(self._layers):
      tensor = layer(tensor, out_dtype=out_dtype if i == n - 1 else torch.bfloat16)
    return tensor

  @property
  def layers(self):
    return self._layers


class AnalysisTransform(_Sequential):
  """bmshj2018.py:73-104."""

  def __init__(self, num_filters, impl=0):
    conv = lambda i, act: signal_conv.SignalConv2D(
        num_filters, (5, 5), name="layer_%d" % i, corr=True, strides_down=2,
        padding="same_zeros", use_bias=True, activation=act, impl=impl)
    super(AnalysisTransform, self).__init__(
        [conv(i, gdn.GDN(name="gdn_%d" % i)) for i in range(3)] + [conv(3, None)])


class SynthesisTransform(_Sequential):
  """bmshj2018.py:107-138."""

  def __init__(self, num_filters, impl=0):
    conv = lambda i, f, act: signal_conv.SignalConv2D(
        f, (5, 5), name="layer_%d" % i, corr=False, strides_up=2, padding="same_zeros",
        use_bias=True, activation=act, impl=impl)
    super(SynthesisTransform, self).__init__(
        [conv(i, num_filters, gdn.GDN(name="igdn_%d" % i, inverse=True)) for i in range(3)] +
        [conv(3, 3, None)])


class HyperAnalysisTransform(_Sequential):
  """bmshj2018.py:141-168."""

  def __init__(self, num_filters, impl=0):
    C = signal_conv.SignalConv2D
    super(HyperAnalysisTransform, self).__init__([
        C(num_filters, (3, 3), name="layer_0", corr=True, strides_down=1, padding="same_zeros",
          use_bias=True, activation=torch.relu, impl=impl),
        C(num_filters, (5, 5), name="layer_1", corr=True, strides_down=2, padding="same_zeros",
          use_bias=True, activation=torch.relu, impl=impl),
        C(num_filters, (5, 5), name="layer_2", corr=True, strides_down=2, padding="same_zeros",
          use_bias=False, activation=None, impl=impl),
    ])


class HyperSynthesisTransform(_Sequential):
  """bmshj2018.py:171-198."""

  def __init__(self, num_filters, impl=0):
    C = signal_conv.SignalConv2D
    super(HyperSynthesisTransform, self).__init__([
        C(num_filters, (5, 5), name="layer_0", corr=False, strides_up=2, padding="same_zeros",
          use_bias=True, kernel_parameterizer=None, activation=torch.relu, impl=impl),
        C(num_filters, (5, 5), name="layer_1", corr=False, strides_up=2, padding="same_zeros",
          use_bias=True, kernel_parameterizer=None, activation=torch.relu, impl=impl),
        C(num_filters, (3, 3), name="layer_2", corr=False, strides_up=1, padding="same_zeros",
          use_bias=True, kernel_parameterizer=None, activation=None, impl=impl),
    ])


class HostStrings(object):
  """The byte strings of one batch as they sit in host memory after the device-to-host copy: a
  [S, cap] uint8 slab and the S lengths.  Behaves like a list of `bytes` (materialised on
  access); `total_bytes` needs no copy."""

  def __init__(self, slab, nbytes):
    self._slab, self._nbytes = slab.numpy(), nbytes.numpy()

  def __len__(self):
    return len(self._nbytes)

  def __getitem__(self, i):
    if isinstance(i, slice):
      return [self[j] for j in range(*i.indices(len(self)))]
    if i < 0:
      i += len(self)
    return self._slab[i, :self._nbytes[i]].tobytes()

  def __iter__(self):
    return (self[i] for i in range(len(self)))

  def __eq__(self, other):
    return list(self) == list(other)

  @property
  def total_bytes(self):
    return int(self._nbytes.sum())


class Pipeline(object):
  """Keeps several batches in flight on separate CUDA streams.

  One image = one serial range-coder chain (the reference's string format), so the coder
  kernels are latency bound and occupy a few SMs for milliseconds while the transforms are
  throughput bound.  The pipeline runs batch k's coder concurrently with batch k+1's
  transforms: every batch gets its own stream (its own kernels stay ordered, no events
  needed), the coder packs `coder_streams_per_cta` images into one CTA and the transform
  kernels switch from a persistent grid to `tiles_per_cta` tiles per CTA so that the hardware
  scheduler fills whatever SMs the coder leaves free.  Results are those of compress()/decompress()
  called batch by batch.
  """

  def __init__(self, model, depth=3, tiles_per_cta=8, coder_streams_per_cta=4, back_lag=None,
               host_slots=None, conv_stages=0, coder_smem_kb=0, coder_table_kb=0,
               stage_streams=True, coder_stream_cap=22, stage_priority=True, tail=None):
    """`stage_streams` (default): the TRANSFORM phases of all batches share a few "stage" streams
    (3 for analysis, 2 for the decoder's hyper-synthesis, 3 for synthesis) and only the CODER
    phases get one stream per batch in flight.  torch's caching allocator keeps one pool per
    stream, and the big activations (0.46 GB per 1080p image) are transform-phase temporaries: with
    a stream per batch the memory reserved grows with the batches in flight (134 GB at 18 x 16
    1080p images), with stage streams it stays at ~8 pools however many batches the latency-bound
    coder needs in flight.  Tensors that cross streams are handed over with events +
    `record_stream`.  `stage_streams=False`: one stream per batch for everything.
    `tail` = (batches, coder_streams_per_cta): the coder calls of the LAST `batches` batches of a
    roundtrip_* call use this packing instead — packing trades a stream's latency for SMs, which
    is the wrong trade for the batches whose coder chain is the end of the job (nothing is left
    to use the SMs it saves).

    The stage streams come from torch's HIGH-priority pool and the coder streams from the default
    one: each pool holds 32 streams and hands them out round-robin, so more than 32 draws from
    one pool alias (a coder stream would silently BE a stage stream, and a 100 ms coder kernel
    would sit in front of another batch's convolutions); the device also has only 32 hardware
    queues (CUDA_DEVICE_MAX_CONNECTIONS), hence `coder_stream_cap`: batches k and k + cap share a
    coder stream.  The priority also makes the block scheduler place waiting transform CTAs before
    waiting coder CTAs.

    `conv_stages`, `coder_smem_kb`, `coder_table_kb` (0: library defaults) size the transform
    and coder CTAs so that they can share an SM: a 3-stage implicit-GEMM CTA takes 172 KB of the
    SM's 228 KB of shared memory and 54 K of its 64 K registers, which leaves room for ONE
    one-stream coder CTA of at most ~54 KB (`coder_streams_per_cta=1`, tables beyond
    `coder_table_kb` read through L1) — the latency-bound coder of batch k then runs in the idle
    issue slots of the SMs that convolve batch k+1 instead of taking SMs away from them."""
    from compression_b200 import _lib
    self.model = model
    self.depth = max(1, int(depth))
    # roundtrip_host: how many batches later the host comes back for a batch's strings
    # and how many sets of pinned staging buffers it cycles through
    self.host_slots = max(self.depth + 1, int(host_slots)) if host_slots else self.depth + 1
    self.back_lag = (min(self.host_slots - 1, max(1, int(back_lag))) if back_lag
                     else max(1, self.host_slots // 2))
    n_coder = self.depth
    if stage_streams and self.depth > 1 and coder_stream_cap:
      n_coder = min(self.depth, max(1, int(coder_stream_cap)))
    self.streams = [torch.cuda.Stream(device=model.device) for _ in range(n_coder)]
    self.stage = None
    if stage_streams and self.depth > 1:
      # stage_priority: True = all transform stages one level above the coder streams; a triple
      # gives the levels of (analysis, hyper-synthesis, synthesis) separately (0 = default pool,
      # -1, -2, -3 = higher; torch keeps one pool of 32 streams per level)
      if isinstance(stage_priority, (tuple, list)):
        pa, ph, ps = (int(v) for v in stage_priority)
      else:
        pa = ph = ps = -1 if stage_priority else 0
      mk = lambda n, prio: [torch.cuda.Stream(device=model.device, priority=prio)
                            for _ in range(n)]
      self.stage = {"analysis": mk(3, pa), "hyper": mk(2, ph), "synthesis": mk(3, ps)}
    self._lib = _lib
    self.tiles_per_cta = int(tiles_per_cta) if self.depth > 1 else 0
    self.pack = int(coder_streams_per_cta) if self.depth > 1 else 0
    self.tail = (int(tail[0]), int(tail[1])) if tail and self.depth > 1 else None
    self._pack_now = None
    self.share = (int(conv_stages), int(coder_smem_kb), int(coder_table_kb)) if self.depth > 1 \
        else (0, 0, 0)
    self.h2d_bytes = self.d2h_bytes = 0
    self.host_busy_s = self.host_wait_s = 0.0
    self._cap = None          # (string, side string) slab bytes per image, learnt from traffic

  def __enter__(self):
    if self.depth > 1:
      self._lib.configure(coder_streams_per_cta=self.pack, tiles_per_cta=self.tiles_per_cta,
                          conv_stages=self.share[0], coder_smem_kb=self.share[1],
                          coder_table_kb=self.share[2])
      self._pack_now = self.pack
    return self

  def __exit__(self, *exc):
    self._lib.configure(coder_streams_per_cta=0, tiles_per_cta=0, conv_stages=0, coder_smem_kb=0,
                        coder_table_kb=0)
    return False

  def _pack_for(self, k, n):
    """Coder packing for batch k of n (the library reads it at launch time)."""
    want = self.tail[1] if self.tail and k >= n - self.tail[0] else self.pack
    if want != self._pack_now:
      self._lib.configure(coder_streams_per_cta=want)
      self._pack_now = want

  def _all_streams(self):
    extra = [s for group in self.stage.values() for s in group] if self.stage else []
    return self.streams + extra

  def _fork(self):
    cur = torch.cuda.current_stream(self.model.device)
    for s in self._all_streams():
      s.wait_stream(cur)
    return cur

  def _join(self, cur):
    for s in self._all_streams():
      cur.wait_stream(s)

  def _streams_of(self, k):
    """(analysis, coder, hyper-synthesis, synthesis) streams of batch k."""
    e = self.streams[k % len(self.streams)]
    if not self.stage:
      return e, e, e, e
    g = self.stage
    return (g["analysis"][k % len(g["analysis"])], e, g["hyper"][k % len(g["hyper"])],
            g["synthesis"][k % len(g["synthesis"])])

  @staticmethod
  def _hand(src, dst, *tensors):
    """Hand tensors produced on stream `src` to stream `dst`: dst waits for what src has queued so
    far, and the allocator is told that dst uses the memory (it must not be given to another
    tensor of src's pool before dst is done with it)."""
    if src is dst:
      return
    ev = torch.cuda.Event()
    ev.record(src)
    dst.wait_event(ev)
    for t in tensors:
      if isinstance(t, (tuple, list)):
        for u in t:
          u.record_stream(dst)
      elif t is not None:
        t.record_stream(dst)

  def roundtrip_device(self, batches):
    """batches: device float32 [B,H,W,3] tensors -> list of x_hat (device).  The coder's status
    words of batch k are looked at 2 * depth batches later (long after it finished) and the rest
    on return."""
    from compression_b200.python.ops import range_coding_ops as rc
    model, outs = self.model, []
    cur = self._fork()
    done = {}
    with rc.deferred_checks() as checks:
      for k, x in enumerate(batches):
        checks.tag = k
        if not self.stage:
          st = self.streams[k % self.depth]
          with torch.cuda.stream(st):
            outs.append(model.decompress(model.compress(x, self._stagger(k, st))))
          checks.drain(2 * self.depth)
          continue
        a, e, hs, sy = self._streams_of(k)
        self._pack_for(k, len(batches))
        if k - self.depth in done:            # at most `depth` batches between analysis and x_hat
          a.wait_event(done.pop(k - self.depth))
        with torch.cuda.stream(a):
          y, z, sigma = model.analyse(x)
        self._hand(a, e, y, z, sigma)
        with torch.cuda.stream(e):
          string, side_string = model.encode_strings(y, z, sigma)
          z_hat = model.decode_z(side_string, tuple(z.shape[1:-1]))
        self._hand(e, hs, z_hat)
        with torch.cuda.stream(hs):
          sigma2 = model.hyper_sigma(z_hat, tuple(y.shape[1:-1]))
        self._hand(hs, e, sigma2)
        with torch.cuda.stream(e):
          y_hat = model.decode_y(string, sigma2)
        self._hand(e, sy, y_hat)
        with torch.cuda.stream(sy):
          x_hat = model.synthesize(y_hat, tuple(x.shape[1:-1]))
          ev = done[k] = torch.cuda.Event()
          ev.record(sy)
        x_hat.record_stream(cur)
        outs.append(x_hat)
        del y, z, sigma, z_hat, sigma2, y_hat, string, side_string
        checks.drain(2 * self.depth)
      self._join(cur)
    return outs

  def _stagger(self, k, st):
    """First wave only: batch k starts its transforms when batch k-1 has finished its own, so
    the `depth` batches that start together do not run transforms, then coders, in lockstep
    (the GPU is transform-bound during that time either way).  Later batches inherit the
    stagger from their predecessors on the same stream."""
    if k == 0:
      self._wave_ev = None
    if k >= self.depth:
      return None
    if self._wave_ev is not None:
      st.wait_event(self._wave_ev)
    ev = self._wave_ev = torch.cuda.Event()
    return lambda: ev.record(st)

  def roundtrip_host(self, batches_u8, on_result=None):
    """batches_u8: pinned host uint8 [B,H,W,3] tensors.  Per batch: image H2D (the bytes
    themselves: the first analysis layer reads uint8), compress, strings D2H, strings H2D,
    decompress (the last synthesis layer writes the quantised uint8 image), x_hat D2H.  The host
    waits twice per batch: for batch k's strings after it has queued the compress of the next
    `back_lag` batches, and for its x_hat when the slot is reused, host_slots - back_lag batches
    after the decompress was queued.  Both waits have to span several batch times or the host,
    not the GPU, sets the pace (a decompress is one serial decode per image and takes longer
    than two batch times).

    The coder's status words travel with the strings: a batch whose encode overflowed its slab
    is re-run with the safe slab size before its decode is queued, and a batch is delivered only
    after its decode status has been looked at.  String slabs are sized from the traffic seen
    (1.25 x the longest string so far) instead of the 2-bytes-per-symbol bound, which halves the
    string bytes that cross PCIe.

    Pinned staging buffers are allocated once per pipeline slot and reused (page-locking
    memory synchronises the device).  Results are delivered in order through
    `on_result(k, compressed, x_hat_u8)`: `compressed.string` / `.side_string` are HostStrings
    (sequences of bytes over the slot's pinned slab) and `x_hat_u8` is a view of the slot's
    pinned buffer, all valid during the callback.  Without a callback the results are copied
    out (strings as lists of bytes) and returned as a list."""
    import time
    from compression_b200.python.ops import range_coding_ops as rc
    model, dev = self.model, self.model.device
    n, nslots = len(batches_u8), self.host_slots
    slots = getattr(self, "_slots", None)
    if slots is None or len(slots) != nslots:
      slots = self._slots = [dict() for _ in range(nslots)]
    out_list = [] if on_result is None else None
    self.h2d_bytes = self.d2h_bytes = 0     # per batch, counted from the tensors copied
    waited = [0.0]
    t_start = time.perf_counter()
    cur = self._fork()

    def wait(ev):
      t0 = time.perf_counter()
      ev.synchronize()
      waited[0] += time.perf_counter() - t0

    def pinned(slot, name, shape, dtype):
      # a view of the slot's page-locked storage for `name`, grown (never shrunk) on demand:
      # the slab sizes settle after the first batches and no page-locking happens after that
      need = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
      store = slot.get(name)
      if store is None or store.numel() < need:
        store = slot[name] = torch.empty((max(need, 16),), dtype=torch.uint8, pin_memory=True)
      return store[:need].view(dtype).view(tuple(shape))

    def deliver(slot, checks):
      k = slot.pop("pending", None)
      if k is None:
        return
      wait(slot["done"])
      checks.verify(("back", k))        # corrupt string: raises before anything is delivered
      if on_result is not None:
        on_result(k, slot["host"], slot["x_hat_view"])
      else:
        h = slot["host"]
        out_list.append((Compressed(list(h.string), list(h.side_string), h.x_shape, h.y_shape,
                                    h.z_shape), slot["x_hat_view"].clone()))

    def front_work(k, slot, st, checks, cap):
      checks.tag = ("front", k)
      self._pack_for(k, n)
      a, e, _, _ = self._streams_of(k)
      with torch.cuda.stream(a):
        x = batches_u8[k].to(dev, non_blocking=True)
        if self.stage:
          y, z, sigma = model.analyse(x)
        else:
          c = model.compress(x, self._stagger(k, st) if cap is not False else None,
                             cap=cap or None)
      if self.stage:
        self._hand(a, e, y, z, sigma)
      with torch.cuda.stream(e):
        if self.stage:
          string, side_string = model.encode_strings(y, z, sigma, cap or None)
          c = Compressed(string, side_string, tuple(x.shape[1:-1]), tuple(y.shape[1:-1]),
                         tuple(z.shape[1:-1]))
        staged = []
        for name, (slab, nbytes) in (("y", c.string), ("z", c.side_string)):
          hs = pinned(slot, name + "_slab", slab.shape, torch.uint8)
          hn = pinned(slot, name + "_len", nbytes.shape, torch.int32)
          hs.copy_(slab, non_blocking=True)
          hn.copy_(nbytes, non_blocking=True)
          staged.append((hs, hn))
        ev = torch.cuda.Event()
        ev.record(st)
      return c, staged, ev

    def issue_front(k, checks):
      slot, st = slots[k % nslots], self._streams_of(k)[1]
      deliver(slot, checks)              # the slot's previous batch must be out of its buffers
      slot["front"] = front_work(k, slot, st, checks, self._cap)
      if k == 0:
        self.h2d_bytes += batches_u8[k].numel()

    def issue_back(k, checks):
      slot, st = slots[k % nslots], self._streams_of(k)[1]
      c, staged, ev = slot.pop("front")
      wait(ev)                           # the strings and their status are in host memory now
      try:
        checks.verify(("front", k))
      except self._lib.CudaError:
        # a string outgrew the learnt slab size: once more with the safe default (rare)
        self._cap = None
        c, staged, ev = front_work(k, slot, st, checks, False)
        wait(ev)
        checks.verify(("front", k))
      longest = (int(staged[0][1].max()), int(staged[1][1].max()))
      if self._cap is None or longest[0] > self._cap[0] / 1.125 or longest[1] > self._cap[1] / 1.125:
        want = tuple(int(1.25 * v) + 64 + (-(int(1.25 * v) + 64)) % 16 for v in longest)
        self._cap = want if self._cap is None else (max(want[0], self._cap[0]),
                                                    max(want[1], self._cap[1]))
      host = Compressed(HostStrings(*staged[0]), HostStrings(*staged[1]), c.x_shape, c.y_shape,
                        c.z_shape)
      checks.tag = ("back", k)
      self._pack_for(k, n)
      _, e, hstream, sy = self._streams_of(k)
      with torch.cuda.stream(e):
        # back to the device straight from the pinned slabs the strings arrived in
        dev_strings = []
        for hs, hn in staged:
          dev_strings.append((hs.to(dev, non_blocking=True), hn.to(dev, non_blocking=True)))
        dc = Compressed(dev_strings[0], dev_strings[1], c.x_shape, c.y_shape, c.z_shape)
        if not self.stage:
          x_hat = model.decompress(dc, out_dtype=torch.uint8)
        else:
          z_hat = model.decode_z(dc.side_string, dc.z_shape)
      if self.stage:
        self._hand(e, hstream, z_hat)
        with torch.cuda.stream(hstream):
          sigma2 = model.hyper_sigma(z_hat, dc.y_shape)
        self._hand(hstream, e, sigma2)
        with torch.cuda.stream(e):
          y_hat = model.decode_y(dc.string, sigma2)
        self._hand(e, sy, y_hat)
      with torch.cuda.stream(sy):
        if self.stage:
          x_hat = model.synthesize(y_hat, dc.x_shape, torch.uint8)
        out = pinned(slot, "x_hat", x_hat.shape, torch.uint8)
        out.copy_(x_hat, non_blocking=True)
        done = torch.cuda.Event()
        done.record(sy)
      slot["host"], slot["done"], slot["pending"], slot["x_hat_view"] = host, done, k, out
      if k == n - 1:                     # bytes per batch at the steady-state slab size
        strings = sum(hs.numel() + 4 * hn.numel() for hs, hn in staged)
        self.d2h_bytes = strings + out.numel()
        self.h2d_bytes = batches_u8[k].numel() + strings

    with rc.deferred_checks() as checks:
      for k in range(n + self.back_lag):
        if k < n:
          issue_front(k, checks)
        j = k - self.back_lag
        if 0 <= j < n:
          issue_back(j, checks)
      self._join(cur)
      first = max(0, n - nslots)
      for k in range(first, n):           # remaining deliveries, oldest first
        deliver(slots[k % nslots], checks)
    cur.synchronize()
    total = time.perf_counter() - t_start
    self.host_wait_s, self.host_busy_s = waited[0], total - waited[0]
    return out_list


class Bmshj2018Model(object):
  """Random-init (or externally loaded) hyperprior model on one GPU."""

  def __init__(self, num_filters=192, device=None, seed=0, impl=0):
    self.num_filters = int(num_filters)
    self.device = torch.device(device) if device is not None else torch.device(
        "cuda", torch.cuda.current_device())
    g = torch.Generator().manual_seed(seed)
    nf = self.num_filters
    self.analysis_transform = AnalysisTransform(nf, impl).build(3, self.device, g)
    self.synthesis_transform = SynthesisTransform(nf, impl).build(nf, self.device, g)
    self.hyper_analysis_transform = HyperAnalysisTransform(nf, impl).build(nf, self.device, g)
    self.hyper_synthesis_transform = HyperSynthesisTransform(nf, impl).build(nf, self.device, g)
    self.entropy_bottleneck = entropy_models.EntropyBottleneck()
    self.entropy_bottleneck.build(nf, self.device, g)
    # the reference requires the update op to have run before compressing
    # (docs/entropy_bottleneck.md:121-124)
    self.entropy_bottleneck.update()
    self.scale_table = np.exp(np.linspace(np.log(SCALES_MIN), np.log(SCALES_MAX), SCALES_LEVELS))
    self._gc_tables = entropy_models.GaussianConditional(None, self.scale_table)
    self._gc_tables.build_tables(self.device)

  def _conditional(self, sigma):
    gc = entropy_models.GaussianConditional(sigma, self.scale_table)
    gc.adopt_tables(self._gc_tables)
    return gc

  # ---- bmshj2018.py:301-312 ----------------------------------------------------------------
  def analyse(self, x):
    """x -> (y, z, sigma): everything compress() needs before entropy coding."""
    y = self.analysis_transform(x)
    z = self.hyper_analysis_transform(tensor_ops.abs_to_bf16(y))
    eb = self.entropy_bottleneck
    z_sym = eb._quantize_symbols(z)
    z_hat = eb._dequantize(z_sym, torch.bfloat16)
    sigma = self.hyper_synthesis_transform(z_hat)
    sigma = sigma[:, :y.shape[1], :y.shape[2], :].contiguous()
    return y, z, sigma

  # ---- the phases of compress / decompress, so that a pipeline can put the transform phases and
  # the coder phases of a batch on different CUDA streams (see Pipeline) ----------------------
  def encode_strings(self, y, z, sigma, cap=None):
    """entropy coding of compress(): (y, z, sigma) -> (string, side_string) on the device."""
    cap_y, cap_z = cap if cap is not None else (None, None)
    side_string = self.entropy_bottleneck.compress_device(z, cap=cap_z)
    string = self._conditional(sigma).compress_device(y, cap=cap_y)
    return string, side_string

  def decode_z(self, side_string, z_shape):
    nf = self.num_filters
    return self.entropy_bottleneck.decompress_device(
        side_string, tuple(z_shape) + (nf,), channels=nf, out_dtype=torch.bfloat16)

  def hyper_sigma(self, z_hat, y_shape):
    sigma = self.hyper_synthesis_transform(z_hat)
    return sigma[:, :y_shape[0], :y_shape[1], :].contiguous()

  def decode_y(self, string, sigma):
    return self._conditional(sigma).decompress_device(string, out_dtype=torch.bfloat16)

  def synthesize(self, y_hat, x_shape, out_dtype=torch.float32):
    x_hat = self.synthesis_transform(y_hat, out_dtype=out_dtype)
    return x_hat[:, :x_shape[0], :x_shape[1], :]

  def compress(self, x, after_analysis=None, cap=None):
    """x: float32 [B,H,W,3] in [0,1], or the uint8 image itself (a byte b stands for b / 255 as
    in read_png, bmshj2018.py:53-57; the first analysis layer reads the bytes directly).
    `after_analysis`: called once the transforms are queued and before the entropy coder is
    (the pipeline records an event there to stagger its first wave of batches).  `cap`:
    (bytes per image of the string slab, of the side-string slab) or None for the safe default
    of 2 bytes per symbol; a too small slab is reported through the coder's status words."""
    x = x.to(self.device)
    y, z, sigma = self.analyse(x)
    if after_analysis is not None:
      after_analysis()
    string, side_string = self.encode_strings(y, z, sigma, cap)
    return Compressed(string, side_string, tuple(x.shape[1:-1]), tuple(y.shape[1:-1]),
                      tuple(z.shape[1:-1]))

  # ---- bmshj2018.py:383-399 ----------------------------------------------------------------
  def decompress(self, c, out_dtype=torch.float32):
    """-> x_hat [B,H,W,3]: float32, or with out_dtype=torch.uint8 the quantised image
    (quantize_image, bmshj2018.py:60-63, fused into the last synthesis layer)."""
    z_hat = self.decode_z(c.side_string, c.z_shape)
    sigma = self.hyper_sigma(z_hat, c.y_shape)
    y_hat = self.decode_y(c.string, sigma)
    return self.synthesize(y_hat, c.x_shape, out_dtype)

  # ---- verbose branch of compress (bmshj2018.py:314-333) ----------------------------------
  def evaluate(self, x):
    """Returns dict with x_hat, likelihoods and the estimated bits per pixel."""
    y, z, sigma = self.analyse(x)
    _, z_lik = self.entropy_bottleneck(z, training=False)
    y_hat, y_lik = self._conditional(sigma)(y, training=False)
    x_hat = self.synthesis_transform(y_hat.to(torch.bfloat16))
    x_hat = x_hat[:, :x.shape[1], :x.shape[2], :]
    num_pixels = float(np.prod(x.shape[:-1]))
    bpp = (torch.log(y_lik).sum() + torch.log(z_lik).sum()) / (-np.log(2) * num_pixels)
    return dict(x_hat=x_hat, y=y, z=z, sigma=sigma, y_hat=y_hat, y_likelihoods=y_lik,
                z_likelihoods=z_lik, eval_bpp=float(bpp))

  def pipeline(self, depth=3, tiles_per_cta=8, coder_streams_per_cta=4, back_lag=None,
               host_slots=None, **share):
    return Pipeline(self, depth, tiles_per_cta, coder_streams_per_cta, back_lag, host_slots,
                    **share)

  # ---- host-side strings --------------------------------------------------------------------
  @staticmethod
  def to_host(c):
    from compression_b200.python.ops import range_coding_ops as rc
    return Compressed(rc.strings_to_host(*c.string), rc.strings_to_host(*c.side_string),
                      c.x_shape, c.y_shape, c.z_shape)

  def to_device(self, c):
    from compression_b200.python.ops import range_coding_ops as rc
    if isinstance(c.string, tuple):
      return c
    return Compressed(rc.strings_to_device(c.string, self.device),
                      rc.strings_to_device(c.side_string, self.device),
                      c.x_shape, c.y_shape, c.z_shape)

  def conv_layers(self):
    return (self.analysis_transform.layers + self.synthesis_transform.layers +
            self.hyper_analysis_transform.layers + self.hyper_synthesis_transform.layers)

  # ---- .tfci files (bmshj2018.py:339-347 write, :369-377 read): one image per file ---------
  @staticmethod
  def pack_tfci(c, image=0, model_id=None):
    """Host-side Compressed -> bytes of the reference's container for image `image`: tensors
    [string, side_string, x_shape, y_shape, z_shape], each of the strings a length-1 vector."""
    from compression_b200.python.util.packed_tensors import PackedTensors
    packed = PackedTensors()
    if model_id is not None:
      packed.model = model_id
    packed.pack(["string", "string", "int32", "int32", "int32"],
                [[c.string[image]], [c.side_string[image]], list(c.x_shape), list(c.y_shape),
                 list(c.z_shape)])
    return packed.string

  @staticmethod
  def unpack_tfci(data):
    """bytes of a .tfci file -> host-side Compressed holding that one image."""
    from compression_b200.python.util.packed_tensors import PackedTensors
    s, ss, xs, ys, zs = PackedTensors(data).unpack(["string", "string", "int32", "int32", "int32"])
    return Compressed([bytes(s[0])], [bytes(ss[0])], tuple(int(v) for v in xs),
                      tuple(int(v) for v in ys), tuple(int(v) for v in zs))

  # ---- export (host copies, for checkpoints and for the CPU checker in tests/bench) --------
  def export_weights(self):
    """Effective float32 weights keyed like oracle/cpu_codec.py expects."""
    w = {}
    for prefix, seq in (("a", self.analysis_transform), ("s", self.synthesis_transform),
                        ("ha", self.hyper_analysis_transform),
                        ("hs", self.hyper_synthesis_transform)):
      for i, l in enumerate(seq.layers):
        name = "%s%d" % (prefix, i)
        w[name + "/kernel"] = l.kernel.float().cpu().numpy()
        if l.bias is not None:
          w[name + "/bias"] = l.bias.float().cpu().numpy()
        if isinstance(l.activation, gdn.GDN):
          beta, gamma = l.activation.effective()
          w[name + "/beta"] = beta.cpu().numpy()
          w[name + "/gamma"] = gamma.cpu().numpy()
    eb = self.entropy_bottleneck
    for i, m in enumerate(eb.matrices):
      w["eb/matrix_%d" % i] = m.cpu().numpy()
      w["eb/bias_%d" % i] = eb.biases[i].cpu().numpy()
    for i, f in enumerate(eb.factors):
      w["eb/factor_%d" % i] = f.cpu().numpy()
    w["eb/quantiles"] = eb.quantiles.cpu().numpy()
    return w

  def export_tables(self):
    eb, gc = self.entropy_bottleneck, self._gc_tables
    return dict(eb_cdf=eb._quantized_cdf.cpu().numpy(), eb_len=eb._cdf_length.cpu().numpy(),
                eb_off=eb._offset.cpu().numpy(), gc_cdf=gc._quantized_cdf.cpu().numpy(),
                gc_len=gc._cdf_length.cpu().numpy(), gc_off=gc._offset.cpu().numpy())


def _trainer_base():
  from compression_b200.examples import bls2017
  return bls2017.Bls2017Trainer


class Bmshj2018Trainer(_trainer_base()):
  """One synchronous data-parallel training step of the hyperprior model
  (`/root/reference/examples/bmshj2018.py:201-282`), on the machinery of
  `bls2017.Bls2017Trainer`: all trainables in one flat bucket (+ the quantiles in a second),
  conv gradients through the CUDA kernels, one all-reduce per bucket, fused TF-Adam updates,
  CDF table rebuild per rank.  The Gaussian likelihood and its gradient
  (entropy_models.py:916-932, :965-969, with the lower bounds on the scale, :822-825, and on the
  likelihood, :181-184) are one CUDA kernel each way (`_GaussianLikelihood`)."""

  def _gaussian_likelihood(self, y_tilde, sigma):
    gc = self.model._gc_tables
    return entropy_models._GaussianLikelihood.apply(
        y_tilde.float(), sigma.float(), gc._effective_bound(), max(gc.likelihood_bound, 0.0))

  def losses(self, x, noise=None):
    """x float32 [B,H,W,3] in [0,1]; noise = (noise_y, noise_z) or None for fresh U(-.5,.5)
    -> dict(loss, bpp, mse, aux)."""
    model = self.model
    num_pixels = float(np.prod(x.shape[:-1]))
    y = self._transform(model.analysis_transform, x)
    z = self._transform(model.hyper_analysis_transform, torch.abs(y))
    noise_y, noise_z = noise if noise is not None else (None, None)
    z_tilde = z + (noise_z if noise_z is not None else torch.rand_like(z) - .5)
    z_lik = self._likelihood(z_tilde)
    sigma = self._transform(model.hyper_synthesis_transform, z_tilde)
    sigma = sigma[:, :y.shape[1], :y.shape[2], :]
    y_tilde = y + (noise_y if noise_y is not None else torch.rand_like(y) - .5)
    y_lik = self._gaussian_likelihood(y_tilde, sigma)
    x_tilde = self._transform(model.synthesis_transform, y_tilde)
    x_tilde = x_tilde[:, :x.shape[1], :x.shape[2], :]
    bpp = (torch.log(y_lik).sum() + torch.log(z_lik).sum()) / (-np.log(2) * num_pixels)
    mse = torch.mean((x - x_tilde) ** 2) * 255 ** 2
    return dict(loss=self.lmbda * mse + bpp, bpp=bpp, mse=mse,
                aux=model.entropy_bottleneck.losses[0], x_tilde=x_tilde)
