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


def quantize_image(image):
  """bmshj2018.py:60-63."""
  return torch.clamp(torch.round(image * 255), 0, 255).to(torch.uint8)


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


class Pipeline(object):
  """Keeps several batches in flight on separate CUDA streams.

  One image = one serial range-coder chain (the reference's string format), so the coder
  kernels are latency bound and occupy a few SMs for milliseconds while the transforms are
  throughput bound.  The pipeline runs batch k's coder concurrently with batch k+1's
  transforms: every batch gets its own stream (its own kernels stay ordered, no events
  needed), the coder packs `coder_streams_per_cta` images into one CTA and the persistent
  transform kernels leave `coder_sms` SMs to it.  Results are those of compress()/decompress()
  called batch by batch.
  """

  def __init__(self, model, depth=3, coder_sms=32, coder_streams_per_cta=4):
    from compression_b200 import _lib
    self.model = model
    self.depth = max(1, int(depth))
    self.streams = [torch.cuda.Stream(device=model.device) for _ in range(self.depth)]
    self._lib = _lib
    self._sms = torch.cuda.get_device_properties(model.device).multi_processor_count
    self.coder_sms = int(coder_sms) if self.depth > 1 else 0
    self.pack = int(coder_streams_per_cta) if self.depth > 1 else 0

  def __enter__(self):
    if self.depth > 1:
      self._lib.configure(sm_budget=max(1, self._sms - self.coder_sms),
                          coder_streams_per_cta=self.pack)
    return self

  def __exit__(self, *exc):
    self._lib.configure(sm_budget=0, coder_streams_per_cta=0)
    return False

  def _fork(self):
    cur = torch.cuda.current_stream(self.model.device)
    for s in self.streams:
      s.wait_stream(cur)
    return cur

  def _join(self, cur):
    for s in self.streams:
      cur.wait_stream(s)

  def roundtrip_device(self, batches):
    """batches: device float32 [B,H,W,3] tensors -> list of x_hat (device)."""
    from compression_b200.python.ops import range_coding_ops as rc
    model, outs = self.model, []
    cur = self._fork()
    with rc.deferred_checks():
      for k, x in enumerate(batches):
        with torch.cuda.stream(self.streams[k % self.depth]):
          outs.append(model.decompress(model.compress(x)))
      self._join(cur)
    return outs

  def roundtrip_host(self, batches_u8):
    """batches_u8: pinned host uint8 [B,H,W,3] -> list of (Compressed with host strings,
    x_hat uint8 on the host).  Per batch: image H2D, compress, strings D2H, strings H2D,
    decompress, x_hat D2H; the host only waits for batch k's strings after it has queued
    batch k+depth-1's compress."""
    from compression_b200.python.ops import range_coding_ops as rc
    model, dev = self.model, self.model.device
    front, results = {}, [None] * len(batches_u8)
    self.h2d_bytes = self.d2h_bytes = 0     # per batch, counted from the tensors copied
    cur = self._fork()

    def issue_front(k):
      st = self.streams[k % self.depth]
      with torch.cuda.stream(st):
        x = batches_u8[k].to(dev, non_blocking=True).float() / 255
        c = model.compress(x)
        pinned = []
        for slab, nbytes in (c.string, c.side_string):
          hs = torch.empty(slab.shape, dtype=torch.uint8, pin_memory=True)
          hn = torch.empty(nbytes.shape, dtype=torch.int32, pin_memory=True)
          hs.copy_(slab, non_blocking=True)
          hn.copy_(nbytes, non_blocking=True)
          pinned.append((hs, hn))
        if k == 0:
          self.h2d_bytes += batches_u8[k].numel()
          self.d2h_bytes += sum(hs.numel() + 4 * hn.numel() for hs, hn in pinned)
        ev = torch.cuda.Event()
        ev.record(st)
      front[k] = (c, pinned, ev)

    def issue_back(k):
      c, pinned, ev = front.pop(k)
      ev.synchronize()
      strings = []
      for hs, hn in pinned:
        a, nb = hs.numpy(), hn.numpy()
        strings.append([a[i, :nb[i]].tobytes() for i in range(len(nb))])
      host = Compressed(strings[0], strings[1], c.x_shape, c.y_shape, c.z_shape)
      st = self.streams[k % self.depth]
      with torch.cuda.stream(st):
        dc = model.to_device(host)
        x_hat = quantize_image(model.decompress(dc))
        out = torch.empty(x_hat.shape, dtype=torch.uint8, pin_memory=True)
        out.copy_(x_hat, non_blocking=True)
        if k == 0:
          self.h2d_bytes += sum(t.numel() * t.element_size() for pair in (dc.string, dc.side_string)
                                for t in pair)
          self.d2h_bytes += out.numel()
      results[k] = (host, out)

    with rc.deferred_checks():
      n = len(batches_u8)
      for k in range(n + self.depth - 1):
        if k < n:
          issue_front(k)
        j = k - (self.depth - 1)
        if 0 <= j < n:
          issue_back(j)
      self._join(cur)
    cur.synchronize()
    return results


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

  def compress(self, x):
    x = x.to(self.device)
    y, z, sigma = self.analyse(x)
    side_string = self.entropy_bottleneck.compress_device(z)
    string = self._conditional(sigma).compress_device(y)
    return Compressed(string, side_string, tuple(x.shape[1:-1]), tuple(y.shape[1:-1]),
                      tuple(z.shape[1:-1]))

  # ---- bmshj2018.py:383-399 ----------------------------------------------------------------
  def decompress(self, c):
    nf = self.num_filters
    z_hat = self.entropy_bottleneck.decompress_device(
        c.side_string, tuple(c.z_shape) + (nf,), channels=nf, out_dtype=torch.bfloat16)
    sigma = self.hyper_synthesis_transform(z_hat)
    sigma = sigma[:, :c.y_shape[0], :c.y_shape[1], :].contiguous()
    y_hat = self._conditional(sigma).decompress_device(c.string, out_dtype=torch.bfloat16)
    x_hat = self.synthesis_transform(y_hat)
    return x_hat[:, :c.x_shape[0], :c.x_shape[1], :]

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

  def pipeline(self, depth=3, coder_sms=32, coder_streams_per_cta=4):
    return Pipeline(self, depth, coder_sms, coder_streams_per_cta)

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
