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
