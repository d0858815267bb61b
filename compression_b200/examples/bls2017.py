"""Factorised-prior codec of Ballé, Laparra & Simoncelli (2017) — mirror of the model in
`/root/reference/examples/bls2017.py` (transforms :70-127, train :130-202, compress :205-271,
decompress :274-305) over the B200 kernels.

  Bls2017Model.compress / decompress   inference path: convs with GDN fused in the epilogue,
                                       EntropyBottleneck symbols + batched range coder
  Bls2017Trainer.step                  one training step: forward with uniform noise, R-D loss,
                                       backward (conv gradients through the CUDA kernels),
                                       gradient all-reduce, two Adam updates, CDF table update
"""
import collections

import numpy as np
import torch

from compression_b200.python.layers import entropy_models
from compression_b200.python.layers import gdn
from compression_b200.python.layers import signal_conv
from compression_b200.python.ops import math_ops
from compression_b200.python.ops import train_ops

Compressed = collections.namedtuple("Compressed", ["string", "x_shape", "y_shape"])


quantize_image = signal_conv.quantize_image     # bls2017.py:57-60


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
  """bls2017.py:70-97."""

  def __init__(self, num_filters, impl=0):
    C = signal_conv.SignalConv2D
    super(AnalysisTransform, self).__init__([
        C(num_filters, (9, 9), name="layer_0", corr=True, strides_down=4, padding="same_zeros",
          use_bias=True, activation=gdn.GDN(name="gdn_0"), impl=impl),
        C(num_filters, (5, 5), name="layer_1", corr=True, strides_down=2, padding="same_zeros",
          use_bias=True, activation=gdn.GDN(name="gdn_1"), impl=impl),
        C(num_filters, (5, 5), name="layer_2", corr=True, strides_down=2, padding="same_zeros",
          use_bias=False, activation=None, impl=impl),
    ])


class SynthesisTransform(_Sequential):
  """bls2017.py:100-127."""

  def __init__(self, num_filters, impl=0):
    C = signal_conv.SignalConv2D
    super(SynthesisTransform, self).__init__([
        C(num_filters, (5, 5), name="layer_0", corr=False, strides_up=2, padding="same_zeros",
          use_bias=True, activation=gdn.GDN(name="igdn_0", inverse=True), impl=impl),
        C(num_filters, (5, 5), name="layer_1", corr=False, strides_up=2, padding="same_zeros",
          use_bias=True, activation=gdn.GDN(name="igdn_1", inverse=True), impl=impl),
        C(3, (9, 9), name="layer_2", corr=False, strides_up=4, padding="same_zeros",
          use_bias=True, activation=None, impl=impl),
    ])


class Bls2017Model(object):

  def __init__(self, num_filters=128, device=None, seed=0, impl=0):
    self.num_filters = int(num_filters)
    self.device = torch.device(device) if device is not None else torch.device(
        "cuda", torch.cuda.current_device())
    g = torch.Generator().manual_seed(seed)
    nf = self.num_filters
    self.analysis_transform = AnalysisTransform(nf, impl).build(3, self.device, g)
    self.synthesis_transform = SynthesisTransform(nf, impl).build(nf, self.device, g)
    self.entropy_bottleneck = entropy_models.EntropyBottleneck()
    self.entropy_bottleneck.build(nf, self.device, g)
    self.entropy_bottleneck.update()

  # ---- bls2017.py:205-271 ------------------------------------------------------------------
  def compress(self, x):
    """x: float32 [B,H,W,3] in [0,1] or the uint8 image (a byte b stands for b / 255)."""
    x = x.to(self.device)
    y = self.analysis_transform(x)
    string = self.entropy_bottleneck.compress_device(y)
    return Compressed(string, tuple(x.shape[1:-1]), tuple(y.shape[1:-1]))

  # ---- bls2017.py:274-305 ------------------------------------------------------------------
  def decompress(self, c, out_dtype=torch.float32):
    nf = self.num_filters
    y_hat = self.entropy_bottleneck.decompress_device(
        c.string, tuple(c.y_shape) + (nf,), channels=nf, out_dtype=torch.bfloat16)
    x_hat = self.synthesis_transform(y_hat, out_dtype=out_dtype)
    return x_hat[:, :c.x_shape[0], :c.x_shape[1], :]

  def evaluate(self, x):
    """The --verbose branch of compress (bls2017.py:226-244)."""
    y = self.analysis_transform(x)
    y_hat, lik = self.entropy_bottleneck(y, training=False)
    x_hat = self.synthesis_transform(y_hat.to(torch.bfloat16))[:, :x.shape[1], :x.shape[2], :]
    num_pixels = float(np.prod(x.shape[:-1]))
    bpp = torch.log(lik).sum() / (-np.log(2) * num_pixels)
    mse = torch.mean((x * 255 - torch.clamp(torch.round(x_hat * 255), 0, 255)) ** 2)
    return dict(x_hat=x_hat, y=y, y_hat=y_hat, likelihoods=lik, eval_bpp=float(bpp),
                mse=float(mse))

  @staticmethod
  def to_host(c):
    from compression_b200.python.ops import range_coding_ops as rc
    return Compressed(rc.strings_to_host(*c.string), c.x_shape, c.y_shape)

  def to_device(self, c):
    from compression_b200.python.ops import range_coding_ops as rc
    if isinstance(c.string, tuple):
      return c
    return Compressed(rc.strings_to_device(c.string, self.device), c.x_shape, c.y_shape)

  # ---- .tfci files (bls2017.py:247-254 write, :276-282 read): one image per file -----------
  @staticmethod
  def pack_tfci(c, image=0, model_id=None):
    from compression_b200.python.util.packed_tensors import PackedTensors
    packed = PackedTensors()
    if model_id is not None:
      packed.model = model_id
    packed.pack(["string", "int32", "int32"],
                [[c.string[image]], list(c.x_shape), list(c.y_shape)])
    return packed.string

  @staticmethod
  def unpack_tfci(data):
    from compression_b200.python.util.packed_tensors import PackedTensors
    s, xs, ys = PackedTensors(data).unpack(["string", "int32", "int32"])
    return Compressed([bytes(s[0])], tuple(int(v) for v in xs), tuple(int(v) for v in ys))

  def conv_layers(self):
    return self.analysis_transform.layers + self.synthesis_transform.layers

  # ---- export (host copies, for checkpoints and for the CPU checker in tests/bench) --------
  def export_weights(self):
    """Effective float32 weights keyed like oracle/cpu_codec.py (CpuBls2017) expects."""
    w = {}
    for prefix, seq in (("a", self.analysis_transform), ("s", self.synthesis_transform)):
      for i, l in enumerate(seq.layers):
        name = "%s%d" % (prefix, i)
        w[name + "/kernel"] = l.kernel.detach().float().cpu().numpy()
        if l.bias is not None:
          w[name + "/bias"] = l.bias.detach().float().cpu().numpy()
        if isinstance(l.activation, gdn.GDN):
          beta, gamma = l.activation.effective()
          w[name + "/beta"] = beta.detach().cpu().numpy()
          w[name + "/gamma"] = gamma.detach().cpu().numpy()
    eb = self.entropy_bottleneck
    for i, m in enumerate(eb.matrices):
      w["eb/matrix_%d" % i] = m.detach().cpu().numpy()
      w["eb/bias_%d" % i] = eb.biases[i].detach().cpu().numpy()
    for i, f in enumerate(eb.factors):
      w["eb/factor_%d" % i] = f.detach().cpu().numpy()
    w["eb/quantiles"] = eb.quantiles.detach().cpu().numpy()
    return w

  def export_tables(self):
    eb = self.entropy_bottleneck
    return dict(eb_cdf=eb._quantized_cdf.cpu().numpy(), eb_len=eb._cdf_length.cpu().numpy(),
                eb_off=eb._offset.cpu().numpy())


class _tf32_matmul(object):
  """Scoped `torch.backends.cuda.matmul.allow_tf32`."""

  def __enter__(self):
    self._old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True

  def __exit__(self, *exc):
    torch.backends.cuda.matmul.allow_tf32 = self._old
    return False


class _GdnContraction(torch.autograd.Function):
  """[P, C] x [C, C] contraction of GDN (gdn.py:183-185) and its two gradients as TF32
  tensor-core GEMMs (fp32 accumulate).  The transforms around it run on bf16 operands, so TF32
  here keeps the step's precision policy; everything else in the step (the density network in
  particular) stays in fp32 SIMT arithmetic."""

  @staticmethod
  def forward(ctx, x2, gamma):
    ctx.save_for_backward(x2, gamma)
    with _tf32_matmul():
      return x2 @ gamma

  @staticmethod
  def backward(ctx, g):
    x2, gamma = ctx.saved_tensors
    with _tf32_matmul():
      return g @ gamma.t(), x2.t() @ g


class _GdnTrain(torch.autograd.Function):
  """GDN / IGDN of the training graph (gdn.py:171-202): three elementwise CUDA passes
  (csrc/gdn_train.cu) around the C x C contraction and its two gradients.  With 128 channels
  the contraction runs on the tcgen05 kernels as a 1x1 convolution over bfloat16 operands
  (x^2 and gnorm rounded to bfloat16, float32 accumulation: the arithmetic of the inference
  epilogue) — norm = conv1x1(x^2, gamma), gx2 = conv1x1(gnorm, gamma^T), d gamma = the 1x1
  filter gradient of (x^2, gnorm); other widths keep the TF32 library GEMMs."""

  @staticmethod
  def forward(ctx, x, beta, gamma, inverse):
    import ctypes
    from compression_b200 import _lib
    shape, c = x.shape, x.shape[-1]
    x2d = x.reshape(-1, c).contiguous()
    p = x2d.shape[0]
    L, st = _lib.lib(), _lib.stream_ptr(x.device)
    umma = c == 128 and x.dim() == 4 and p >= 128
    if umma:
      x2 = torch.empty((p, c), dtype=torch.bfloat16, device=x.device)
      _lib.check(L.tfc_gdn_train_square(_lib.ptr(x2d), ctypes.c_int64(p), ctypes.c_int32(c),
                                        _lib.ptr(x2), st))
      n, h, w, _ = shape
      nraw = train_ops._conv_raw(x2.view(shape), gamma.view(1, 1, c, c), None, 1, 1, 1, 1, 0, 0,
                                 h, w, torch.float32).view(p, c)
    else:
      x2 = x2d * x2d
      with _tf32_matmul():
        nraw = x2 @ gamma
    b = beta.detach().float().contiguous()
    if umma:
      # y feeds the next layer's convolution, which reads bfloat16: written as such
      y = torch.empty((p, c), dtype=torch.bfloat16, device=x.device)
      _lib.check(L.tfc_gdn_train_forward_bf16(
          _lib.ptr(x2d), _lib.ptr(nraw), _lib.ptr(b), ctypes.c_int64(p), ctypes.c_int32(c),
          int(bool(inverse)), _lib.ptr(y), _lib.ptr(nraw), st))
    else:
      y = torch.empty_like(x2d)
      _lib.check(L.tfc_gdn_train_forward(
          _lib.ptr(x2d), _lib.ptr(nraw), _lib.ptr(b), ctypes.c_int64(p), ctypes.c_int32(c),
          int(bool(inverse)), _lib.ptr(y), _lib.ptr(nraw), st))
    ctx.save_for_backward(x2d, x2, nraw, gamma)
    ctx.inverse, ctx.shape, ctx.umma = bool(inverse), shape, umma
    return y.reshape(shape)

  @staticmethod
  def backward(ctx, gy):
    import ctypes
    from compression_b200 import _lib
    from compression_b200.python.layers.signal_conv import ConvDesc
    x2d, x2, norm, gamma = ctx.saved_tensors
    p, c = x2d.shape
    gx = torch.empty_like(x2d)
    gnorm = torch.empty_like(x2d)
    L = _lib.lib()
    st = _lib.stream_ptr(x2d.device)
    if ctx.umma:
      n, h, w, _ = ctx.shape
      gy2d = gy.reshape(p, c)
      if gy2d.dtype not in (torch.float32, torch.bfloat16):
        gy2d = gy2d.float()
      gy2d = gy2d.contiguous()
      gnb = torch.empty((p, c), dtype=torch.bfloat16, device=x2d.device)
      _lib.check(L.tfc_gdn_train_backward_pre2(
          _lib.ptr(gy2d), 1 if gy2d.dtype == torch.bfloat16 else 0, _lib.ptr(x2d), _lib.ptr(norm),
          ctypes.c_int64(p), ctypes.c_int32(c), int(ctx.inverse), _lib.ptr(gx), _lib.ptr(gnorm),
          _lib.ptr(gnb), st))
      gx2 = train_ops._conv_raw(gnb.view(ctx.shape), gamma.t().contiguous().view(1, 1, c, c), None,
                                1, 1, 1, 1, 0, 0, h, w, torch.float32).view(p, c)
      ggamma = torch.empty((1, 1, c, c), dtype=torch.float32, device=x2d.device)
      d = ConvDesc(n=n, h=h, w=w, cin=c, ho=h, wo=w, cout=c, kh=1, kw=1, sd=1, su=1, pad_top=0,
                   pad_left=0, act=0, in_dtype=1, out_dtype=0, impl=0, flags=0)
      _lib.check(L.tfc_signal_conv2d_wgrad(ctypes.byref(d), _lib.ptr(x2), _lib.ptr(gnb),
                                           _lib.ptr(ggamma), st))
      ggamma = ggamma.view(c, c)
    else:
      gy2d = gy.reshape(p, c).float().contiguous()
      _lib.check(L.tfc_gdn_train_backward_pre(
          _lib.ptr(gy2d), _lib.ptr(x2d), _lib.ptr(norm), ctypes.c_int64(p), ctypes.c_int32(c),
          int(ctx.inverse), _lib.ptr(gx), _lib.ptr(gnorm), st))
      with _tf32_matmul():
        gx2 = gnorm @ gamma.t()
        ggamma = x2.t() @ gnorm
    _lib.check(L.tfc_gdn_train_backward_post(
        _lib.ptr(gx), _lib.ptr(x2d), _lib.ptr(gx2), ctypes.c_int64(p), ctypes.c_int32(c), st))
    gbeta = torch.empty((c,), dtype=torch.float32, device=x2d.device)
    _lib.check(L.tfc_column_sums(_lib.ptr(gnorm), ctypes.c_int64(p), ctypes.c_int32(c),
                                 _lib.ptr(gbeta), st))
    return gx.reshape(ctx.shape), gbeta, ggamma, None


class Bls2017Trainer(object):
  """One synchronous data-parallel training step (bls2017.py:156-185).

  All trainable variables live in two flat buckets (main: kernels, biases, GDN parameters,
  density network; aux: the quantiles), each with its own Adam state: the gradient exchange
  is one all-reduce per bucket and each update one kernel.  `updates[0]` (the CDF table
  rebuild) is a pure function of the then identical parameters, so every rank runs it locally.
  """

  def __init__(self, model, lmbda=0.01, main_lr=1e-4, aux_lr=1e-3):
    self.model, self.lmbda = model, float(lmbda)
    self.main_lr, self.aux_lr = float(main_lr), float(aux_lr)
    eb = model.entropy_bottleneck
    slots = []                       # (object, attribute[, index]) of every trainable
    for l in model.conv_layers():
      slots.append((l, "kernel_var", None))
      if l.bias_var is not None:
        slots.append((l, "bias_var", None))
      if isinstance(l.activation, gdn.GDN):
        slots.append((l.activation, "reparam_beta", None))
        slots.append((l.activation, "reparam_gamma", None))
    for name in ("matrices", "biases", "factors"):
      for i in range(len(getattr(eb, name))):
        slots.append((eb, name, i))
    self._slots = slots
    self.main = train_ops.FlatBucket([self._get(s) for s in slots], model.device)
    for s, v in zip(slots, self.main.views):
      self._set(s, v)
    self.aux = train_ops.FlatBucket([eb.quantiles], model.device)
    eb.quantiles = self.aux.views[0]
    self._invalidate()

  @staticmethod
  def _get(slot):
    obj, name, i = slot
    v = getattr(obj, name)
    return v if i is None else v[i]

  @staticmethod
  def _set(slot, value):
    obj, name, i = slot
    if i is None:
      setattr(obj, name, value)
    else:
      getattr(obj, name)[i] = value

  def _invalidate(self):
    for l in self.model.conv_layers():
      l.invalidate()

  @property
  def num_parameters(self):
    return int(self.main.param.numel() + self.aux.param.numel())

  # ---- forward with the training-time graph (bls2017.py:156-171) ----------------------------
  @staticmethod
  def _gdn(act, x):
    """gdn.py:171-202 in torch ops: the 1x1 contraction is a plain GEMM."""
    beta, gamma = act.beta_parameterizer(act.reparam_beta), act.gamma_parameterizer(act.reparam_gamma)
    if x.dtype == torch.float32 and x.shape[-1] % 4 == 0:
      return _GdnTrain.apply(x, beta, gamma.float().contiguous(), act.inverse)
    norm = _GdnContraction.apply((x * x).reshape(-1, x.shape[-1]), gamma) + beta
    norm = norm.reshape(x.shape)
    return x * (torch.sqrt(norm) if act.inverse else torch.rsqrt(norm))

  def _transform(self, seq, t):
    for layer in seq.layers:
      t = train_ops.signal_conv2d(layer, t if t.shape[-1] == 3 else t.to(torch.bfloat16))
      if isinstance(layer.activation, gdn.GDN):
        t = self._gdn(layer.activation, t)
      elif layer.activation is not None:
        t = layer.activation(t)                     # torch.relu of the hyper transforms
    return t

  def _likelihood(self, y_tilde):
    """entropy_models.py:626-658 + the bound of :181-184, differentiable."""
    eb = self.model.entropy_bottleneck
    c = y_tilde.shape[-1]
    if len(eb.filters) >= 1 and len(set(eb.filters)) == 1 and eb.filters[0] <= 4 and \
        y_tilde.dtype == torch.float32:
      # one kernel each way instead of ~40 launches of batched matmuls with inner size <= 3
      return eb.likelihood_training(y_tilde)
    v = y_tilde.reshape(-1, c).t().reshape(c, 1, -1)
    lower = eb._logits_cumulative(v - .5)
    upper = eb._logits_cumulative(v + .5)
    sign = -torch.sign(lower + upper).detach()
    lik = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))
    lik = math_ops.lower_bound(lik, eb.likelihood_bound)
    return lik.reshape(c, -1).t().reshape(y_tilde.shape)

  def losses(self, x, noise=None):
    """x float32 [B,H,W,3] in [0,1] -> dict(loss, bpp, mse, aux)."""
    model = self.model
    num_pixels = float(np.prod(x.shape[:-1]))
    y = self._transform(model.analysis_transform, x)
    if noise is None:
      noise = torch.rand_like(y) - .5                         # entropy_models.py:605-607
    y_tilde = y + noise
    lik = self._likelihood(y_tilde)
    x_tilde = self._transform(model.synthesis_transform, y_tilde)
    bpp = torch.log(lik).sum() / (-np.log(2) * num_pixels)
    mse = torch.mean((x - x_tilde) ** 2) * 255 ** 2
    return dict(loss=self.lmbda * mse + bpp, bpp=bpp, mse=mse,
                aux=model.entropy_bottleneck.losses[0], x_tilde=x_tilde)

  def _forward_backward(self, x, noise):
    self.main.zero_grad()
    self.aux.zero_grad()
    out = self.losses(x, noise)
    out["loss"].backward()
    out["aux"].backward()
    return out

  def _replay(self, x, noise):
    """Forward + backward of one step as ONE CUDA graph launch (captured on first use per input
    shape): the step is ~600 small launches and host-bound when they are issued one by one.
    The gradient all-reduce, the Adam updates and the table rebuild stay outside the graph."""
    graphs = self.__dict__.setdefault("_graphs", {})
    key = (tuple(x.shape), x.dtype, noise is not None)
    rec = graphs.get(key)
    if rec is None:
      sx = x.clone()
      sn = noise.clone() if noise is not None else None
      side = torch.cuda.Stream(device=x.device)
      side.wait_stream(torch.cuda.current_stream(x.device))
      with torch.cuda.stream(side):       # warm-up: allocator pools, lazily set attributes
        for _ in range(2):
          self._forward_backward(sx, sn)
      torch.cuda.current_stream(x.device).wait_stream(side)
      graph = torch.cuda.CUDAGraph()
      with torch.cuda.graph(graph):
        out = self._forward_backward(sx, sn)
      stats = {k: v.detach() for k, v in out.items() if k != "x_tilde"}
      rec = graphs[key] = (graph, sx, sn, stats)
    graph, sx, sn, stats = rec
    sx.copy_(x)
    if sn is not None:
      sn.copy_(noise)
    graph.replay()
    return stats

  def step(self, x, noise=None, graph=False):
    """train_op = group(main_step, aux_step, updates[0]) (bls2017.py:173-180).  `graph=True`
    replays the forward + backward as a captured CUDA graph (same kernels, same results up to
    the order of the filter gradient's atomic adds)."""
    out = self._replay(x, noise) if graph else self._forward_backward(x, noise)
    scale = self.main.all_reduce()
    self.aux.all_reduce()
    train_ops.adam_step(self.main, self.main_lr, scale)
    train_ops.adam_step(self.aux, self.aux_lr, scale)
    self._invalidate()
    self.model.entropy_bottleneck.update()
    return {k: float(v.detach()) for k, v in out.items() if k != "x_tilde"}
