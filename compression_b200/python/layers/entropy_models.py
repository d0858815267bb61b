"""Entropy models — mirror of `tensorflow_compression/python/layers/entropy_models.py`.

Classes and methods keep the reference's names and argument meaning:
  EntropyModel (:41-295), EntropyBottleneck (:298-704), SymmetricConditional (:707-950),
  GaussianConditional / LogisticConditional / LaplacianConditional (:953-1001).
Tensors are torch CUDA tensors, channels_last.  `compress` returns a list of `bytes` (one per
batch element, like the string tensor of the reference); `compress_device` /
`decompress_device` keep the strings in HBM as (slab [B, cap] uint8, nbytes [B] int32).
"""
import ctypes

import numpy as np
import scipy.special
import scipy.stats
import torch

from compression_b200 import _lib
from compression_b200.python.ops import math_ops
from compression_b200.python.ops import range_coding_ops as rc

__all__ = ["EntropyModel", "EntropyBottleneck", "SymmetricConditional", "GaussianConditional",
           "LogisticConditional", "LaplacianConditional"]

_i64 = ctypes.c_int64
_i32 = ctypes.c_int32
_f32 = ctypes.c_float


class EntropyModel(object):
  """Entropy model (base class) — entropy_models.py:41-295."""

  def __init__(self, tail_mass=2 ** -8, likelihood_bound=1e-9, range_coder_precision=16,
               dtype=torch.float32, name=None):
    self._tail_mass = float(tail_mass)
    if not 0 < self.tail_mass < 1:
      raise ValueError("`tail_mass` must be between 0 and 1, got {}.".format(self.tail_mass))
    self._likelihood_bound = float(likelihood_bound)
    self._range_coder_precision = int(range_coder_precision)
    if dtype != torch.float32:
      raise NotImplementedError("entropy models compute in float32")
    self.dtype = dtype
    self.name = name
    self.built = False

  tail_mass = property(lambda self: self._tail_mass)
  likelihood_bound = property(lambda self: self._likelihood_bound)
  range_coder_precision = property(lambda self: self._range_coder_precision)

  def _pmf_to_cdf(self, pmf, tail_mass, pmf_length, max_length):
    """entropy_models.py:131-146: per row, `pmf[:length] ++ [tail]` goes through
    pmf_to_quantized_cdf and is zero-padded to max_length + 2.  Rows of equal length share
    one launch of the batched kernel."""
    pmf = torch.as_tensor(pmf, dtype=torch.float32)
    dev = pmf.device if pmf.is_cuda else torch.device("cuda", torch.cuda.current_device())
    pmf = pmf.to(dev)
    tail = torch.as_tensor(tail_mass, dtype=torch.float32).to(dev).reshape(-1, 1)
    lengths = np.asarray(torch.as_tensor(pmf_length).cpu(), dtype=np.int64)
    rows = pmf.shape[0]
    cdf = torch.zeros((rows, int(max_length) + 2), dtype=torch.int32, device=dev)
    for n in np.unique(lengths):
      sel = torch.as_tensor(np.nonzero(lengths == n)[0], device=dev)
      prob = torch.cat([pmf[sel, :int(n)], tail[sel]], dim=1).contiguous()
      cdf[sel, :int(n) + 2] = rc.pmf_to_quantized_cdf(prob, self.range_coder_precision)
    return cdf

  # subclasses provide: _quantize_symbols(x), _dequantize(symbols, out_dtype), _likelihood,
  # _prepare_indexes, and the table triple (_quantized_cdf, _cdf_length, _offset)

  def compress_device(self, inputs, cap=None):
    """entropy_models.py:189-249 — one stream per batch element, strings stay on the device."""
    inputs = self._check_inputs(inputs)
    symbols, indexes, shared = self._symbols_and_indexes(inputs)
    return rc.unbounded_index_range_encode_batch(
        symbols, indexes, self._quantized_cdf, self._cdf_length, self._offset,
        precision=self.range_coder_precision, overflow_width=4, debug_level=0, cap=cap)

  def compress(self, inputs):
    out, nbytes = self.compress_device(inputs)
    return rc.strings_to_host(out, nbytes)

  def _decode_symbols(self, strings, indexes, shared):
    if isinstance(strings, (tuple, list)) and len(strings) == 2 and \
        isinstance(strings[0], torch.Tensor):
      enc, nbytes = strings
    else:
      enc, nbytes = rc.strings_to_device(list(strings), indexes.device)
    return rc.unbounded_index_range_decode_batch(
        enc, nbytes, indexes, self._quantized_cdf, self._cdf_length, self._offset,
        precision=self.range_coder_precision, overflow_width=4, debug_level=0,
        shared_index=shared)

  def _check_inputs(self, inputs):
    if not isinstance(inputs, torch.Tensor):
      inputs = torch.as_tensor(np.asarray(inputs))
    if not inputs.dtype.is_floating_point:
      raise ValueError("{} can't take integer inputs.".format(type(self).__name__))
    _lib.require_cuda(inputs, "inputs")
    return inputs.float().contiguous()

  def compute_output_shape(self, input_shape):
    return tuple(input_shape), tuple(input_shape)


class _EbLikelihood(torch.autograd.Function):
  """EntropyBottleneck._likelihood + likelihood bound (entropy_models.py:626-658, :181-184) as
  one CUDA kernel each way; `params` is the packed effective parameter matrix [channels, 43]."""

  @staticmethod
  def forward(ctx, x, params, eb):
    lik = torch.empty_like(x)
    p = params.detach().contiguous()
    nl, fw = eb._filter_args()
    _lib.check(_lib.lib().tfc_eb_call(
        _lib.ptr(x), _i64(x.numel()), _i32(eb.channels), _lib.ptr(p), nl, fw,
        ctypes.c_void_p(0), _f32(max(eb.likelihood_bound, 0.0)), 1, ctypes.c_void_p(0),
        _lib.ptr(lik), _lib.stream_ptr(x.device)))
    ctx.save_for_backward(x, p)
    ctx.eb = eb
    return lik

  @staticmethod
  def backward(ctx, grad):
    x, p = ctx.saved_tensors
    eb = ctx.eb
    grad = grad.contiguous()
    gx = torch.empty_like(x)
    gp = torch.empty_like(p)
    nl, fw = eb._filter_args()
    _lib.check(_lib.lib().tfc_eb_likelihood_backward(
        _lib.ptr(x), _i64(x.numel()), _i32(eb.channels), _lib.ptr(p), nl, fw,
        _f32(max(eb.likelihood_bound, 0.0)), _lib.ptr(grad), _lib.ptr(gx), _lib.ptr(gp),
        _lib.stream_ptr(x.device)))
    return gx, gp, None


class _GaussianLikelihood(torch.autograd.Function):
  """GaussianConditional._likelihood with both lower bounds (entropy_models.py:822-825, :916-932,
  :965-969, :181-184) as one CUDA kernel each way: forward tfc_gaussian_call (mode 1), backward
  tfc_gaussian_likelihood_backward (the lower_bound gradient rule of math_ops.py:61-74 on the
  scale and on the likelihood)."""

  @staticmethod
  def forward(ctx, values, sigma, scale_bound, likelihood_bound):
    values, sigma = values.contiguous(), sigma.contiguous()
    lik = torch.empty_like(values)
    _lib.check(_lib.lib().tfc_gaussian_call(
        _lib.ptr(values), _lib.ptr(sigma), _i64(values.numel()), _f32(scale_bound),
        _f32(likelihood_bound), 1, ctypes.c_void_p(0), _lib.ptr(lik),
        _lib.stream_ptr(values.device)))
    ctx.save_for_backward(values, sigma)
    ctx.bounds = (scale_bound, likelihood_bound)
    return lik

  @staticmethod
  def backward(ctx, grad):
    values, sigma = ctx.saved_tensors
    grad = grad.contiguous()
    gv, gs = torch.empty_like(values), torch.empty_like(sigma)
    _lib.check(_lib.lib().tfc_gaussian_likelihood_backward(
        _lib.ptr(values), _lib.ptr(sigma), _lib.ptr(grad), _i64(values.numel()),
        _f32(ctx.bounds[0]), _f32(ctx.bounds[1]), _lib.ptr(gv), _lib.ptr(gs),
        _lib.stream_ptr(values.device)))
    return gv, gs, None, None


def _wants_grad(*tensors):
  return torch.is_grad_enabled() and any(
      isinstance(t, torch.Tensor) and t.requires_grad for t in tensors)


class EntropyBottleneck(EntropyModel):
  """Entropy bottleneck layer (factorised prior) — entropy_models.py:298-704."""

  def __init__(self, init_scale=10, filters=(3, 3, 3), data_format="channels_last", **kwargs):
    super(EntropyBottleneck, self).__init__(**kwargs)
    self._init_scale = float(init_scale)
    self._filters = tuple(int(f) for f in filters)
    self._data_format = str(data_format)
    if self.data_format not in ("channels_first", "channels_last"):
      raise ValueError("Unknown data format: '{}'.".format(self.data_format))
    if len(set(self._filters)) > 1:
      raise NotImplementedError("non-uniform `filters` are not built")
    self._losses, self._updates = [], []

  init_scale = property(lambda self: self._init_scale)
  filters = property(lambda self: self._filters)
  data_format = property(lambda self: self._data_format)

  # ---- variables (entropy_models.py:461-515) -------------------------------------------------
  def build(self, channels, device=None, generator=None):
    if channels is None:
      raise ValueError("The channel dimension of the inputs must be defined.")
    dev = device or torch.device("cuda", torch.cuda.current_device())
    c = self.channels = int(channels)
    f = (1,) + self.filters + (1,)
    scale = self.init_scale ** (1 / (len(self.filters) + 1))
    self.matrices, self.biases, self.factors = [], [], []
    for i in range(len(self.filters) + 1):
      init = float(np.log(np.expm1(1 / scale / f[i + 1])))
      self.matrices.append(torch.full((c, f[i + 1], f[i]), init, dtype=torch.float32, device=dev))
      b = torch.rand((c, f[i + 1], 1), generator=generator, dtype=torch.float32) - .5
      self.biases.append(b.to(dev))
      if i < len(self.filters):
        self.factors.append(torch.zeros((c, f[i + 1], 1), dtype=torch.float32, device=dev))
    q = torch.tensor([[[-self.init_scale, 0., self.init_scale]]], dtype=torch.float32)
    self.quantiles = q.repeat(c, 1, 1).to(dev)
    # placeholder tables valid before the first update (entropy_models.py:563-583)
    half = torch.full((c, 2), .5, dtype=torch.float32, device=dev)
    self._quantized_cdf = rc.pmf_to_quantized_cdf(half, self.range_coder_precision)
    self._cdf_length = torch.full((c,), 3, dtype=torch.int32, device=dev)
    self.built = True
    self.refresh()

  def variables(self):
    return list(self.matrices) + list(self.biases) + list(self.factors) + [self.quantiles]

  def _versions(self):
    return tuple((id(v), v._version) for v in self.variables())

  def _sync(self):
    """The packed density parameters and the medians are caches of the variables: re-derive
    them when a variable was replaced or written in place (an optimizer step) since."""
    if getattr(self, "_derived_from", None) != self._versions():
      self.refresh()

  def refresh(self):
    """Re-derives everything that is a pure function of the variables except the CDF tables
    (those change only through `update()`, like the reference's update op)."""
    self._derived_from = self._versions()
    q = self.quantiles
    self._medians = q[:, 0, 1].detach().contiguous()
    minima = torch.clamp(torch.ceil(self._medians - q[:, 0, 0]).to(torch.int32), min=0)
    maxima = torch.clamp(torch.ceil(q[:, 0, 2] - self._medians).to(torch.int32), min=0)
    self._offset = (-minima).contiguous()
    self._pmf_start = self._medians - minima.float()
    self._pmf_length = maxima + minima + 1
    self._params = self._pack_params()

  def _pack_params(self, detach=True):
    """Reparameterised density-net weights in the per-channel layout tfc_eb_call expects
    (`detach=False`: as a differentiable function of the variables, for the training step)."""
    parts = []
    for i in range(len(self.filters) + 1):
      parts.append(torch.nn.functional.softplus(self.matrices[i]).reshape(self.channels, -1))
      parts.append(self.biases[i].reshape(self.channels, -1))
      if i < len(self.factors):
        parts.append(torch.tanh(self.factors[i]).reshape(self.channels, -1))
    packed = torch.cat(parts, dim=1).float().contiguous()
    return packed.detach() if detach else packed

  def likelihood_training(self, x):
    """Likelihood of `x` as given (the caller has added its uniform noise), differentiable with
    respect to x and the density variables: forward tfc_eb_call (mode 1), backward
    tfc_eb_likelihood_backward.  Needs at least one hidden layer."""
    return _EbLikelihood.apply(x.contiguous(), self._pack_params(detach=False), self)

  def _filter_args(self):
    return _i32(len(self.filters)), _i32(self.filters[0] if self.filters else 1)

  def _logits_cumulative(self, inputs, stop_gradient=False):
    """entropy_models.py:402-436 in torch ops (differentiable; used by the aux loss).
    inputs: (channels, 1, batch)."""
    logits = inputs
    for i in range(len(self.filters) + 1):
      m = torch.nn.functional.softplus(self.matrices[i])
      b = self.biases[i]
      if stop_gradient:
        m, b = m.detach(), b.detach()
      logits = torch.matmul(m, logits) + b
      if i < len(self.factors):
        fac = torch.tanh(self.factors[i])
        if stop_gradient:
          fac = fac.detach()
        logits = logits + fac * torch.tanh(logits)
    return logits

  def _logits_cumulative_kernel(self, x):
    """Same function through the CUDA kernel; x: (channels, samples) float32."""
    x = x.float().contiguous()
    out = torch.empty_like(x)
    nl, fw = self._filter_args()
    _lib.check(_lib.lib().tfc_eb_logits_cumulative(
        _lib.ptr(x), _i32(x.shape[0]), _i32(x.shape[1]), _lib.ptr(self._params), nl, fw,
        _lib.ptr(out), _lib.stream_ptr(x.device)))
    return out

  @property
  def losses(self):
    """[aux loss] = sum |logits_cumulative(quantiles) - target| (entropy_models.py:502-518)."""
    t = getattr(self, "_aux_target", None)
    if t is None or t.device != self.quantiles.device:
      target = float(np.log(2 / self.tail_mass - 1))
      t = self._aux_target = torch.tensor([-target, 0., target], dtype=torch.float32,
                                          device=self.quantiles.device)
    logits = self._logits_cumulative(self.quantiles, stop_gradient=True)
    return [torch.sum(torch.abs(logits - t))]

  # ---- the CDF tables and the update op ------------------------------------------------------
  # The table is [channels, max_length + 2] with max_length = max(_pmf_length), a function of the
  # quantiles: data-dependent, and the reference's update op runs in EVERY training step
  # (bls2017.py:173-180).  Reading max_length back each step stalls the step on the host, so
  # update() builds the tables for the max_length it saw LAST (a host integer), queues an
  # asynchronous copy of the true one, and the check is settled by whoever reads the tables
  # next: if the guess was wrong (the quantiles moved a table boundary) the tables are rebuilt
  # with the right size before anybody sees them.  Observable results are those of the
  # synchronous form; a training loop that only updates never waits.
  @property
  def _quantized_cdf(self):
    self._settle_tables()
    return self._tables[0]

  @_quantized_cdf.setter
  def _quantized_cdf(self, value):        # tables set from outside (build, a checkpoint) are final
    self._pending_max = None
    self._tables = (value, getattr(self, "_tables", (None, None))[1])

  @property
  def _cdf_length(self):
    self._settle_tables()
    return self._tables[1]

  @_cdf_length.setter
  def _cdf_length(self, value):
    self._pending_max = None
    self._tables = (getattr(self, "_tables", (None, None))[0], value)

  def _settle_tables(self):
    pending = getattr(self, "_pending_max", None)
    if pending is None:
      return
    host, event, used = pending
    self._pending_max = None
    event.synchronize()
    true = int(host.item())
    self._max_length_seen = true
    if true != used:
      self._build_tables(true)          # rare: sized for the real max_length this time

  def _build_tables(self, max_length):
    """entropy_models.py:540-559 + :131-146 for the current variables, tables of width
    max_length + 2; no host round trip (row lengths stay on the device)."""
    dev = self._medians.device
    c, L = self.channels, int(max_length)
    samples = torch.arange(L, dtype=torch.float32, device=dev)
    samples = samples[None, :] + self._pmf_start[:, None]
    lower = self._logits_cumulative_kernel(samples - .5)
    upper = self._logits_cumulative_kernel(samples + .5)
    sign = -torch.sign(lower + upper)
    pmf = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))
    # tail mass as the reference takes it (:556-560): below the first sample of the row and
    # above the LAST sample of the padded grid
    tail = torch.sigmoid(lower[:, :1]) + torch.sigmoid(-upper[:, -1:])
    length = torch.clamp(self._pmf_length, max=L).to(torch.int64)
    # row r of the op's input: pmf[r, :len_r] ++ [tail_r]
    prob = torch.cat([pmf, torch.zeros((c, 1), dtype=pmf.dtype, device=dev)], dim=1)
    prob.scatter_(1, length[:, None], tail)
    prob = prob.contiguous()
    n = (length + 1).to(torch.int32).contiguous()
    cdf = torch.empty((c, L + 2), dtype=torch.int32, device=dev)
    _lib.check(_lib.lib().tfc_pmf_to_quantized_cdf_ragged(
        _lib.ptr(prob), _i64(c), _i32(L + 1), _lib.ptr(n), _i32(L + 1),
        int(self.range_coder_precision), _lib.ptr(cdf), _i32(L + 2), _lib.stream_ptr(dev)))
    self._tables = (cdf, (self._pmf_length + 2).to(torch.int32).contiguous())

  def update(self):
    """The reference's update op (entropy_models.py:588-596): rebuild quantized_cdf and
    cdf_length from the current variables."""
    self.refresh()
    pending = getattr(self, "_pending_max", None)
    if pending is not None and pending[1].query():       # an older check that nobody needed
      self._max_length_seen = int(pending[0].item())
    self._pending_max = None
    dev_max = self._pmf_length.max()
    guess = getattr(self, "_max_length_seen", None)
    if guess is None:
      guess = self._max_length_seen = int(dev_max.item())    # first update: one round trip
      self._build_tables(guess)
      return
    self._build_tables(guess)
    ring = self.__dict__.setdefault("_max_ring", [])
    if len(ring) < 4:
      ring.append(torch.empty((1,), dtype=torch.int32, pin_memory=True))
    host = ring[self.__dict__.setdefault("_max_ring_at", 0) % len(ring)]
    self._max_ring_at += 1
    host.copy_(dev_max.reshape(1).to(torch.int32), non_blocking=True)
    event = torch.cuda.Event()
    event.record(torch.cuda.current_stream(self._medians.device))
    self._pending_max = (host, event, guess)

  @property
  def updates(self):
    return [self.update]

  # ---- forward ------------------------------------------------------------------------------
  def _ensure_built(self, inputs):
    if self.data_format != "channels_last":
      raise NotImplementedError("EntropyBottleneck kernels are channels_last")
    if not self.built:
      self.build(inputs.shape[-1], inputs.device)
    if inputs.shape[-1] != self.channels:
      raise ValueError("built for %d channels, got %d" % (self.channels, inputs.shape[-1]))

  def __call__(self, inputs, training=False):
    """entropy_models.py:148-187 -> (values, likelihood)."""
    x = self._check_inputs(inputs)
    self._ensure_built(x)
    self._sync()
    mode = 0
    if training:  # additive uniform noise (entropy_models.py:605-607)
      x = x + (torch.rand_like(x) - .5)
      mode = 1
      if _wants_grad(x, *self.variables()):
        # the reference's differentiable training entry point: gradients flow to the inputs
        # and to the density variables through the likelihood kernels
        if not self.filters:
          raise NotImplementedError("the likelihood gradient kernel needs a hidden layer")
        return x, self.likelihood_training(x)
    values = torch.empty_like(x) if not training else x
    lik = torch.empty_like(x)
    nl, fw = self._filter_args()
    _lib.check(_lib.lib().tfc_eb_call(
        _lib.ptr(x), _i64(x.numel()), _i32(self.channels), _lib.ptr(self._params), nl, fw,
        _lib.ptr(self._medians), _f32(max(self.likelihood_bound, 0.0)), mode,
        _lib.ptr(values) if not training else ctypes.c_void_p(0), _lib.ptr(lik),
        _lib.stream_ptr(x.device)))
    return values, lik

  def _quantize_symbols(self, x):
    self._sync()
    sym = torch.empty(x.shape, dtype=torch.int32, device=x.device)
    _lib.check(_lib.lib().tfc_eb_symbols(
        _lib.ptr(x), _i64(x.numel()), _i32(self.channels), _lib.ptr(self._medians),
        _lib.ptr(sym), _lib.stream_ptr(x.device)))
    return sym

  def _dequantize(self, symbols, out_dtype=torch.float32):
    self._sync()
    out = torch.empty(symbols.shape, dtype=out_dtype, device=symbols.device)
    _lib.check(_lib.lib().tfc_eb_dequantize(
        _lib.ptr(symbols), _i64(symbols.numel()), _i32(self.channels), _lib.ptr(self._medians),
        _lib.ptr(out), 1 if out_dtype == torch.bfloat16 else 0,
        _lib.stream_ptr(symbols.device)))
    return out

  def _prepare_indexes(self, shape, channels=None, device=None):
    """entropy_models.py:660-681: the channel id at every position, shared by the batch."""
    if not self.built:
      self.build(channels if channels is not None else shape[-1], device)
    shape = tuple(int(s) for s in shape)
    idx = torch.arange(self.channels, dtype=torch.int32, device=self._medians.device)
    return idx.expand(shape).contiguous()

  def _symbols_and_indexes(self, x):
    self._ensure_built(x)
    return self._quantize_symbols(x), self._prepare_indexes(x.shape[1:]), True

  def decompress_device(self, strings, shape, channels=None, out_dtype=torch.float32):
    shape = tuple(int(s) for s in (shape.tolist() if hasattr(shape, "tolist") else shape))
    indexes = self._prepare_indexes(shape, channels)
    symbols = self._decode_symbols(strings, indexes, True)
    return self._dequantize(symbols, out_dtype)

  def decompress(self, strings, shape, channels=None):
    """entropy_models.py:684-704."""
    return self.decompress_device(strings, shape, channels)


class SymmetricConditional(EntropyModel):
  """Symmetric conditional entropy model (base class) — entropy_models.py:707-950."""

  def __init__(self, scale, scale_table, scale_bound=None, mean=None, indexes=None, **kwargs):
    super(SymmetricConditional, self).__init__(**kwargs)
    self._scale = scale
    self._scale_table = tuple(sorted(float(s) for s in scale_table))
    if any(s <= 0 for s in self.scale_table):
      raise ValueError("`scale_table` must be an iterable of positive numbers.")
    self._scale_bound = None if scale_bound is None else float(scale_bound)
    self._mean = mean      # entropy_models.py:734-735: same shape as the inputs, or None
    self._indexes = None
    if indexes is not None:
      if indexes.dtype != torch.int32:
        raise ValueError("`indexes` must have `int32` dtype.")
      self._indexes = indexes.contiguous()

  scale = property(lambda self: self._scale)
  scale_table = property(lambda self: self._scale_table)
  scale_bound = property(lambda self: self._scale_bound)
  mean = property(lambda self: self._mean)

  @property
  def indexes(self):
    if not self.built:
      self.build()
    return self._indexes

  def _standardized_cumulative_np(self, inputs):
    """float32 numpy version used for the (static) table construction."""
    raise NotImplementedError("Must inherit from SymmetricConditional.")

  def _standardized_cumulative(self, inputs):
    """torch version (likelihoods of the non-Gaussian subclasses)."""
    raise NotImplementedError("Must inherit from SymmetricConditional.")

  def _standardized_quantile(self, quantile):
    raise NotImplementedError("Must inherit from SymmetricConditional.")

  def _effective_bound(self):
    """entropy_models.py:822-825: lower bound applied to `scale` (0 = disabled)."""
    table0 = float(np.float32(self.scale_table[0]))
    if self.scale_bound is None:
      return table0
    return float(np.float32(self.scale_bound)) if self.scale_bound > 0 else 0.0

  def build_tables(self, device=None):
    """The static part of `build` (entropy_models.py:827-865, :883): CDF table per scale."""
    dev = device or torch.device("cuda", torch.cuda.current_device())
    table = np.array(self.scale_table, dtype=np.float64)
    multiplier = -self._standardized_quantile(self.tail_mass / 2)
    pmf_center = np.ceil(table * multiplier).astype(int)
    pmf_length = 2 * pmf_center + 1
    max_length = int(np.max(pmf_length))
    samples = np.abs(np.arange(max_length, dtype=int) - pmf_center[:, None]).astype(np.float32)
    samples_scale = np.array(self.scale_table, dtype=np.float32)[:, None]
    upper = self._standardized_cumulative_np((np.float32(.5) - samples) / samples_scale)
    lower = self._standardized_cumulative_np((np.float32(-.5) - samples) / samples_scale)
    pmf = (upper - lower).astype(np.float32)
    tail_mass = (np.float32(2) * lower[:, :1]).astype(np.float32)
    self._quantized_cdf = self._pmf_to_cdf(torch.from_numpy(pmf).to(dev),
                                           torch.from_numpy(tail_mass).to(dev), pmf_length,
                                           max_length)
    self._cdf_length = torch.as_tensor(pmf_length + 2, dtype=torch.int32).to(dev)
    self._offset = torch.as_tensor(-pmf_center, dtype=torch.int32).to(dev)
    self._table_dev = torch.as_tensor(np.array(self.scale_table, np.float32)).to(dev)
    self._tables_built = True

  def adopt_tables(self, other):
    """Share the (scale-independent) tables of another instance built on the same table."""
    assert other.scale_table == self.scale_table
    for n in ("_quantized_cdf", "_cdf_length", "_offset", "_table_dev"):
      setattr(self, n, getattr(other, n))
    self._tables_built = True

  def compute_indexes(self, scale):
    """entropy_models.py:869-881: (levels-1) - sum_k [max(scale, bound) <= table[k]]."""
    scale = scale.float().contiguous()
    idx = torch.empty(scale.shape, dtype=torch.int32, device=scale.device)
    _lib.check(_lib.lib().tfc_gaussian_symbols_indexes(
        ctypes.c_void_p(0), _lib.ptr(scale), _i64(scale.numel()), _lib.ptr(self._table_dev),
        _i32(len(self.scale_table)), _f32(self._effective_bound()), ctypes.c_void_p(0),
        _lib.ptr(idx), _lib.stream_ptr(scale.device)))
    return idx

  def build(self, input_shape=None):
    del input_shape
    dev = None
    if isinstance(self._scale, torch.Tensor):
      _lib.require_cuda(self._scale, "scale")
      dev = self._scale.device
    elif self._indexes is not None:
      dev = self._indexes.device
    if not getattr(self, "_tables_built", False):
      self.build_tables(dev)
    if self._indexes is None:
      self._indexes = self.compute_indexes(self._scale)
    self.built = True

  def _centered(self, x):
    """inputs minus `mean` (entropy_models.py:895-897, :917-919)."""
    return x if self._mean is None else x - self._mean.to(x.dtype)

  def _uncentered(self, values):
    return values if self._mean is None else values + self._mean.to(values.dtype)

  def _bounded_scale(self):
    b = self._effective_bound()
    s = self._scale.float()
    return math_ops.lower_bound(s, b) if b > 0 else s

  def __call__(self, inputs, training=False):
    x = self._check_inputs(inputs)
    if not self.built:
      self.build()
    x = self._centered(x)
    values = x + (torch.rand_like(x) - .5) if training else torch.floor(x + .5)
    a = torch.abs(values)
    s = self._bounded_scale()
    lik = self._standardized_cumulative((.5 - a) / s) - self._standardized_cumulative((-.5 - a) / s)
    if self.likelihood_bound > 0:
      lik = math_ops.lower_bound(lik, self.likelihood_bound)
    return self._uncentered(values), lik

  def _symbols_and_indexes(self, x):
    if not self.built:
      self.build()
    sym = torch.floor(self._centered(x) + .5).to(torch.int32)
    return sym, self._indexes, False

  def _dequantize(self, symbols, out_dtype=torch.float32):
    out = torch.empty(symbols.shape, dtype=out_dtype, device=symbols.device)
    _lib.check(_lib.lib().tfc_symbols_to_float(
        _lib.ptr(symbols), _i64(symbols.numel()), _lib.ptr(out),
        1 if out_dtype == torch.bfloat16 else 0, _lib.stream_ptr(symbols.device)))
    return self._uncentered(out)

  def _prepare_indexes(self, shape=None):
    del shape
    return self.indexes

  def decompress_device(self, strings, out_dtype=torch.float32):
    indexes = self._prepare_indexes()
    symbols = self._decode_symbols(strings, indexes, False)
    return self._dequantize(symbols, out_dtype)

  def decompress(self, strings):
    """entropy_models.py:941-950."""
    return self.decompress_device(strings)


class GaussianConditional(SymmetricConditional):
  """Conditional Gaussian entropy model — entropy_models.py:953-972.  Likelihood, quantisation
  and the scale -> index lookup each run as one fused CUDA kernel."""

  def _standardized_cumulative_np(self, inputs):
    const = np.float32(-(2 ** -0.5))
    return (np.float32(.5) * scipy.special.erfc(const * inputs.astype(np.float32))).astype(
        np.float32)

  def _standardized_cumulative(self, inputs):
    return .5 * torch.special.erfc(-(2 ** -0.5) * inputs)

  def _standardized_quantile(self, quantile):
    return scipy.stats.norm.ppf(quantile)

  def __call__(self, inputs, training=False):
    x = self._check_inputs(inputs)
    if not self.built:
      self.build()
    mode = 0
    x = self._centered(x).contiguous()
    if training:
      x = x + (torch.rand_like(x) - .5)
      mode = 1
    sigma = self._scale.float().contiguous()
    if sigma.shape != x.shape:
      raise ValueError("`scale` %s and inputs %s differ in shape" % (tuple(sigma.shape),
                                                                     tuple(x.shape)))
    if training and _wants_grad(x, sigma):
      lik = _GaussianLikelihood.apply(x, sigma, self._effective_bound(),
                                      max(self.likelihood_bound, 0.0))
      return self._uncentered(x), lik
    values = torch.empty_like(x) if not training else x
    lik = torch.empty_like(x)
    _lib.check(_lib.lib().tfc_gaussian_call(
        _lib.ptr(x), _lib.ptr(sigma), _i64(x.numel()), _f32(self._effective_bound()),
        _f32(max(self.likelihood_bound, 0.0)), mode,
        _lib.ptr(values) if not training else ctypes.c_void_p(0), _lib.ptr(lik),
        _lib.stream_ptr(x.device)))
    return self._uncentered(values), lik

  def _symbols_and_indexes(self, x):
    """Quantise and look the scale index up in one pass over (y, sigma)."""
    if not getattr(self, "_tables_built", False):
      self.build_tables(x.device)
    x = self._centered(x).contiguous()
    sym = torch.empty(x.shape, dtype=torch.int32, device=x.device)
    if self._indexes is not None:
      idx, sigma, idx_ptr = self._indexes, x, ctypes.c_void_p(0)
    else:
      sigma = self._scale.float().contiguous()
      idx = torch.empty(x.shape, dtype=torch.int32, device=x.device)
      idx_ptr = _lib.ptr(idx)
    _lib.check(_lib.lib().tfc_gaussian_symbols_indexes(
        _lib.ptr(x), _lib.ptr(sigma), _i64(x.numel()), _lib.ptr(self._table_dev),
        _i32(len(self.scale_table)), _f32(self._effective_bound()), _lib.ptr(sym), idx_ptr,
        _lib.stream_ptr(x.device)))
    self._indexes = idx
    self.built = True
    return sym, idx, False


class LogisticConditional(SymmetricConditional):
  """entropy_models.py:975-986."""

  def _standardized_cumulative_np(self, inputs):
    x = inputs.astype(np.float32)
    return (np.float32(1) / (np.float32(1) + np.exp(-x))).astype(np.float32)

  def _standardized_cumulative(self, inputs):
    return torch.sigmoid(inputs)

  def _standardized_quantile(self, quantile):
    return scipy.stats.logistic.ppf(quantile)


class LaplacianConditional(SymmetricConditional):
  """entropy_models.py:989-1001."""

  def _standardized_cumulative_np(self, inputs):
    x = inputs.astype(np.float32)
    e = np.exp(-np.abs(x)).astype(np.float32)
    return (np.where(x > 0, np.float32(2) - e, e) / np.float32(2)).astype(np.float32)

  def _standardized_cumulative(self, inputs):
    e = torch.exp(-torch.abs(inputs))
    return torch.where(inputs > 0, 2 - e, e) / 2

  def _standardized_quantile(self, quantile):
    return scipy.stats.laplace.ppf(quantile)
