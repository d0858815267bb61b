"""SignalConv2D — mirror of `tensorflow_compression/python/layers/signal_conv.py`.

Same constructor arguments as the reference (`signal_conv.py:191-201`).  `call` reproduces the
reference's dispatch (`:709-814`): kernel flipping for odd supports (`:723-738`), "same"
padding geometry (`:742-758`, `ops/padding_ops.py`), correlation + downsampling with explicit
padding (`:487-524`) and upsampled convolution via the transposed-conv formulation
(`:629-707`) — and reduces each case to the one normalised form the CUDA kernels implement
(see `tfc_conv_desc` in include/tfc_b200.h).  Combinations the reference itself does not
implement raise NotImplementedError like `:393-401`; `channel_separable` (2-D and 1-D) runs on a
CUDA-core depthwise kernel and 1-D layers run as 2-D layers of height 1 (neither is on the
codecs' hot path); 3-D raises; `same_reflect` and `channels_first` are reduced on the host
side to the same kernels (reflection pre-pad by gather, layout conversion in and out).
"""
import ctypes
import math

import numpy as np
import torch

from compression_b200 import _lib
from compression_b200.python.layers import gdn as gdn_lib
from compression_b200.python.layers import parameterizers
from compression_b200.python.ops import padding_ops

__all__ = ["SignalConv1D", "SignalConv2D", "ConvDesc"]

ACT_NONE, ACT_RELU, ACT_GDN, ACT_IGDN = 0, 1, 2, 3


class ConvDesc(ctypes.Structure):
  """ctypes image of tfc_conv_desc."""
  _fields_ = [(n, ctypes.c_int32) for n in (
      "n", "h", "w", "cin", "ho", "wo", "cout", "kh", "kw", "sd", "su", "pad_top", "pad_left",
      "act", "in_dtype", "out_dtype", "impl", "flags")]

CONV_WEIGHTS_STAGED = 1


def _dtype_code(dt):
  if dt == torch.float32:
    return 0
  if dt == torch.bfloat16:
    return 1
  if dt == torch.uint8:
    return 2   # TFC_U8: an 8-bit image (input: byte / 255; output: quantize_image)
  raise ValueError("unsupported dtype %s (float32, bfloat16 or uint8 images)" % dt)


def quantize_image(image):
  """examples/bmshj2018.py:60-63: saturate_cast<uint8>(round(image * 255))."""
  return torch.clamp(torch.round(image * 255), 0, 255).to(torch.uint8)


def variance_scaling_init(shape, generator=None, device="cpu", dtype=torch.float32):
  """tf.initializers.variance_scaling() defaults (scale 1, fan_in, truncated normal), the
  reference's `kernel_initializer` default (`signal_conv.py:196`)."""
  kh, kw, cin, _ = shape
  fan_in = kh * kw * cin
  stddev = math.sqrt(1.0 / max(1.0, fan_in)) / .87962566103423978
  t = torch.empty(shape, dtype=torch.float32)
  torch.nn.init.trunc_normal_(t, mean=0.0, std=stddev, a=-2 * stddev, b=2 * stddev,
                              generator=generator)
  return t.to(device=device, dtype=dtype)


class SignalConv2D(object):
  """2-D signal convolution layer (channels_last)."""

  _rank = 2

  def __init__(self, filters, kernel_support, corr=False, strides_down=1, strides_up=1,
               padding="valid", extra_pad_end=True, channel_separable=False,
               data_format="channels_last", activation=None, use_bias=False, use_explicit=True,
               kernel_initializer=None, bias_initializer=None, kernel_regularizer=None,
               bias_regularizer=None,
               kernel_parameterizer=parameterizers.RDFTParameterizer(),
               bias_parameterizer=None, name=None, dtype=torch.float32, impl=0):
    def pair(v, what):
      v = tuple(int(x) for x in (v if isinstance(v, (tuple, list)) else (v,) * 2))
      if len(v) != 2 or any(x < 1 for x in v):
        raise ValueError("`%s` must be a positive int or a pair of them" % what)
      return v
    self.filters = int(filters)
    self.kernel_support = pair(kernel_support, "kernel_support")
    self.corr = bool(corr)
    self.strides_down = pair(strides_down, "strides_down")
    self.strides_up = pair(strides_up, "strides_up")
    self.padding = str(padding).lower()
    if self.padding not in ("valid", "same_zeros", "same_reflect"):
      raise ValueError("Unsupported padding mode: '{}'".format(padding))
    self.extra_pad_end = bool(extra_pad_end)
    self.channel_separable = bool(channel_separable)
    self.data_format = str(data_format)
    if self.data_format not in ("channels_first", "channels_last"):
      raise ValueError("Unknown data format: '{}'.".format(data_format))
    self.activation = activation
    self.use_bias = bool(use_bias)
    self.use_explicit = bool(use_explicit)
    self.kernel_initializer = kernel_initializer
    self.bias_initializer = bias_initializer
    self.kernel_regularizer = kernel_regularizer
    self.bias_regularizer = bias_regularizer
    # each layer needs its own parameterizer state (the reference's default object is
    # stateless; ours remembers the variable shape)
    self.kernel_parameterizer = (type(kernel_parameterizer)(kernel_parameterizer.dc)
                                 if isinstance(kernel_parameterizer,
                                               parameterizers.RDFTParameterizer)
                                 else kernel_parameterizer)
    self.bias_parameterizer = bias_parameterizer
    self.name = name
    self.dtype = dtype
    self.impl = int(impl)
    self.out_dtype = torch.float32  # dtype of the produced activations (float32 | bfloat16)
    self.built = False
    self._eff = None
    self._geom = {}
    # (device, stream) -> [buffer, effective weights staged in it, GDN params, kernel path that
    # staged them] — the repacked layout belongs to the path (tfc_signal_conv2d_path)
    self._ws = {}

  # ---- reference's capability gaps ------------------------------------------------------
  def _raise_notimplemented(self):
    raise NotImplementedError(
        "The provided combination of SignalConv arguments is not currently implemented "
        "(kernel_support={}, corr={}, strides_down={}, strides_up={}, channel_separable={}, "
        "filters={}).".format(self.kernel_support, self.corr, self.strides_down,
                              self.strides_up, self.channel_separable, self.filters))

  # ---- variables ---------------------------------------------------------------------------
  def build(self, input_channels, device="cuda", generator=None):
    cin = int(input_channels)
    kshape = self.kernel_support + (cin, self.filters)
    # channel_separable (signal_conv.py:405-411): `filters` per input channel, cin * filters outputs
    self.out_channels = cin * self.filters if self.channel_separable else self.filters
    if self.channel_separable and max(self.strides_up) > 1 and self.filters != 1:
      self._raise_notimplemented()           # the reference's transposed depthwise op (:590-596)
    init = (self.kernel_initializer(kshape) if self.kernel_initializer is not None
            else variance_scaling_init(kshape, generator))
    init = torch.as_tensor(init, dtype=self.dtype).to(device)
    kp = self.kernel_parameterizer
    if kp is None:
      self.kernel_var = init          # variable "kernel"
    else:
      self.kernel_var = kp.init(init)  # variable "kernel_rdft"
    if self.use_bias:
      b0 = (self.bias_initializer((self.out_channels,)) if self.bias_initializer is not None
            else torch.zeros(self.out_channels))
      b0 = torch.as_tensor(b0, dtype=self.dtype).to(device)
      bp = self.bias_parameterizer
      self.bias_var = b0 if bp is None else bp.init(b0)
    else:
      self.bias_var = None
    if isinstance(self.activation, gdn_lib.GDN) and not self.activation.built:
      self.activation.build(self.out_channels, device)
    self.input_channels = cin
    self.built = True
    self._eff = None
    self._geom = {}

  def invalidate(self):
    """Call after changing kernel_var / bias_var so the effective weights are recomputed."""
    self._eff = None
    if isinstance(self.activation, gdn_lib.GDN):
      self.activation.invalidate()

  @property
  def kernel(self):
    kp = self.kernel_parameterizer
    return self.kernel_var if kp is None else kp(self.kernel_var)

  @property
  def bias(self):
    if self.bias_var is None:
      return None
    bp = self.bias_parameterizer
    return self.bias_var if bp is None else bp(self.bias_var)

  # ---- geometry: the reference's dispatch reduced to the normalised form ---------------------
  def _plan(self, mode=None):
    """Returns (flip, pad_begin (top, left), out_len_fn) for the current arguments, the input
    being channels_last.  `mode` overrides the padding mode (the pre-padded same_reflect path
    runs as "valid" / "same_zeros" over the reflected image, see `_reflect`)."""
    mode = mode or self.padding
    if mode == "same_reflect":
      raise NotImplementedError("same_reflect is planned on the pre-padded input: use __call__")
    if self.strides_down[0] != self.strides_down[1] or self.strides_up[0] != self.strides_up[1]:
      raise NotImplementedError("anisotropic strides are not built")
    k = self.kernel_support
    su, sd = self.strides_up[0], self.strides_down[0]
    corr, flip = self.corr, False
    odd = all(s % 2 == 1 for s in k)
    if not corr and su == 1 and odd:        # signal_conv.py:723-728
      corr, flip = True, True
    elif corr and su != 1 and odd:          # signal_conv.py:733-738
      corr, flip = False, True
    if mode == "valid":
      padding = [(0, 0), (0, 0)]
    else:
      padding = padding_ops.same_padding_for_kernel(k, corr, (su, su))
    if corr and su == 1:
      # conv2d / convolution with (explicit or zero) padding, then stride sd (:487-524,444-485)
      pad_begin = (padding[0][0], padding[1][0])
      def out_len(l, i):
        return (l + padding[i][0] + padding[i][1] - k[i]) // sd + 1
      return flip, pad_begin, out_len
    if not corr:
      # conv2d_backprop_input with explicit padding p == zero-stuffed correlation with the
      # flipped kernel and padding k-1-p (:629-707); prepadding is zero in EXPLICIT mode
      if mode == "valid":
        p = [(k[0] - 1, k[0] - 1), (k[1] - 1, k[1] - 1)]
      else:
        p = [(k[i] // 2, (k[i] - 1) // 2) for i in range(2)]
      pad_begin = (k[0] - 1 - p[0][0], k[1] - 1 - p[1][0])
      def out_len(l, i):
        full = l * su + (k[i] - 1) - (p[i][0] + p[i][1])
        if not self.extra_pad_end:
          full -= su - 1
        return (full - 1) // sd + 1
      return (not flip), pad_begin, out_len   # true convolution == correlation, kernel flipped
    self._raise_notimplemented()

  def compute_output_shape(self, input_shape):
    """signal_conv.py:816-845."""
    n, h, w, _ = input_shape
    out = []
    for i, s in enumerate((h, w)):
      s = s * self.strides_up[i] if self.extra_pad_end else (s - 1) * self.strides_up[i] + 1
      if self.padding == "valid":
        s -= self.kernel_support[i] - 1
      out.append((s - 1) // self.strides_down[i] + 1)
    channels = getattr(self, "out_channels", None)
    if channels is None:
      channels = input_shape[-1] * self.filters if self.channel_separable else self.filters
    return (n, out[0], out[1], channels)

  def _activation_code(self):
    a = self.activation
    if a is None:
      return ACT_NONE, None
    if isinstance(a, gdn_lib.GDN):
      return (ACT_IGDN if a.inverse else ACT_GDN), None
    if a in (torch.relu, torch.nn.functional.relu) or a == "relu":
      return ACT_RELU, None
    return ACT_NONE, a  # arbitrary callable: applied after the kernel

  def effective_weights(self):
    """(kernel [kh,kw,cin,cout] in correlation form, bias|None), float32 contiguous; cached."""
    if self._eff is None:
      flip, _, _ = self._plan("same_zeros" if self.padding == "same_reflect" else None)
      w = self.kernel.float()
      if flip:
        w = torch.flip(w, dims=(0, 1))
      b = self.bias
      self._eff = (w.contiguous(), None if b is None else b.float().contiguous())
    return self._eff

  # ---- forward -----------------------------------------------------------------------------
  def _reflect(self, x):
    """same_reflect (signal_conv.py:744-758): the reference pre-pads with tf.pad(REFLECT) and
    then runs its VALID ops; for upsampled convolutions the pre-padding is folded into the
    explicit padding of conv2d_backprop_input (`:651-666`), which equals the same_zeros result
    over the reflected image cropped by `prepad * strides_up`.  Returns (padded x, mode to plan
    with, crop (top, left) of the output or None)."""
    k, su = self.kernel_support, self.strides_up[0]
    corr = self.corr
    odd = all(s % 2 == 1 for s in k)
    if not corr and su == 1 and odd:
      corr = True
    elif corr and su != 1 and odd:
      corr = False
    pad = padding_ops.same_padding_for_kernel(k, corr, (su, su))
    n, h, w, _ = x.shape
    if max(pad[0]) >= h or max(pad[1]) >= w:
      raise ValueError("same_reflect needs the image to be larger than the padding")

    def idx(length, lo, hi):
      i = torch.arange(-lo, length + hi, device=x.device)
      i = torch.where(i < 0, -i, i)
      return torch.where(i >= length, 2 * (length - 1) - i, i)

    x = x.index_select(1, idx(h, *pad[0])).index_select(2, idx(w, *pad[1]))
    if corr and su == 1:
      return x, "valid", None
    if self.strides_down[0] != 1:
      self._raise_notimplemented()
    return x, "same_zeros", (pad[0][0] * su, pad[1][0] * su)

  def _call_separable(self, x, pad_begin, ho, wo, crop, want_shape, out_dtype):
    """channel_separable=True: the depthwise kernel (float32), then the activation as its own
    op — the reference's structure (tf.nn.depthwise_conv2d_native, bias_add, activation)."""
    n, h, w, cin = x.shape
    if x.dtype == torch.uint8:
      x = x.float() / 255
    x = x.float().contiguous()
    wgt, bias = self.effective_weights()
    d = ConvDesc(n=n, h=h, w=w, cin=cin, ho=ho, wo=wo, cout=self.out_channels,
                 kh=self.kernel_support[0], kw=self.kernel_support[1],
                 sd=self.strides_down[0], su=self.strides_up[0], pad_top=pad_begin[0],
                 pad_left=pad_begin[1], act=ACT_NONE, in_dtype=_dtype_code(torch.float32),
                 out_dtype=_dtype_code(torch.float32), impl=0, flags=0)
    out = torch.empty((n, ho, wo, self.out_channels), dtype=torch.float32, device=x.device)
    _lib.check(_lib.lib().tfc_depthwise_conv2d(
        ctypes.byref(d), _lib.ptr(x), _lib.ptr(wgt), _lib.ptr(bias), _lib.ptr(out),
        _lib.stream_ptr(x.device)))
    if crop is not None:
      out = out[:, crop[0]:crop[0] + want_shape[1], crop[1]:crop[1] + want_shape[2], :].contiguous()
    if self.activation is not None:
      a = self.activation
      out = torch.relu(out) if a == "relu" else a(out)
    out_dtype = out_dtype or self.out_dtype
    if out_dtype == torch.uint8:
      return quantize_image(out)
    return out if out.dtype == out_dtype else out.to(out_dtype)

  def __call__(self, inputs, out_dtype=None):
    """inputs: [batch, height, width, channels] (channels_last) or [batch, channels, height,
    width] (channels_first; converted on the way in and out — the kernels are channels_last)."""
    if self.data_format == "channels_first":
      if inputs.dim() != 4:
        raise ValueError("SignalConv2D expects a 4-D input")
      out = self._call_nhwc(inputs.permute(0, 2, 3, 1), out_dtype)
      return out.permute(0, 3, 1, 2).contiguous()
    return self._call_nhwc(inputs, out_dtype)

  def _call_nhwc(self, inputs, out_dtype=None):
    x = inputs
    _lib.require_cuda(x, "inputs")
    if x.dim() != 4:
      raise ValueError("SignalConv2D expects [batch, height, width, channels]")
    if not self.built:
      self.build(x.shape[-1], x.device)
    if x.shape[-1] != self.input_channels:
      raise ValueError("layer built for %d input channels, got %d"
                       % (self.input_channels, x.shape[-1]))
    if x.dtype not in (torch.float32, torch.bfloat16, torch.uint8):
      x = x.float()
    if x.dtype == torch.uint8 and self.padding == "same_reflect":
      x = x.float() / 255
    want_shape = tuple(self.compute_output_shape(x.shape))
    mode, crop = None, None
    if self.padding == "same_reflect":
      x, mode, crop = self._reflect(x)
    x = x.contiguous()
    _, pad_begin, out_len = self._plan(mode)
    n, h, w, cin = x.shape
    ho, wo = out_len(h, 0), out_len(w, 1)
    if crop is None:
      assert (n, ho, wo, self.out_channels) == want_shape, \
          "internal: output geometry disagrees with compute_output_shape"
    if self.channel_separable:
      return self._call_separable(x, pad_begin, ho, wo, crop, want_shape, out_dtype)
    act, post = self._activation_code()
    out_dtype = out_dtype or self.out_dtype
    wgt, bias = self.effective_weights()
    beta = gamma = None
    if act in (ACT_GDN, ACT_IGDN):
      beta, gamma = self.activation.effective()
    d = ConvDesc(n=n, h=h, w=w, cin=cin, ho=ho, wo=wo, cout=self.filters,
                 kh=self.kernel_support[0], kw=self.kernel_support[1],
                 sd=self.strides_down[0], su=self.strides_up[0], pad_top=pad_begin[0],
                 pad_left=pad_begin[1], act=act, in_dtype=_dtype_code(x.dtype),
                 out_dtype=_dtype_code(out_dtype), impl=self.impl, flags=0)
    L = _lib.lib()
    # kernel path and workspace size of this geometry (two C calls, cached per layer)
    gkey = (n, h, w, d.in_dtype, d.out_dtype, mode, post is not None, self.impl)
    geom = self._geom.get(gkey)
    if geom is None:
      L.tfc_signal_conv2d_workspace.restype = ctypes.c_size_t
      path = int(L.tfc_signal_conv2d_path(ctypes.byref(d)))
      # 8-bit images are read / written by the thin-input / thin-output kernels only (paths 4,
      # 3); anywhere else the scaling and the quantisation are torch ops around a float32 call
      scale_before = quantize_after = False
      if x.dtype == torch.uint8 and path != 4:
        scale_before, d.in_dtype = True, _dtype_code(torch.float32)
        path = int(L.tfc_signal_conv2d_path(ctypes.byref(d)))
      if out_dtype == torch.uint8 and (path not in (3, 6) or post is not None):
        quantize_after, d.out_dtype = True, _dtype_code(torch.float32)
        path = int(L.tfc_signal_conv2d_path(ctypes.byref(d)))
      need = int(L.tfc_signal_conv2d_workspace(ctypes.byref(d)))
      geom = self._geom[gkey] = (path, need, scale_before, quantize_after)
    path, need, scale_before, quantize_after = geom
    if scale_before:
      x = x.float() / 255
      d.in_dtype = _dtype_code(torch.float32)
    if quantize_after:
      out_dtype = torch.float32
      d.out_dtype = _dtype_code(torch.float32)
    # One workspace per CUDA stream (concurrent batches on different streams must not share
    # scratch); the repacked weights stay valid in it until the effective weights change.
    stream = _lib.stream_ptr(x.device)
    key = (x.device.index, stream.value)
    rec = self._ws.get(key)
    if rec is None or rec[0].numel() < need:
      rec = [torch.empty((max(need, 256),), dtype=torch.uint8, device=x.device), None, None, -1]
      self._ws[key] = rec
    gdn_params = self.activation._cache if act in (ACT_GDN, ACT_IGDN) else None
    # The staged weights are reusable only by the kernel path that wrote them: the path depends
    # on the input dtype and on h % sd / w % sd as well as on the layer, every path has its own
    # layout, and the CUDA-core path (0) uses the workspace as scratch.
    if (path > 0 and rec[3] == path and rec[1] is self._eff and rec[2] is gdn_params
        and rec[1] is not None):
      d.flags |= CONV_WEIGHTS_STAGED
    out = torch.empty((n, ho, wo, self.filters), dtype=out_dtype, device=x.device)
    _lib.check(L.tfc_signal_conv2d(
        ctypes.byref(d), _lib.ptr(x), _lib.ptr(wgt), _lib.ptr(bias), _lib.ptr(beta),
        _lib.ptr(gamma), _lib.ptr(out), _lib.ptr(rec[0]), ctypes.c_size_t(rec[0].numel()),
        stream))
    rec[1], rec[2], rec[3] = (self._eff, gdn_params, path) if path > 0 else (None, None, -1)
    if crop is not None:
      out = out[:, crop[0]:crop[0] + want_shape[1], crop[1]:crop[1] + want_shape[2], :].contiguous()
      assert tuple(out.shape) == want_shape
    if post is not None:
      out = post(out)
    if quantize_after:
      out = quantize_image(out)
    return out


class SignalConv1D(SignalConv2D):
  """1-D signal convolution layer (signal_conv.py:848-863): inputs [batch, length, channels]
  (channels_last) or [batch, channels, length].  Runs as the 2-D layer over an image of height 1
  with a kernel of height 1 — the reference inserts the same extra dimension where TF has no 1-D
  op (:459-471, :498-516, :565-589); an upsampled layer computes `strides_up` rows of which the
  first is the 1-D result.  Not on the codecs' hot path."""

  _rank = 1

  def __init__(self, filters, kernel_support, corr=False, strides_down=1, strides_up=1, **kwargs):
    def one(v, what):
      v = tuple(v) if isinstance(v, (tuple, list)) else (v,)
      if len(v) != 1:
        raise ValueError("`%s` must be an int or a 1-tuple" % what)
      return int(v[0])
    k, sd, su = one(kernel_support, "kernel_support"), one(strides_down, "strides_down"), \
        one(strides_up, "strides_up")
    super(SignalConv1D, self).__init__(filters, (1, k), corr=corr, strides_down=(sd, sd),
                                       strides_up=(su, su), **kwargs)

  def compute_output_shape(self, input_shape):
    if len(input_shape) == 3:
      n, l, c = input_shape
      full = super(SignalConv1D, self).compute_output_shape((n, 1, l, c))
      return (n, full[2], full[3])
    return super(SignalConv1D, self).compute_output_shape(input_shape)

  def __call__(self, inputs, out_dtype=None):
    if inputs.dim() != 3:
      raise ValueError("SignalConv1D expects a 3-D input")
    if self.data_format == "channels_first":
      out = self._call_nhwc(inputs.permute(0, 2, 1)[:, None], out_dtype)
      return out[:, 0].permute(0, 2, 1).contiguous()
    out = self._call_nhwc(inputs[:, None], out_dtype)
    return out[:, 0].contiguous()
