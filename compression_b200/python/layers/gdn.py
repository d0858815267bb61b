"""Generalized divisive normalization — mirror of `tensorflow_compression/python/layers/gdn.py`.

`y[i] = x[i] / sqrt(beta[i] + sum_j gamma[j, i] * x[j]^2)`  (IGDN multiplies instead).
When a GDN object is the `activation` of a SignalConv2D the conv kernel applies it in its
epilogue; called on its own it runs the stand-alone kernel (`tfc_gdn`).
"""
import ctypes

import torch

from compression_b200 import _lib
from compression_b200.python.layers import parameterizers

__all__ = ["GDN"]

_default_beta_param = parameterizers.NonnegativeParameterizer(minimum=1e-6)
_default_gamma_param = parameterizers.NonnegativeParameterizer()


class GDN(object):
  """gdn.py:37-209.  Supported here: channels_last, rectify=False (what the codecs use)."""

  def __init__(self, inverse=False, rectify=False, gamma_init=.1, data_format="channels_last",
               beta_parameterizer=_default_beta_param,
               gamma_parameterizer=_default_gamma_param, name=None, dtype=torch.float32):
    self.inverse = bool(inverse)
    self.rectify = bool(rectify)
    self.gamma_init = float(gamma_init)
    self.data_format = str(data_format)
    self.beta_parameterizer = beta_parameterizer
    self.gamma_parameterizer = gamma_parameterizer
    self.name = name
    self.dtype = dtype
    self.built = False
    if self.data_format not in ("channels_first", "channels_last"):
      raise ValueError("Unknown data format: '{}'.".format(self.data_format))
    if self.rectify or self.data_format != "channels_last":
      raise NotImplementedError("GDN: only channels_last, rectify=False is built")

  def build(self, num_channels, device="cuda"):
    c = int(num_channels)
    beta0 = torch.ones(c, dtype=self.dtype, device=device)
    gamma0 = torch.eye(c, dtype=self.dtype, device=device) * self.gamma_init
    bp, gp = self.beta_parameterizer, self.gamma_parameterizer
    self.reparam_beta = bp.init(beta0) if bp is not None else beta0
    self.reparam_gamma = gp.init(gamma0) if gp is not None else gamma0
    self.num_channels = c
    self.built = True
    self._cache = None

  def invalidate(self):
    self._cache = None

  @property
  def beta(self):
    return self.effective()[0]

  @property
  def gamma(self):
    return self.effective()[1]

  def effective(self):
    """(beta [C], gamma [C_in(k), C_out(c)]) after reparameterisation, float32 contiguous."""
    if self._cache is None:
      bp, gp = self.beta_parameterizer, self.gamma_parameterizer
      beta = bp(self.reparam_beta) if bp is not None else self.reparam_beta
      gamma = gp(self.reparam_gamma) if gp is not None else self.reparam_gamma
      self._cache = (beta.float().contiguous(), gamma.float().contiguous())
    return self._cache

  def __call__(self, inputs, out_dtype=None):
    x = inputs
    _lib.require_cuda(x, "inputs")
    if not self.built:
      self.build(x.shape[-1], x.device)
    if x.shape[-1] != self.num_channels:
      raise ValueError("GDN built for %d channels, got %d" % (self.num_channels, x.shape[-1]))
    x = x.float().contiguous()
    out_dtype = out_dtype or torch.float32
    out = torch.empty(x.shape, dtype=out_dtype, device=x.device)
    beta, gamma = self.effective()
    pixels = x.numel() // self.num_channels
    _lib.check(_lib.lib().tfc_gdn(
        _lib.ptr(x), ctypes.c_int64(pixels), ctypes.c_int32(self.num_channels), _lib.ptr(beta),
        _lib.ptr(gamma), int(self.inverse), 0, 1 if out_dtype == torch.bfloat16 else 0,
        _lib.ptr(out), _lib.stream_ptr(x.device)))
    return out
