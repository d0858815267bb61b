"""Parameterizers — mirror of `tensorflow_compression/python/layers/parameterizers.py`.

A parameterizer owns the stored variable(s) of a parameter and produces the effective value.
Here it is an object with `init(value) -> stored` and `__call__(stored) -> effective`, both on
torch tensors, so the same object serves weight preparation (inference) and, through autograd,
the training step.
"""
import torch

from compression_b200.python.ops import math_ops
from compression_b200.python.ops import spectral_ops

__all__ = ["Parameterizer", "StaticParameterizer", "RDFTParameterizer",
           "NonnegativeParameterizer"]


class Parameterizer(object):
  """Base class (parameterizers.py:35-48)."""
  suffix = ""

  def init(self, value):
    return value

  def __call__(self, stored):
    return stored


class StaticParameterizer(Parameterizer):
  """A parameter fixed to its initial value (parameterizers.py:51-78)."""


class RDFTParameterizer(Parameterizer):
  """Kernel stored as its real DFT (parameterizers.py:81-144): effective kernel
  `reshape(M @ rdft)` with `M = irdft_matrix(kernel_support)`; 1x1 kernels are stored as is."""
  suffix = "_rdft"

  def __init__(self, dc=True):
    self.dc = bool(dc)

  _matrices = {}     # (support, device, dtype) -> irdft matrix on the device (built once)

  def _matrix(self, shape, like):
    key = (tuple(shape[:-2]), str(like.device), like.dtype)
    m = RDFTParameterizer._matrices.get(key)
    if m is None:
      m = torch.from_numpy(spectral_ops.irdft_matrix(shape[:-2])).to(like.device, like.dtype)
      RDFTParameterizer._matrices[key] = m
    return m if self.dc else m[:, 1:]

  def init(self, value):
    shape = tuple(value.shape)
    self.var_shape = shape
    if all(s == 1 for s in shape[:-2]):
      return value
    flat = value.reshape(-1, shape[-2] * shape[-1])
    return self._matrix(shape, value).t() @ flat

  def __call__(self, stored):
    shape = self.var_shape
    if all(s == 1 for s in shape[:-2]):
      return stored
    return (self._matrix(shape, stored) @ stored).reshape(shape)


class NonnegativeParameterizer(Parameterizer):
  """`max(v, sqrt(minimum + offset^2))^2 - offset^2` (parameterizers.py:147-200)."""
  prefix = "reparam_"

  def __init__(self, minimum=0, reparam_offset=2 ** -18):
    self.minimum = float(minimum)
    self.reparam_offset = float(reparam_offset)

  def init(self, value):
    pedestal = torch.tensor(self.reparam_offset ** 2, dtype=value.dtype, device=value.device)
    return torch.sqrt(torch.maximum(value + pedestal, pedestal))

  def __call__(self, stored):
    pedestal = self.reparam_offset ** 2
    bound = (self.minimum + self.reparam_offset ** 2) ** .5
    v = math_ops.lower_bound(stored, bound)
    return torch.square(v) - pedestal
