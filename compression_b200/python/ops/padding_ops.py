"""Padding geometry — mirror of `tensorflow_compression/python/ops/padding_ops.py:27-56`."""

__all__ = ["same_padding_for_kernel"]


def same_padding_for_kernel(shape, corr, strides_up=None):
  """Amount of (begin, end) zero padding per dimension that makes a 'valid' correlation
  (`corr=True`) or convolution produce an output the size of the (upsampled) input."""
  rank = len(shape)
  if strides_up is None:
    strides_up = rank * (1,)
  out = []
  for k, s in zip(shape, strides_up):
    begin, end = (k // 2, (k - 1) // 2) if corr else ((k - 1) // 2, k // 2)
    # with upsampling only every s-th padded sample is real
    out.append(((begin - 1) // s + 1, (end - 1) // s + 1))
  return out
