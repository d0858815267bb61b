"""`lower_bound` / `upper_bound` — mirror of `tensorflow_compression/python/ops/math_ops.py`
(:77-184 forward, :30-74 gradients) as torch autograd functions."""
import torch

__all__ = ["lower_bound", "upper_bound"]


class _Bound(torch.autograd.Function):

  @staticmethod
  def forward(ctx, inputs, bound, lower, gradient):
    ctx.lower, ctx.gradient = lower, gradient
    if not isinstance(bound, torch.Tensor):
      # a Python number stays on the host (no H2D copy per call; CUDA-graph capturable)
      ctx.scalar = float(bound)
      ctx.save_for_backward(inputs)
      return torch.clamp(inputs, min=ctx.scalar) if lower else torch.clamp(inputs, max=ctx.scalar)
    ctx.scalar = None
    bound = bound.to(dtype=inputs.dtype, device=inputs.device)
    ctx.save_for_backward(inputs, bound)
    return torch.maximum(inputs, bound) if lower else torch.minimum(inputs, bound)

  @staticmethod
  def backward(ctx, grad):
    if ctx.scalar is not None:
      (inputs,), bound = ctx.saved_tensors, ctx.scalar
    else:
      inputs, bound = ctx.saved_tensors
    if ctx.gradient == "identity":
      return grad, None, None, None
    ok = inputs >= bound if ctx.lower else inputs <= bound
    if ctx.gradient == "identity_if_towards":
      # pass the gradient also where the update would move the input back inside the bound
      ok = ok | ((grad < 0) if ctx.lower else (grad > 0))
    return grad * ok.to(grad.dtype), None, None, None


def _check(gradient):
  if gradient not in ("disconnected", "identity", "identity_if_towards"):
    raise ValueError("Gradient can't be '{}'.".format(gradient))


def lower_bound(inputs, bound, gradient="identity_if_towards", name=None):
  """Same as `max(inputs, bound)`, gradient per math_ops.py:61-74."""
  del name
  _check(gradient)
  return _Bound.apply(inputs, bound, True, gradient)


def upper_bound(inputs, bound, gradient="identity_if_towards", name=None):
  del name
  _check(gradient)
  return _Bound.apply(inputs, bound, False, gradient)
