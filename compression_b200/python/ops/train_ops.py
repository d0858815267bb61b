"""Training-side ops of the bls2017 step (`examples/bls2017.py:130-202`): the convolution as a
torch.autograd.Function over the CUDA kernels, flat parameter buckets and the fused Adam update.

forward   tfc_signal_conv2d (tcgen05 implicit GEMM, no activation, float32 out)
backward  input gradient  = tfc_signal_conv2d with strides_down/strides_up swapped, the kernel
                            flipped and transposed, padding k-1-pad (a strided correlation's
                            gradient is a transposed convolution and vice versa)
          filter gradient = tfc_signal_conv2d_wgrad
          bias gradient   = sum over pixels
Everything else in the step (GDN's 1x1 contraction, the density network of the entropy
bottleneck, the loss) is differentiated by autograd over torch ops.
"""
import ctypes

import torch

from compression_b200 import _lib
from compression_b200.python.layers.signal_conv import ConvDesc, _dtype_code

__all__ = ["signal_conv2d", "FlatBucket", "adam_step"]

_ws = {}


def _workspace(need, device):
  key = (device.index, _lib.stream_ptr(device).value)
  buf = _ws.get(key)
  if buf is None or buf.numel() < need:
    buf = _ws[key] = torch.empty((max(need, 256),), dtype=torch.uint8, device=device)
  return buf


def _conv_raw(x, w, bias, kh, kw, sd, su, pad_top, pad_left, ho, wo, out_dtype):
  """Normalised convolution of tfc_b200.h on raw tensors (no activation)."""
  x = x.contiguous()
  w = w.float().contiguous()
  n, h, wd, cin = x.shape
  cout = w.shape[3]
  d = ConvDesc(n=n, h=h, w=wd, cin=cin, ho=ho, wo=wo, cout=cout, kh=kh, kw=kw, sd=sd, su=su,
               pad_top=pad_top, pad_left=pad_left, act=0, in_dtype=_dtype_code(x.dtype),
               out_dtype=_dtype_code(out_dtype), impl=0, flags=0)
  L = _lib.lib()
  L.tfc_signal_conv2d_workspace.restype = ctypes.c_size_t
  ws = _workspace(int(L.tfc_signal_conv2d_workspace(ctypes.byref(d))), x.device)
  out = torch.empty((n, ho, wo, cout), dtype=out_dtype, device=x.device)
  _lib.check(L.tfc_signal_conv2d(
      ctypes.byref(d), _lib.ptr(x), _lib.ptr(w), _lib.ptr(bias), ctypes.c_void_p(0),
      ctypes.c_void_p(0), _lib.ptr(out), _lib.ptr(ws), ctypes.c_size_t(ws.numel()),
      _lib.stream_ptr(x.device)))
  return out


class _SignalConv(torch.autograd.Function):

  @staticmethod
  def forward(ctx, x, w, bias, geom):
    kh, kw, sd, su, pt, pl, ho, wo = geom
    if x.dtype not in (torch.float32, torch.bfloat16):
      x = x.float()
    b = None if bias is None else bias.float().contiguous()
    out = _conv_raw(x, w, b, kh, kw, sd, su, pt, pl, ho, wo, torch.float32)
    ctx.save_for_backward(x, w)
    ctx.geom, ctx.has_bias = geom, bias is not None
    return out

  @staticmethod
  def backward(ctx, dout):
    x, w = ctx.saved_tensors
    kh, kw, sd, su, pt, pl, ho, wo = ctx.geom
    n, h, wd, cin = x.shape
    dy = dout.to(torch.bfloat16).contiguous()
    dx = dw = db = None
    if ctx.needs_input_grad[0]:
      wt = torch.flip(w, dims=(0, 1)).transpose(2, 3).contiguous()     # [kh,kw,cout,cin]
      # the gradient leaves the kernel in the dtype of the activation it belongs to (bfloat16
      # between the layers of a transform): no cast pass after it
      dx = _conv_raw(dy, wt, None, kh, kw, su, sd, kh - 1 - pt, kw - 1 - pl, h, wd, x.dtype)
    if ctx.needs_input_grad[1]:
      xb = x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16)
      xb = xb.contiguous()
      dw = torch.empty(tuple(w.shape), dtype=torch.float32, device=x.device)
      d = ConvDesc(n=n, h=h, w=wd, cin=cin, ho=ho, wo=wo, cout=w.shape[3], kh=kh, kw=kw, sd=sd,
                   su=su, pad_top=pt, pad_left=pl, act=0, in_dtype=1, out_dtype=0, impl=0,
                   flags=0)
      _lib.check(_lib.lib().tfc_signal_conv2d_wgrad(
          ctypes.byref(d), _lib.ptr(xb), _lib.ptr(dy), _lib.ptr(dw), _lib.stream_ptr(x.device)))
      dw = dw.to(w.dtype)
    if ctx.has_bias and ctx.needs_input_grad[2]:
      do = dout.float().contiguous()
      db = torch.empty((do.shape[-1],), dtype=torch.float32, device=do.device)
      _lib.check(_lib.lib().tfc_column_sums(
          _lib.ptr(do), ctypes.c_int64(do.numel() // do.shape[-1]), ctypes.c_int32(do.shape[-1]),
          _lib.ptr(db), _lib.stream_ptr(do.device)))
    return dx, dw, db, None


def signal_conv2d(layer, x):
  """Differentiable SignalConv2D *without* its activation: `layer` supplies the geometry
  (signal_conv.py:709-782) and the parameterised kernel / bias (with their autograd graph);
  returns float32 [n, ho, wo, filters]."""
  flip, pad_begin, out_len = layer._plan()
  w = layer.kernel
  if flip:
    w = torch.flip(w, dims=(0, 1))
  n, h, wd, _ = x.shape
  geom = (layer.kernel_support[0], layer.kernel_support[1], layer.strides_down[0],
          layer.strides_up[0], pad_begin[0], pad_begin[1], out_len(h, 0), out_len(wd, 1))
  return _SignalConv.apply(x, w, layer.bias, geom)


class FlatBucket(object):
  """Parameters as views of ONE flat float32 buffer, gradients as views of another: the
  gradient all-reduce is a single NCCL call and the optimizer a single kernel."""

  def __init__(self, tensors, device):
    self.shapes = [tuple(t.shape) for t in tensors]
    self.sizes = [int(t.numel()) for t in tensors]
    total = sum(self.sizes)
    self.param = torch.empty((total,), dtype=torch.float32, device=device)
    self.grad = torch.zeros((total,), dtype=torch.float32, device=device)
    self.m = torch.zeros_like(self.param)
    self.v = torch.zeros_like(self.param)
    self.step = 0
    self.views = []
    at = 0
    for t, n, shape in zip(tensors, self.sizes, self.shapes):
      self.param[at:at + n].copy_(t.detach().float().reshape(-1))
      v = self.param[at:at + n].view(shape).requires_grad_()
      v.grad = self.grad[at:at + n].view(shape)
      self.views.append(v)
      at += n

  def zero_grad(self):
    self.grad.zero_()
    at = 0
    for v, n, shape in zip(self.views, self.sizes, self.shapes):
      v.grad = self.grad[at:at + n].view(shape)     # autograd may have replaced the view
      at += n

  def all_reduce(self):
    """Sum over ranks (the 1/world factor goes into the optimizer's grad_scale)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
      dist.all_reduce(self.grad, op=dist.ReduceOp.SUM)
      return 1.0 / dist.get_world_size()
    return 1.0


def adam_step(bucket, lr, grad_scale=1.0, beta1=0.9, beta2=0.999, eps=1e-8):
  """tf.train.AdamOptimizer(learning_rate=lr) on the whole bucket (bls2017.py:173-178)."""
  bucket.step += 1
  _lib.check(_lib.lib().tfc_adam_step(
      _lib.ptr(bucket.param), _lib.ptr(bucket.grad), _lib.ptr(bucket.m), _lib.ptr(bucket.v),
      ctypes.c_int64(bucket.param.numel()), ctypes.c_float(lr), ctypes.c_float(beta1),
      ctypes.c_float(beta2), ctypes.c_float(eps), ctypes.c_int64(bucket.step),
      ctypes.c_float(grad_scale), _lib.stream_ptr(bucket.param.device)))
