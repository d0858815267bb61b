"""bls2017 training step: convolution gradients through the CUDA kernels against
torch autograd on a float64 restatement of the same layer, the fused Adam kernel against
tf.train.AdamOptimizer's formulas, and the whole step against a float32 torch restatement."""
import ctypes

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _static(value):
  return lambda shape: torch.as_tensor(value)


def _ref_conv(x, k, b, corr, sd, su):
  """Differentiable float64 restatement of SignalConv2D (same_zeros, odd kernels), following
  signal_conv.py:709-782 like oracle/tfc_np.py:signal_conv2d.  x NHWC, k [kh,kw,cin,cout]."""
  kh, kw = k.shape[:2]
  xn = x.permute(0, 3, 1, 2)
  if corr and su == 1:
    out = F.conv2d(F.pad(xn, (kw // 2, kw // 2, kh // 2, kh // 2)), k.permute(3, 2, 0, 1), stride=sd)
  else:
    assert not corr and sd == 1
    n, _, h, w = xn.shape
    p = kh // 2
    full = F.conv_transpose2d(xn, k.permute(2, 3, 0, 1), stride=su)
    full = F.pad(full, (0, su, 0, su))
    out = full[:, :, p:p + h * su, p:p + w * su]
  out = out.permute(0, 2, 3, 1)
  return out + b if b is not None else out


CASES = [
    # cin, cout, k, corr, sd, su, H, W, input dtype
    (64, 64, 5, True, 2, 1, 16, 24, torch.bfloat16),
    (3, 128, 9, True, 4, 1, 32, 32, torch.float32),      # first analysis layer of bls2017
    (128, 128, 5, True, 2, 1, 16, 16, torch.bfloat16),
    (128, 128, 5, False, 1, 2, 8, 12, torch.bfloat16),
    (128, 3, 9, False, 1, 4, 8, 8, torch.bfloat16),      # last synthesis layer of bls2017
    (192, 192, 3, True, 1, 1, 10, 14, torch.bfloat16),
    # filter gradient on tcgen05 (MN-major operands): M side 128, N side 64 / 128 / 192
    (128, 192, 3, True, 1, 1, 10, 14, torch.bfloat16),
    (128, 64, 5, True, 2, 1, 12, 20, torch.bfloat16),
    (192, 128, 5, False, 1, 2, 6, 10, torch.bfloat16),
    (128, 128, 5, True, 2, 1, 64, 64, torch.bfloat16),   # several tiles per image and per CTA
]


@pytest.mark.parametrize("cin,cout,k,corr,sd,su,H,W,in_dtype", CASES)
def test_conv_gradients_match_autograd(tfc, cin, cout, k, corr, sd, su, H, W, in_dtype):
  from compression_b200.python.ops import train_ops
  rng = np.random.default_rng(21)
  x = rng.integers(-3, 4, size=(2, H, W, cin)).astype(np.float32)
  kern = (rng.integers(-4, 5, size=(k, k, cin, cout)) / 4.0).astype(np.float32)
  kern *= (rng.random(size=kern.shape) < 0.25)
  b = (rng.integers(-4, 5, size=cout) / 2.0).astype(np.float32)
  layer = tfc.SignalConv2D(cout, (k, k), corr=corr, strides_down=sd, strides_up=su,
                           padding="same_zeros", use_bias=True, kernel_parameterizer=None,
                           kernel_initializer=_static(kern), bias_initializer=_static(b))
  layer.build(cin, "cuda")
  layer.kernel_var.requires_grad_()
  layer.bias_var.requires_grad_()
  xg = torch.from_numpy(x).cuda().to(in_dtype).requires_grad_()
  out = train_ops.signal_conv2d(layer, xg)
  dout = rng.integers(-2, 3, size=tuple(out.shape)).astype(np.float32)
  out.backward(torch.from_numpy(dout).cuda())

  xr = torch.from_numpy(x).double().requires_grad_()
  kr = torch.from_numpy(kern).double().requires_grad_()
  br = torch.from_numpy(b).double().requires_grad_()
  ref = _ref_conv(xr, kr, br, corr, sd, su)
  assert tuple(ref.shape) == tuple(out.shape)
  ref.backward(torch.from_numpy(dout).double())

  def close(got, want, what, tol=1e-5):
    got, want = got.detach().double().cpu(), want.detach()
    scale = want.abs().max().item() + 1e-9
    err = (got - want).abs().max().item() / scale
    assert err <= tol, (what, err, scale)

  close(out, ref, "forward")
  close(layer.kernel_var.grad, kr.grad, "filter gradient")
  close(layer.bias_var.grad, br.grad, "bias gradient")
  # the gradient of a bfloat16 activation is itself stored in bfloat16 (8 significand bits)
  close(xg.grad.float(), xr.grad, "input gradient", 1e-5 if in_dtype == torch.float32 else 4e-3)


def test_adam_kernel_follows_tf_adam(tfc):
  from compression_b200.python.ops import train_ops
  rng = np.random.default_rng(5)
  p0 = rng.normal(size=1000).astype(np.float32)
  bucket = train_ops.FlatBucket([torch.from_numpy(p0)], torch.device("cuda"))
  p, m, v = p0.astype(np.float64), np.zeros(1000), np.zeros(1000)
  lr, b1, b2, eps = 1e-3, 0.9, 0.999, 1e-8
  for t in range(1, 6):
    g = rng.normal(size=1000).astype(np.float32)
    bucket.grad.copy_(torch.from_numpy(g))
    train_ops.adam_step(bucket, lr, grad_scale=0.5)
    gs = g.astype(np.float64) * 0.5
    m = b1 * m + (1 - b1) * gs
    v = b2 * v + (1 - b2) * gs * gs
    p -= lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t) * m / (np.sqrt(v) + eps)
    np.testing.assert_allclose(bucket.param.cpu().numpy(), p, rtol=2e-5, atol=2e-6)


def _ref_losses(trainer, x, noise):
  """The same graph in float32 torch ops only (no CUDA kernels of ours)."""
  from compression_b200.python.layers import gdn
  model = trainer.model

  def run(seq, t):
    for layer in seq.layers:
      flip, _, _ = layer._plan()
      t = _ref_conv(t, layer.kernel, layer.bias, layer.corr, layer.strides_down[0],
                    layer.strides_up[0])
      if isinstance(layer.activation, gdn.GDN):
        act = layer.activation
        beta = act.beta_parameterizer(act.reparam_beta)
        gamma = act.gamma_parameterizer(act.reparam_gamma).float()
        norm = ((t * t).reshape(-1, t.shape[-1]) @ gamma + beta).reshape(t.shape)   # gdn.py:183-202
        t = t * (torch.sqrt(norm) if act.inverse else torch.rsqrt(norm))
    return t

  y = run(model.analysis_transform, x)
  y_tilde = y + noise
  lik = trainer._likelihood(y_tilde)
  x_tilde = run(model.synthesis_transform, y_tilde)
  num_pixels = float(np.prod(x.shape[:-1]))
  bpp = torch.log(lik).sum() / (-np.log(2) * num_pixels)
  mse = torch.mean((x - x_tilde) ** 2) * 255 ** 2
  return trainer.lmbda * mse + bpp


def test_training_step_gradients_and_descent(tfc):
  from compression_b200.examples import bls2017
  model = bls2017.Bls2017Model(128, "cuda", seed=3)
  trainer = bls2017.Bls2017Trainer(model, lmbda=0.01)
  assert trainer.num_parameters > 1_700_000        # SURVEY: ~1.77 M trainables
  g = torch.Generator().manual_seed(4)
  x = torch.rand((4, 64, 64, 3), generator=g).cuda()
  noise = (torch.rand((4, 4, 4, 128), generator=g) - .5).cuda()
  # gradients of the kernel-backed graph vs the torch-only graph (bf16 operands vs float32)
  trainer.main.zero_grad()
  out = trainer.losses(x, noise)
  out["loss"].backward()
  got = trainer.main.grad.clone()
  trainer.main.zero_grad()
  ref = _ref_losses(trainer, x, noise)
  ref.backward()
  want = trainer.main.grad.clone()
  assert abs(float(out["loss"]) - float(ref)) <= 1e-3 * abs(float(ref))
  cos = torch.dot(got, want) / (got.norm() * want.norm())
  assert cos > 0.9995, float(cos)                  # measured 1.0000 (relative L2 error 5e-4)
  assert (got - want).norm() <= 0.01 * want.norm()
  # a few steps on a fixed batch: the R-D loss goes down, the aux loss too, tables stay valid
  first = trainer.step(x, noise)
  for _ in range(30):
    last = trainer.step(x, noise)
  assert np.isfinite(last["loss"]) and last["loss"] < first["loss"]
  assert last["aux"] < first["aux"]
  c = model.compress(x)
  x_hat = model.decompress(c)
  assert tuple(x_hat.shape) == (4, 64, 64, 3) and bool(torch.isfinite(x_hat).all())


def test_eb_likelihood_kernel_gradients_match_autograd(tfc):
  """EntropyBottleneck.likelihood_training (one CUDA kernel each way) against the same function
  in torch ops under autograd (entropy_models.py:626-658 + the lower_bound gradient rule)."""
  from compression_b200.python.layers import entropy_models
  from compression_b200.python.ops import math_ops
  g = torch.Generator().manual_seed(9)
  eb = entropy_models.EntropyBottleneck()
  eb.build(12, "cuda", g)
  for group in (eb.matrices, eb.biases, eb.factors):
    for i in range(len(group)):
      group[i] = (group[i] + 0.3 * torch.randn(group[i].shape, generator=g).cuda()).requires_grad_()
  x = (torch.randn((3, 5, 7, 12), generator=g) * 6).cuda().requires_grad_()
  x.data[0, 0, 0, :] = 40.0                           # likelihood below the bound: gradient rule
  up = torch.randn((3, 5, 7, 12), generator=g).cuda()

  def torch_path(v4):
    c = v4.shape[-1]
    v = v4.reshape(-1, c).t().reshape(c, 1, -1)
    lower = eb._logits_cumulative(v - .5)
    upper = eb._logits_cumulative(v + .5)
    sign = -torch.sign(lower + upper).detach()
    lik = torch.abs(torch.sigmoid(sign * upper) - torch.sigmoid(sign * lower))
    lik = math_ops.lower_bound(lik, eb.likelihood_bound)
    return lik.reshape(c, -1).t().reshape(v4.shape)

  results = []
  for fn in (torch_path, eb.likelihood_training):
    for t in [x] + eb.matrices + eb.biases + eb.factors:
      t.grad = None
    lik = fn(x)
    (lik * up).sum().backward()
    results.append([lik.detach().clone(), x.grad.clone()] +
                   [t.grad.clone() for t in eb.matrices + eb.biases + eb.factors])
  names = ["likelihood", "d/dx"] + ["d/dparam %d" % i for i in range(len(results[0]) - 2)]
  for name, want, got in zip(names, results[0], results[1]):
    scale = float(want.abs().max()) + 1e-12
    err = float((got - want).abs().max()) / scale
    assert err <= 2e-4, (name, err, scale)


def test_training_step_cuda_graph_matches_eager(tfc):
  """step(graph=True) replays forward + backward as one CUDA graph: same parameters after a few
  steps as the launch-by-launch step (up to the order of the filter gradient's atomic adds)."""
  from compression_b200.examples import bls2017
  g = torch.Generator().manual_seed(5)
  x = torch.rand((4, 64, 64, 3), generator=g).cuda()
  noise = (torch.rand((4, 4, 4, 128), generator=g) - .5).cuda()
  params, losses = [], []
  for use_graph in (False, True):
    trainer = bls2017.Bls2017Trainer(bls2017.Bls2017Model(128, "cuda", seed=3), lmbda=0.01)
    for k in range(4):
      out = trainer.step(x * (1 - 0.1 * k), noise, graph=use_graph)   # inputs change per step
    params.append(trainer.main.param.clone())
    losses.append(out["loss"])
  assert abs(losses[0] - losses[1]) <= 1e-3 * abs(losses[0])
  # Adam normalises the step, so a gradient near zero whose sign the summation order decides
  # moves its parameter by up to lr = 1e-4 per step either way: bounded difference for those
  # few, agreement on average
  diff = (params[0] - params[1]).abs()
  assert float(diff.max()) <= 8e-4, float(diff.max())
  assert float(diff.mean()) <= 2e-5, float(diff.mean())


def test_bls2017_compress_decompress_roundtrip(tfc, oracle):
  """BASELINE configs[0] wiring: strings decode (CPU reference coder) to the coded symbols."""
  from compression_b200.examples import bls2017
  model = bls2017.Bls2017Model(128, "cuda", seed=0)
  g = torch.Generator().manual_seed(8)
  x = torch.rand((2, 64, 96, 3), generator=g).cuda()
  c = model.compress(x)
  assert c.y_shape == (4, 6)
  host = model.to_host(c)
  eb = model.entropy_bottleneck
  y = model.analysis_transform(x)
  sym = eb._quantize_symbols(y).cpu().numpy()
  idx = np.broadcast_to(np.arange(128, dtype=np.int32), sym.shape[1:])
  cdf, ln, off = (eb._quantized_cdf.cpu().numpy(), eb._cdf_length.cpu().numpy(),
                  eb._offset.cpu().numpy())
  for b in range(2):
    assert host.string[b] == oracle.ui_encode(sym[b], idx, cdf, ln, off, 16, 4)
  x_hat = model.decompress(model.to_device(host))
  assert tuple(x_hat.shape) == (2, 64, 96, 3)
  y_hat = eb.decompress_device(c.string, (4, 6, 128), 128)
  assert torch.equal(y_hat, eb._dequantize(eb._quantize_symbols(y)))


def _ref_transform(trainer, seq, t):
  """A transform in float32 torch ops only (no CUDA kernels of ours)."""
  from compression_b200.python.layers import gdn
  for layer in seq.layers:
    t = _ref_conv(t, layer.kernel, layer.bias, layer.corr, layer.strides_down[0],
                  layer.strides_up[0])
    if isinstance(layer.activation, gdn.GDN):
      t = trainer._gdn(layer.activation, t)
    elif layer.activation is not None:
      t = layer.activation(t)
  return t


def _torch_gaussian_likelihood(y, sigma, scale_bound, likelihood_bound):
  """entropy_models.py:822-825, :916-932, :965-969, :181-184 in torch ops (autograd)."""
  from compression_b200.python.ops import math_ops
  scale = math_ops.lower_bound(sigma, float(np.float32(scale_bound)))
  values = torch.abs(y)
  const = -(2 ** -0.5)
  upper = 0.5 * torch.erfc(const * ((0.5 - values) / scale))
  lower = 0.5 * torch.erfc(const * ((-0.5 - values) / scale))
  return math_ops.lower_bound(upper - lower, likelihood_bound)


def test_gaussian_likelihood_kernel_gradients_match_autograd(tfc):
  """`_GaussianLikelihood` (forward tfc_gaussian_call mode 1, backward
  tfc_gaussian_likelihood_backward) against the same function in float64 torch ops under
  autograd, including both lower_bound gradient rules (scale below the bound, likelihood below
  the bound, gradients of either sign)."""
  from compression_b200.python.layers import entropy_models
  g = torch.Generator().manual_seed(12)
  shape = (3, 5, 7, 16)
  y = (torch.randn(shape, generator=g) * 3).cuda()
  sigma = (torch.rand(shape, generator=g) * 3 - 0.3).cuda()      # some below the bound 0.11
  y[0, 0, 0, :] = 60.0                                           # likelihood below its bound
  y[0, 0, 1, :8] = 0.0                                           # |.| at zero
  up = torch.randn(shape, generator=g).cuda()
  bound, lb = float(np.float32(0.11)), 1e-9
  yk, sk = y.clone().requires_grad_(), sigma.clone().requires_grad_()
  lik = entropy_models._GaussianLikelihood.apply(yk, sk, bound, lb)
  (lik * up).sum().backward()
  yr, sr = y.double().requires_grad_(), sigma.double().requires_grad_()
  ref = _torch_gaussian_likelihood(yr, sr, bound, lb)
  (ref * up.double()).sum().backward()
  for name, got, want in (("likelihood", lik.detach(), ref.detach()), ("d/dy", yk.grad, yr.grad),
                          ("d/dsigma", sk.grad, sr.grad)):
    scale = float(want.abs().max()) + 1e-12
    err = float((got.double() - want).abs().max()) / scale
    assert err <= 2e-5, (name, err, scale)
  # elements clamped by a bound pass only negative gradients (math_ops.py:61-74)
  low = sigma < bound
  assert bool(((sk.grad[low] == 0) | (sk.grad[low] < 0)).all())
  assert bool((sk.grad[low] != 0).any()) and bool((sk.grad[low] == 0).any())


def test_bmshj2018_training_step_gradients_and_descent(tfc):
  """SURVEY §8f rank 3 (bmshj2018.py:201-282): hyperprior R-D loss through the CUDA conv
  kernels against the same graph in float32 torch ops, then a few descending steps."""
  from compression_b200.examples import bmshj2018
  model = bmshj2018.Bmshj2018Model(64, "cuda", seed=5)
  trainer = bmshj2018.Bmshj2018Trainer(model, lmbda=0.01)
  g = torch.Generator().manual_seed(6)
  x = torch.rand((2, 64, 64, 3), generator=g).cuda()
  noise = ((torch.rand((2, 4, 4, 64), generator=g) - .5).cuda(),
           (torch.rand((2, 1, 1, 64), generator=g) - .5).cuda())
  trainer.main.zero_grad()
  out = trainer.losses(x, noise)
  out["loss"].backward()
  got = trainer.main.grad.clone()

  trainer.main.zero_grad()
  y = _ref_transform(trainer, model.analysis_transform, x)
  z = _ref_transform(trainer, model.hyper_analysis_transform, torch.abs(y))
  z_tilde = z + noise[1]
  sigma = _ref_transform(trainer, model.hyper_synthesis_transform, z_tilde)[:, :4, :4, :]
  y_tilde = y + noise[0]
  x_tilde = _ref_transform(trainer, model.synthesis_transform, y_tilde)
  bpp = (torch.log(_torch_gaussian_likelihood(y_tilde, sigma, 0.11, 1e-9)).sum() +
         torch.log(trainer._likelihood(z_tilde)).sum()) / (-np.log(2) * 2 * 64 * 64)
  ref = 0.01 * torch.mean((x - x_tilde) ** 2) * 255 ** 2 + bpp
  ref.backward()
  want = trainer.main.grad.clone()
  assert abs(float(out["loss"]) - float(ref)) <= 2e-2 * abs(float(ref))
  cos = torch.dot(got, want) / (got.norm() * want.norm())
  assert cos > 0.99, float(cos)
  assert (got - want).norm() <= 0.15 * want.norm()

  first = trainer.step(x, noise)
  for _ in range(20):
    last = trainer.step(x, noise)
  assert np.isfinite(last["loss"]) and last["loss"] < first["loss"]
  c = model.compress(x)
  x_hat = model.decompress(c)
  assert tuple(x_hat.shape) == (2, 64, 64, 3) and bool(torch.isfinite(x_hat).all())


@pytest.mark.gpu
@pytest.mark.parametrize("pixels,channels", [(1, 4), (37, 128), (64 * 128 * 128 + 5, 128),
                                             (64 * 64 * 64, 192), (100003, 3), (4099, 5),
                                             (9 * 296 * 2 + 1, 1024)])
@pytest.mark.parametrize("misalign", [0, 1])
def test_column_sums_match_torch(pixels, channels, misalign):
  """tfc_column_sums (the bias gradients of the training step): float4 kernel with four passes
  in flight and the 16-byte vector atomic, its scalar-atomic variant when the destination is
  only 4-byte aligned (a slot inside a flat gradient bucket), and the scalar kernel for odd
  channel counts, against a float64 sum."""
  import ctypes
  from compression_b200 import _lib
  dev = torch.device("cuda", 0)
  g = torch.Generator(device=dev).manual_seed(pixels % 1000 + channels)
  x = torch.randn((pixels, channels), generator=g, device=dev, dtype=torch.float32)
  buf = torch.full((channels + 4,), 7.0, device=dev, dtype=torch.float32)
  out = buf[misalign:misalign + channels]
  _lib.check(_lib.lib().tfc_column_sums(
      _lib.ptr(x), ctypes.c_int64(pixels), ctypes.c_int32(channels),
      ctypes.c_void_p(buf.data_ptr() + 4 * misalign), _lib.stream_ptr(dev)))
  want = x.double().sum(0)
  tol = 1e-6 * float(x.abs().double().sum(0).max()) + 1e-6
  assert float((out.double() - want).abs().max()) <= tol
  assert float(buf[misalign + channels]) == 7.0 and (misalign == 0 or float(buf[0]) == 7.0)
