"""GPU parity tests of the transform and entropy-model kernels (through the layer classes and
the C ABI) against oracle/tfc_np.py and the reference's own test properties."""
import numpy as np
import pytest
import torch

from oracle import tfc_np
from tests.golden import tfc_goldens as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tfc():
  import compression_b200 as tfc_mod
  return tfc_mod


def _static(kernel):
  """StaticParameterizer-style fixed kernel (signal_conv_test.py:96-97)."""
  return lambda shape: torch.as_tensor(kernel, dtype=torch.float32)


CONV_CASES = [
    # corr, kernel, sd, su, padding, extra_pad_end
    (True, (3, 3), 1, 1, "valid", True), (False, (3, 3), 1, 1, "valid", True),
    (True, (5, 5), 2, 1, "valid", True), (False, (5, 5), 1, 2, "valid", True),
    (True, (5, 5), 1, 2, "valid", True), (False, (5, 5), 1, 2, "valid", False),
    (False, (4, 4), 1, 2, "valid", True), (True, (2, 4), 2, 1, "valid", True),
    (True, (5, 5), 2, 1, "same_zeros", True), (False, (5, 5), 1, 2, "same_zeros", True),
    (True, (3, 3), 1, 1, "same_zeros", True), (False, (3, 3), 1, 1, "same_zeros", True),
    (True, (9, 9), 4, 1, "same_zeros", True), (False, (9, 9), 1, 4, "same_zeros", True),
    (True, (5, 5), 1, 2, "same_zeros", True), (False, (5, 5), 2, 2, "same_zeros", True),
    (False, (4, 4), 1, 2, "same_zeros", True), (False, (5, 5), 1, 2, "same_zeros", False),
    # same_reflect (signal_conv.py:744-758; signal_conv_test.py sweeps it like same_zeros)
    (True, (3, 3), 1, 1, "same_reflect", True), (False, (3, 3), 1, 1, "same_reflect", True),
    (True, (5, 5), 2, 1, "same_reflect", True), (False, (5, 5), 1, 2, "same_reflect", True),
    (True, (5, 5), 1, 2, "same_reflect", True), (False, (4, 4), 1, 2, "same_reflect", True),
    (True, (2, 4), 2, 1, "same_reflect", True),
]


@pytest.mark.parametrize("impl", [1])
@pytest.mark.parametrize("corr,ks,sd,su,padding,epe", CONV_CASES)
def test_signal_conv_matches_oracle(tfc, corr, ks, sd, su, padding, epe, impl):
  """Integer-valued data as in signal_conv_test.py:81-129 (exact in float32), every dispatch
  branch of signal_conv.py:709-782 that the 2-D same_zeros/valid layer can take."""
  rng = np.random.default_rng(0)
  x = rng.integers(0, 32, size=(2, 13, 11, 3)).astype(np.float32)
  k = rng.integers(0, 16, size=ks + (3, 5)).astype(np.float32)
  b = rng.integers(-8, 8, size=5).astype(np.float32)
  layer = tfc.SignalConv2D(5, ks, corr=corr, strides_down=sd, strides_up=su, padding=padding,
                           extra_pad_end=epe, use_bias=True, kernel_parameterizer=None,
                           kernel_initializer=_static(k), bias_initializer=_static(b),
                           impl=impl)
  got = layer(torch.from_numpy(x).cuda())
  want = tfc_np.signal_conv2d(x, k, b, corr, sd, su, padding, epe).numpy()
  assert tuple(got.shape) == want.shape == tuple(layer.compute_output_shape(x.shape))
  np.testing.assert_allclose(got.cpu().numpy(), want, rtol=0, atol=1e-3)


def test_signal_conv_not_implemented(tfc):
  """signal_conv.py:393-401 / signal_conv_test.py:185-219."""
  x = torch.zeros((1, 8, 8, 2), device="cuda")
  with pytest.raises(NotImplementedError):
    tfc.SignalConv2D(2, (4, 4), corr=True, strides_up=2, padding="valid")(x)
  with pytest.raises(NotImplementedError):        # upsampling + downsampling over a reflection
    tfc.SignalConv2D(2, (5, 5), corr=False, strides_up=2, strides_down=2,
                     padding="same_reflect")(x)
  with pytest.raises(NotImplementedError):        # upsampled separable layers: one filter only (:590-596)
    tfc.SignalConv2D(2, (3, 3), corr=False, strides_up=2, channel_separable=True)(x)


def test_signal_conv_channels_first(tfc):
  """data_format="channels_first" (signal_conv.py:216-218, :790-800): same numbers as the
  channels_last layer on the transposed tensors, bias and activation included."""
  rng = np.random.default_rng(3)
  x = rng.integers(0, 16, size=(2, 12, 10, 3)).astype(np.float32)
  k = rng.integers(-4, 5, size=(5, 5, 3, 6)).astype(np.float32)
  b = rng.integers(-8, 8, size=6).astype(np.float32)
  outs = {}
  for fmt in ("channels_last", "channels_first"):
    layer = tfc.SignalConv2D(6, (5, 5), corr=True, strides_down=2, padding="same_zeros",
                             data_format=fmt, use_bias=True, activation=torch.relu,
                             kernel_parameterizer=None, kernel_initializer=_static(k),
                             bias_initializer=_static(b))
    xin = torch.from_numpy(x).cuda()
    outs[fmt] = layer(xin if fmt == "channels_last" else xin.permute(0, 3, 1, 2).contiguous())
  assert tuple(outs["channels_first"].shape) == (2, 6, 6, 5)
  assert torch.equal(outs["channels_first"].permute(0, 2, 3, 1), outs["channels_last"])


def test_rdft_parameterizer_roundtrip(tfc):
  """parameterizers_test.py:46-52: the RDFT-parameterised kernel equals its initial value."""
  k = np.random.default_rng(1).normal(size=(5, 5, 3, 4)).astype(np.float32)
  layer = tfc.SignalConv2D(4, (5, 5), corr=True, padding="same_zeros",
                           kernel_initializer=_static(k))
  layer.build(3, "cuda")
  assert tuple(layer.kernel_var.shape) == (25, 12)
  np.testing.assert_allclose(layer.kernel.cpu().numpy(), k, atol=1e-5)


@pytest.mark.parametrize("inverse", [False, True])
def test_gdn_default_init_closed_form(tfc, inverse):
  """gdn_test.py:49-75."""
  x = np.random.default_rng(2).uniform(size=(1, 2, 3, 4)).astype(np.float32)
  y = tfc.GDN(inverse=inverse)(torch.from_numpy(x).cuda()).cpu().numpy()
  want = x * np.sqrt(1 + .1 * x ** 2) if inverse else x / np.sqrt(1 + .1 * x ** 2)
  np.testing.assert_allclose(y, want, rtol=0, atol=1e-6)


@pytest.mark.parametrize("inverse", [False, True])
def test_conv_gdn_fused_matches_oracle(tfc, inverse):
  """Analysis/synthesis layer = conv + bias + (I)GDN with a non-trivial gamma."""
  rng = np.random.default_rng(3)
  c = 8
  x = rng.normal(size=(2, 12, 10, c)).astype(np.float32)
  k = (rng.normal(size=(5, 5, c, c)) / 10).astype(np.float32)
  b = rng.normal(size=c).astype(np.float32)
  act = tfc.GDN(inverse=inverse)
  if inverse:
    layer = tfc.SignalConv2D(c, (5, 5), corr=False, strides_up=2, padding="same_zeros",
                             use_bias=True, activation=act, kernel_initializer=_static(k),
                             bias_initializer=_static(b), impl=1)
  else:
    layer = tfc.SignalConv2D(c, (5, 5), corr=True, strides_down=2, padding="same_zeros",
                             use_bias=True, activation=act, kernel_initializer=_static(k),
                             bias_initializer=_static(b), impl=1)
  layer.build(c, "cuda")
  act.reparam_gamma = act.gamma_parameterizer.init(
      torch.as_tensor(rng.uniform(0, .2, size=(c, c)), dtype=torch.float32).cuda())
  act.reparam_beta = act.beta_parameterizer.init(
      torch.as_tensor(rng.uniform(.5, 2, size=c), dtype=torch.float32).cuda())
  act.invalidate()
  got = layer(torch.from_numpy(x).cuda()).cpu().numpy()
  beta = tfc_np.nonnegative_reparam(act.reparam_beta.cpu().numpy(), 1e-6)
  gamma = tfc_np.nonnegative_reparam(act.reparam_gamma.cpu().numpy(), 0)
  pre = tfc_np.signal_conv2d(x, layer.kernel.cpu().numpy(), b, not inverse,
                             1 if inverse else 2, 2 if inverse else 1, "same_zeros")
  want = tfc_np.gdn(pre, beta, gamma, inverse).numpy()
  np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-4)


# ---- entropy models ------------------------------------------------------------------------
def _eb_oracle_params(eb):
  m, b, f = tfc_np.eb_params64([t.cpu().numpy() for t in eb.matrices],
                               [t.cpu().numpy() for t in eb.biases],
                               [t.cpu().numpy() for t in eb.factors])
  return m, b, f


@pytest.mark.parametrize("filters", [(3, 3, 3), (), (3,), (2, 2), (4, 4, 4, 4)])
def test_entropy_bottleneck_call(tfc, filters):
  """values exact; likelihood max-abs <= 1e-3 (north-star tolerance; measured ~1e-6) against
  the float64 restatement of entropy_models.py:626-658."""
  rng = np.random.default_rng(4)
  c = 6
  eb = tfc.EntropyBottleneck(filters=filters)
  eb.build(c, "cuda", torch.Generator().manual_seed(1))
  for t in eb.factors:                       # non-trivial tanh factors
    t.copy_(torch.as_tensor(rng.normal(size=tuple(t.shape)) * .5, dtype=torch.float32))
  eb.quantiles[:, 0, 1] = torch.as_tensor(rng.uniform(-.4, .4, size=c), dtype=torch.float32)
  eb.refresh()
  x = rng.uniform(-25, 25, size=(3, 7, 5, c)).astype(np.float32)
  values, lik = eb(torch.from_numpy(x).cuda(), training=False)
  m, b, f = _eb_oracle_params(eb)
  med = eb._medians.cpu().numpy()
  want_v, want_l = tfc_np.eb_call(x, med, m, b, f)
  np.testing.assert_array_equal(values.cpu().numpy(), want_v)
  assert np.abs(lik.cpu().numpy() - want_l).max() <= 1e-3
  np.testing.assert_allclose(lik.cpu().numpy(), want_l, rtol=2e-3, atol=1e-7)
  sym = eb._quantize_symbols(torch.from_numpy(x).cuda())
  np.testing.assert_array_equal(sym.cpu().numpy(), tfc_np.eb_symbols(x, med))
  deq = eb._dequantize(sym)
  np.testing.assert_array_equal(deq.cpu().numpy(), want_v)


def test_entropy_bottleneck_tables_and_codec(tfc, oracle):
  """update() builds the same tables as the restatement of entropy_models.py:521-596 (up to
  float32-vs-float64 density evaluation: |cdf diff| <= 2), and compress/decompress round-trips
  bit-exactly against the CPU coder on the layer's own tables."""
  rng = np.random.default_rng(5)
  c = 8
  eb = tfc.EntropyBottleneck()
  eb.build(c, "cuda", torch.Generator().manual_seed(2))
  assert eb._quantized_cdf.cpu().numpy().tolist() == [[0, 32768, 65536]] * c  # :563-583
  eb.update()
  m, b, f = _eb_oracle_params(eb)
  cdf, length, offset, med = tfc_np.eb_tables(
      eb.quantiles.cpu().numpy(), m, b, f, lambda p: oracle.pmf_to_cdf(p[None], 16)[0])
  np.testing.assert_array_equal(eb._cdf_length.cpu().numpy(), length)
  np.testing.assert_array_equal(eb._offset.cpu().numpy(), offset)
  assert length.tolist() == [23] * c               # init_scale=10
  got_cdf = eb._quantized_cdf.cpu().numpy()
  assert np.abs(got_cdf - cdf).max() <= 2
  assert (got_cdf[:, 0] == 0).all() and (got_cdf[:, 22] == 65536).all()
  x = rng.normal(scale=6, size=(5, 9, 4, c)).astype(np.float32)
  x[0, 0, 0, :] = [40, -40, 200, -300, 11, -11, 10, -10]      # escapes
  xd = torch.from_numpy(x).cuda()
  strings = eb.compress(xd)
  sym = tfc_np.eb_symbols(x, med)
  idx = np.broadcast_to(np.arange(c, dtype=np.int32), x.shape[1:])
  for s in range(5):
    assert strings[s] == oracle.ui_encode(sym[s], idx, got_cdf, length, offset, 16, 4)
  dec = eb.decompress(strings, x.shape[1:], channels=c)
  np.testing.assert_array_equal(dec.cpu().numpy(), sym.astype(np.float32) + med)
  assert np.abs(dec.cpu().numpy() - x).max() <= .5 + 1e-5


@pytest.mark.parametrize("kind", ["gaussian", "logistic", "laplacian"])
def test_conditional_tables_golden_and_decode(tfc, kind):
  """G2-G5: tables built on the GPU equal the golden CDF arrays, and the layer decodes the
  reference's stored bit strings (entropy_models_test.py:407-425)."""
  cls = {"gaussian": tfc.GaussianConditional, "logistic": tfc.LogisticConditional,
         "laplacian": tfc.LaplacianConditional}[kind]
  g = G.CONDITIONALS[kind]
  layer = cls(None, G.COND_SCALE_TABLE, indexes=torch.from_numpy(G.COND_INDEXES).cuda())
  layer.build()
  np.testing.assert_array_equal(layer._quantized_cdf.cpu().numpy(), g["cdf"])
  np.testing.assert_array_equal(layer._cdf_length.cpu().numpy(), g["cdf_length"])
  np.testing.assert_array_equal(layer._offset.cpu().numpy(), g["offset"])
  dec = layer.decompress(g["bitstrings"])
  np.testing.assert_array_equal(dec.cpu().numpy(), G.COND_EXPECTED)
  enc = layer.compress(torch.from_numpy(G.COND_EXPECTED).cuda())
  assert enc == g["bitstrings"]


def test_gaussian_conditional_call_indexes_codec(tfc, oracle):
  """Likelihood <= 1e-3 max-abs vs float64 (entropy_models.py:916-932, :965-969); indexes and
  symbols exact (:869-881, :898); strings bit-exact on the 64-level bmshj2018 table."""
  rng = np.random.default_rng(6)
  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))
  shape = (3, 6, 5, 16)
  sigma = np.exp(rng.uniform(np.log(0.05), np.log(300), size=shape)).astype(np.float32)
  sigma.ravel()[:70] = np.concatenate([table.astype(np.float32), [-1, 0, 0.11, 1e-9, 256, 1e4]])
  y = (rng.standard_normal(shape) * sigma * 1.2).astype(np.float32)
  y.ravel()[100:104] = [0.5, -0.5, 1.5, -1.5]            # round-half-up, not rint
  gc = tfc.GaussianConditional(torch.from_numpy(sigma).cuda(), table)
  values, lik = gc(torch.from_numpy(y).cuda(), training=False)
  want_v, want_l = tfc_np.conditional_call(y, sigma, table)
  np.testing.assert_array_equal(values.cpu().numpy(), want_v)
  assert np.abs(lik.cpu().numpy() - want_l).max() <= 1e-3
  np.testing.assert_allclose(lik.cpu().numpy(), want_l, rtol=1e-3, atol=1e-7)
  want_idx = tfc_np.conditional_indexes(sigma, table)
  np.testing.assert_array_equal(gc.indexes.cpu().numpy(), want_idx)
  cdf, length, offset = tfc_np.conditional_tables(
      table, lambda p: oracle.pmf_to_cdf(p[None], 16)[0])
  np.testing.assert_array_equal(gc._cdf_length.cpu().numpy(), length)
  np.testing.assert_array_equal(gc._offset.cpu().numpy(), offset)
  np.testing.assert_array_equal(gc._quantized_cdf.cpu().numpy(), cdf)
  assert int(length.sum()) == 13014 and cdf.shape == (64, 1481)
  gc2 = tfc.GaussianConditional(torch.from_numpy(sigma).cuda(), table)
  strings = gc2.compress(torch.from_numpy(y).cuda())
  sym = tfc_np.conditional_symbols(y)
  for s in range(shape[0]):
    assert strings[s] == oracle.ui_encode(sym[s], want_idx[s], cdf, length, offset, 16, 4)
  dec = gc2.decompress(strings)
  np.testing.assert_array_equal(dec.cpu().numpy(), sym.astype(np.float32))


@pytest.mark.parametrize("kind", ["gaussian", "logistic", "laplacian"])
def test_conditional_quantization_and_codec_with_mean(tfc, kind):
  """entropy_models_test.py:300-320 (test_quantization_mean) and :339-361 (test_codec_mean):
  values are quantised to integers offset by the mean, through call and through the codec;
  the likelihood is that of the centred value."""
  cls = {"gaussian": tfc.GaussianConditional, "logistic": tfc.LogisticConditional,
         "laplacian": tfc.LaplacianConditional}[kind]
  rng = np.random.default_rng(15)
  values = np.linspace(-50, 50, 100).astype(np.float32)
  mean = rng.standard_normal(values.shape).astype(np.float32)
  scale = rng.uniform(1, 10, size=values.shape).astype(np.float32)
  layer = cls(torch.from_numpy(scale).cuda(), [1], mean=torch.from_numpy(mean).cuda())
  quantized, lik = layer(torch.from_numpy(values).cuda(), training=False)
  want = np.floor(values - mean + np.float32(.5)) + mean
  np.testing.assert_allclose(quantized.cpu().numpy(), want, rtol=0, atol=1e-5)
  plain = cls(torch.from_numpy(scale).cuda(), [1])
  _, lik0 = plain(torch.from_numpy(values - mean).cuda(), training=False)
  np.testing.assert_allclose(lik.cpu().numpy(), lik0.cpu().numpy(), rtol=1e-6, atol=1e-9)

  values2 = values[None]
  mean2 = rng.standard_normal(values2.shape).astype(np.float32)
  scale2 = rng.uniform(25, 75, size=values2.shape).astype(np.float32)
  codec = cls(torch.from_numpy(scale2).cuda(), [2.0 ** x for x in range(-10, 10)],
              mean=torch.from_numpy(mean2).cuda())
  strings = codec.compress(torch.from_numpy(values2).cuda())
  decoded = codec.decompress(strings)
  want2 = np.floor(values2 - mean2 + np.float32(.5)) + mean2
  np.testing.assert_allclose(decoded.cpu().numpy(), want2, rtol=0, atol=1e-5)


def test_lower_bound_gradients(tfc):
  """math_ops_test.py:28-102."""
  x = torch.tensor([-1., 1.], requires_grad=True)
  for mode, want in (("identity_if_towards", [0., 1.]), ("identity", [1., 1.]),
                     ("disconnected", [0., 1.])):
    x.grad = None
    tfc.lower_bound(x, 0., gradient=mode).sum().backward()
    assert x.grad.tolist() == want
  x.grad = None
  (-tfc.lower_bound(x, 0.)).sum().backward()    # gradient pushes the input up -> passes
  assert x.grad.tolist() == [-1., -1.]
  with pytest.raises(ValueError):
    tfc.lower_bound(x, 0., gradient="nope")


# ---- tcgen05 implicit-GEMM path (impl=2) ------------------------------------------------------
UMMA_CASES = [
    # cin, cout, kernel, corr, sd, su, H, W, act
    (64, 64, 3, True, 1, 1, 8, 16, None),
    (192, 192, 5, True, 2, 1, 32, 48, None),
    (192, 192, 5, True, 2, 1, 24, 40, "relu"),
    (128, 128, 5, True, 2, 1, 16, 16, None),
    (192, 192, 3, True, 1, 1, 12, 20, "relu"),
    (192, 192, 3, False, 1, 1, 12, 20, None),
    (192, 192, 5, False, 1, 2, 12, 20, "relu"),
    (192, 3, 5, False, 1, 2, 16, 24, None),
    (128, 3, 9, False, 1, 4, 8, 8, None),
    (192, 192, 5, True, 2, 1, 32, 48, "gdn"),
    (192, 192, 5, False, 1, 2, 12, 20, "igdn"),
    (128, 128, 5, True, 2, 1, 20, 36, "gdn"),
    (64, 64, 5, False, 1, 2, 6, 10, "igdn"),
    # sizes that are not multiples of the stride (1080p: 135 -> 68 rows): zero-extended copy
    (192, 192, 5, True, 2, 1, 27, 45, "gdn"),
    (192, 192, 5, True, 2, 1, 33, 48, None),
    (64, 128, 5, True, 2, 1, 16, 21, "relu"),
]


@pytest.mark.parametrize("cin,cout,k,corr,sd,su,H,W,act", UMMA_CASES)
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_umma_conv_matches_oracle(tfc, cin, cout, k, corr, sd, su, H, W, act, out_dtype):
  """bf16-exact integer-valued data: products and fp32 sums are exact, so the tensor-core
  path must agree with the float64 oracle to fp32 rounding (plus the bf16 rounding of x^2 in
  the GDN GEMM / of the output when it is bfloat16)."""
  rng = np.random.default_rng(10)
  n = 3
  x = rng.integers(-4, 5, size=(n, H, W, cin)).astype(np.float32)
  kern = (rng.integers(-8, 9, size=(k, k, cin, cout)) / 8.0).astype(np.float32)
  kern *= (rng.random(size=kern.shape) < 0.3)          # keep magnitudes moderate
  b = (rng.integers(-8, 9, size=cout) / 4.0).astype(np.float32)
  activation = None
  if act == "relu":
    activation = torch.relu
  elif act in ("gdn", "igdn"):
    activation = tfc.GDN(inverse=(act == "igdn"))
  layer = tfc.SignalConv2D(cout, (k, k), corr=corr, strides_down=sd, strides_up=su,
                           padding="same_zeros", use_bias=True, activation=activation,
                           kernel_parameterizer=None, kernel_initializer=_static(kern),
                           bias_initializer=_static(b), impl=2)
  layer.build(cin, "cuda")
  beta = gamma = None
  if act in ("gdn", "igdn"):
    g0 = (rng.integers(0, 5, size=(cout, cout)) / 64.0).astype(np.float32)
    b0 = (rng.integers(4, 12, size=cout) / 4.0).astype(np.float32)
    activation.reparam_gamma = activation.gamma_parameterizer.init(torch.from_numpy(g0).cuda())
    activation.reparam_beta = activation.beta_parameterizer.init(torch.from_numpy(b0).cuda())
    activation.invalidate()
    beta, gamma = [t.cpu().numpy() for t in activation.effective()]
  got = layer(torch.from_numpy(x).cuda().to(torch.bfloat16), out_dtype=out_dtype)
  assert got.dtype == out_dtype
  want = tfc_np.signal_conv2d(x, kern, b, corr, sd, su, "same_zeros")
  if act == "relu":
    want = torch.relu(want)
  elif act in ("gdn", "igdn"):
    want = tfc_np.gdn(want, beta, gamma, act == "igdn")
  want = want.numpy()
  assert tuple(got.shape) == want.shape
  got = got.float().cpu().numpy()
  tol = 2e-2 if (act in ("gdn", "igdn") or out_dtype == torch.bfloat16) else 1e-4
  scale = np.abs(want).max() + 1e-6
  err = np.abs(got - want).max() / scale
  assert err <= tol, (err, scale)


def test_umma_unsupported_shape_reports(tfc):
  x = torch.zeros((1, 8, 8, 100), device="cuda", dtype=torch.bfloat16)
  layer = tfc.SignalConv2D(8, (3, 3), corr=True, padding="same_zeros", impl=2)
  with pytest.raises(NotImplementedError, match="cin must be a multiple of 64"):
    layer(x)


UPTHIN_CASES = [
    # cin, cout, kernel, su, H, W, act, out dtype
    (192, 3, 5, 2, 16, 24, None, torch.float32),
    (192, 3, 5, 2, 13, 31, "relu", torch.float32),     # ragged: tiles overshoot the image
    (128, 3, 5, 2, 6, 14, None, torch.bfloat16),       # exactly one interior tile
    (64, 1, 3, 2, 9, 9, None, torch.float32),
    (192, 4, 5, 2, 20, 33, None, torch.float32),
    (256, 2, 7, 2, 11, 17, None, torch.float32),       # 7*7*2 = 98 accumulator columns
]


@pytest.mark.parametrize("cin,cout,k,su,H,W,act,out_dtype", UPTHIN_CASES)
def test_thin_output_transposed_conv_scatter_gemm(tfc, cin, cout, k, su, H, W, act, out_dtype):
  """Last synthesis layer (192 -> 3, 5x5 up 2; signal_conv.py:629-707) on the scatter-GEMM
  kernel (impl=3), against the float64 oracle and against the implicit-GEMM kernel."""
  rng = np.random.default_rng(12)
  x = rng.integers(-4, 5, size=(3, H, W, cin)).astype(np.float32)
  kern = (rng.integers(-8, 9, size=(k, k, cin, cout)) / 8.0).astype(np.float32)
  b = (rng.integers(-8, 9, size=cout) / 4.0).astype(np.float32)
  outs = {}
  for impl in (3, 2):
    layer = tfc.SignalConv2D(cout, (k, k), corr=False, strides_up=su, padding="same_zeros",
                             use_bias=True, activation=torch.relu if act == "relu" else None,
                             kernel_parameterizer=None, kernel_initializer=_static(kern),
                             bias_initializer=_static(b), impl=impl)
    layer.build(cin, "cuda")
    outs[impl] = layer(torch.from_numpy(x).cuda().to(torch.bfloat16), out_dtype=out_dtype)
    assert outs[impl].dtype == out_dtype
  want = tfc_np.signal_conv2d(x, kern, b, False, 1, su, "same_zeros")
  if act == "relu":
    want = torch.relu(want)
  want = want.numpy()
  got = outs[3].float().cpu().numpy()
  assert got.shape == want.shape
  scale = np.abs(want).max() + 1e-6
  tol = 1e-2 if out_dtype == torch.bfloat16 else 1e-5
  assert np.abs(got - want).max() / scale <= tol
  assert np.abs(got - outs[2].float().cpu().numpy()).max() / scale <= tol


S2D_CASES = [
    # cin, cout, kernel, sd, H, W, act, input dtype
    (3, 192, 5, 2, 32, 48, "gdn", torch.float32),
    (3, 192, 5, 2, 24, 40, None, torch.bfloat16),
    (3, 128, 9, 4, 32, 32, "gdn", torch.float32),
    (3, 64, 3, 1, 12, 20, "relu", torch.float32),
    (12, 64, 5, 2, 16, 24, None, torch.float32),
]


@pytest.mark.parametrize("cin,cout,k,sd,H,W,act,in_dtype", S2D_CASES)
def test_thin_input_conv_space_to_depth(tfc, cin, cout, k, sd, H, W, act, in_dtype):
  """First analysis layer (3 input channels) on the tensor cores via space-to-depth."""
  rng = np.random.default_rng(11)
  x = rng.integers(-4, 5, size=(2, H, W, cin)).astype(np.float32)
  kern = (rng.integers(-8, 9, size=(k, k, cin, cout)) / 8.0).astype(np.float32)
  b = (rng.integers(-8, 9, size=cout) / 4.0).astype(np.float32)
  activation = {"relu": torch.relu, "gdn": tfc.GDN(), None: None}[act]
  layer = tfc.SignalConv2D(cout, (k, k), corr=True, strides_down=sd, padding="same_zeros",
                           use_bias=True, activation=activation, kernel_parameterizer=None,
                           kernel_initializer=_static(kern), bias_initializer=_static(b), impl=2)
  layer.build(cin, "cuda")
  got = layer(torch.from_numpy(x).cuda().to(in_dtype), out_dtype=torch.float32)
  want = tfc_np.signal_conv2d(x, kern, b, True, sd, 1, "same_zeros")
  if act == "relu":
    want = torch.relu(want)
  elif act == "gdn":
    beta, gamma = [t.cpu().numpy() for t in activation.effective()]
    want = tfc_np.gdn(want, beta, gamma, False)
  want = want.numpy()
  assert tuple(got.shape) == want.shape
  err = np.abs(got.cpu().numpy() - want).max() / (np.abs(want).max() + 1e-6)
  assert err <= (2e-2 if act == "gdn" else 1e-4), err


THIN_CASES = [
    # cout, kernel, sd, H, W, act, input dtype
    (192, 5, 2, 32, 48, "gdn", torch.float32),
    (192, 5, 2, 33, 47, "gdn", torch.float32),        # odd sizes: ragged tiles and borders
    (192, 5, 2, 24, 40, None, torch.bfloat16),
    (128, 5, 2, 64, 64, "gdn", torch.bfloat16),
    (128, 9, 4, 32, 32, "gdn", torch.float32),        # first analysis layer of bls2017
    (128, 9, 4, 61, 35, "relu", torch.float32),
    (64, 9, 4, 16, 16, None, torch.float32),
    (100, 5, 1, 12, 20, "relu", torch.float32),       # cout not a multiple of 32, stride 1
    (192, 5, 2, 4, 4, "gdn", torch.float32),          # smaller than one tile
]


@pytest.mark.parametrize("cout,k,sd,H,W,act,in_dtype", THIN_CASES)
def test_thin_input_conv_im2col(tfc, cout, k, sd, H, W, act, in_dtype):
  """First analysis layer on the software-im2col tcgen05 kernel (impl=4; signal_conv.py:487-524
  + gdn.py:183-202) against the float64 oracle."""
  rng = np.random.default_rng(13)
  x = rng.integers(-4, 5, size=(3, H, W, 3)).astype(np.float32)
  kern = (rng.integers(-8, 9, size=(k, k, 3, cout)) / 8.0).astype(np.float32)
  b = (rng.integers(-8, 9, size=cout) / 4.0).astype(np.float32)
  activation = {"relu": torch.relu, "gdn": tfc.GDN(), None: None}[act]
  layer = tfc.SignalConv2D(cout, (k, k), corr=True, strides_down=sd, padding="same_zeros",
                           use_bias=True, activation=activation, kernel_parameterizer=None,
                           kernel_initializer=_static(kern), bias_initializer=_static(b), impl=4)
  layer.build(3, "cuda")
  xin = torch.from_numpy(x).cuda().to(in_dtype)
  got = layer(xin, out_dtype=torch.float32)
  again = layer(xin, out_dtype=torch.float32)        # second call: weights already staged
  assert torch.equal(got, again)
  want = tfc_np.signal_conv2d(x, kern, b, True, sd, 1, "same_zeros")
  if act == "relu":
    want = torch.relu(want)
  elif act == "gdn":
    beta, gamma = [t.cpu().numpy() for t in activation.effective()]
    want = tfc_np.gdn(want, beta, gamma, False)
  want = want.numpy()
  assert tuple(got.shape) == want.shape
  err = np.abs(got.cpu().numpy() - want).max() / (np.abs(want).max() + 1e-6)
  assert err <= (2e-2 if act == "gdn" else 1e-4), err


def test_thin_input_conv_unsupported_reports(tfc):
  layer = tfc.SignalConv2D(8, (3, 3), corr=True, padding="same_zeros", impl=4)
  with pytest.raises(NotImplementedError, match="3 channels and 5x5 / 9x9"):
    layer(torch.zeros((1, 8, 8, 3), device="cuda"))


def test_one_layer_alternating_dtypes_and_sizes(tfc):
  """the repacked weights in a layer's workspace belong to the kernel path that wrote
  them (tfc_signal_conv2d_path).  One layer, one stream: bf16 / fp32 inputs and even / odd sizes
  interleaved — every call must equal a fresh layer's result for the same input."""
  rng = np.random.default_rng(31)
  cin, cout, k = 64, 64, 5
  kern = (rng.integers(-8, 9, size=(k, k, cin, cout)) / 8.0).astype(np.float32)
  b = (rng.integers(-8, 9, size=cout) / 4.0).astype(np.float32)

  def make():
    layer = tfc.SignalConv2D(cout, (k, k), corr=True, strides_down=2, padding="same_zeros",
                             use_bias=True, activation=tfc.GDN(), kernel_parameterizer=None,
                             kernel_initializer=_static(kern), bias_initializer=_static(b))
    layer.build(cin, "cuda")
    return layer

  shared = make()
  cases = [(torch.bfloat16, 16, 24), (torch.float32, 16, 24), (torch.bfloat16, 16, 24),
           (torch.bfloat16, 15, 23), (torch.float32, 15, 23), (torch.bfloat16, 16, 24),
           (torch.bfloat16, 15, 23), (torch.bfloat16, 16, 24)]
  paths = set()
  for dt, h, w in cases:
    x = torch.from_numpy(rng.integers(-3, 4, size=(2, h, w, cin)).astype(np.float32)).cuda().to(dt)
    got = shared(x, out_dtype=torch.float32)
    want = make()(x, out_dtype=torch.float32)
    assert torch.equal(got, want), (dt, h, w)
    paths.add(next(iter(shared._ws.values()))[3])
  assert len(paths) >= 2, paths           # the sequence really crossed kernel paths


def test_uint8_images_first_and_last_layer(tfc):
  """TFC_U8: the first analysis layer reads the image bytes (b / 255, read_png) and the last
  synthesis layer writes quantize_image(x) = saturate_cast<uint8>(round(x * 255)); both must
  equal the float path followed / preceded by the torch ops, also where the thin kernels do not
  apply (fallback)."""
  from compression_b200.python.layers import signal_conv
  g = torch.Generator().manual_seed(5)
  for h, w in ((64, 96), (33, 47)):         # TMA patches | rows that are not 16-byte multiples
    u8 = torch.randint(0, 256, (2, h, w, 3), generator=g, dtype=torch.uint8).cuda()
    a0 = tfc.SignalConv2D(192, (5, 5), corr=True, strides_down=2, padding="same_zeros",
                          use_bias=True, activation=tfc.GDN())
    a0.build(3, "cuda", g)
    got = a0(u8, out_dtype=torch.bfloat16)
    want = a0(u8.float() / 255, out_dtype=torch.bfloat16)
    assert torch.equal(got, want), (h, w)
  # a layer the thin-input kernel does not take (3x3): scaling by torch, same result
  c3 = tfc.SignalConv2D(8, (3, 3), corr=True, padding="same_zeros")
  c3.build(3, "cuda", g)
  assert torch.equal(c3(u8), c3(u8.float() / 255))
  for h, w in ((16, 24), (13, 31)):
    y = (torch.randn((2, h, w, 192), generator=g) * 0.5).cuda().to(torch.bfloat16)
    s3 = tfc.SignalConv2D(3, (5, 5), corr=False, strides_up=2, padding="same_zeros",
                          use_bias=True, bias_initializer=_static(np.array([.3, .5, .7], np.float32)))
    s3.build(192, "cuda", g)
    q = s3(y, out_dtype=torch.uint8)
    f = s3(y, out_dtype=torch.float32)
    assert q.dtype == torch.uint8 and tuple(q.shape) == (2, 2 * h, 2 * w, 3)
    assert torch.equal(q, signal_conv.quantize_image(f))
    assert int(q.min()) == 0 and int(q.max()) == 255      # both saturations exercised
  # fallback: a layer the scatter kernel does not take
  s = tfc.SignalConv2D(8, (3, 3), corr=True, padding="same_zeros")
  s.build(3, "cuda", g)
  x = torch.rand((1, 9, 9, 3), generator=g).cuda()
  assert torch.equal(s(x, out_dtype=torch.uint8), signal_conv.quantize_image(s(x)))


def test_entropy_models_training_true_is_differentiable(tfc):
  """`entropy_bottleneck(y, training=True)` / `conditional(y, training=True)` are the
  reference's differentiable training entry points (entropy_models.py:148-187): the rate term
  must reach the inputs, the scale and the density variables, and the cached parameters must
  follow the variables."""
  from compression_b200.python.layers import entropy_models
  g = torch.Generator().manual_seed(2)
  eb = entropy_models.EntropyBottleneck()
  eb.build(8, "cuda", g)
  for group in (eb.matrices, eb.biases, eb.factors):
    for i in range(len(group)):
      group[i] = group[i].clone().requires_grad_()
  y = (torch.randn((2, 4, 4, 8), generator=g) * 3).cuda().requires_grad_()
  y_tilde, lik = eb(y, training=True)
  assert lik.requires_grad and y_tilde.requires_grad
  torch.log(lik).sum().backward()
  assert y.grad is not None and float(y.grad.abs().sum()) > 0
  assert all(m.grad is not None and float(m.grad.abs().sum()) > 0 for m in eb.matrices)
  # no gradient wanted: the plain kernel path, same likelihood values for the same noise
  with torch.no_grad():
    torch.manual_seed(7)
    _, lik0 = eb(y.detach(), training=True)
  torch.manual_seed(7)
  _, lik1 = eb(y, training=True)
  assert torch.allclose(lik0, lik1.detach(), rtol=1e-5, atol=1e-7)
  # the packed parameters follow an in-place update of a variable
  before = eb(y.detach(), training=False)[1].clone()
  with torch.no_grad():
    eb.biases[0].add_(0.25)
  after = eb(y.detach(), training=False)[1]
  assert not torch.equal(before, after)
  sigma = (torch.rand((2, 4, 4, 8), generator=g) * 2 + 0.05).cuda().requires_grad_()
  gc = entropy_models.GaussianConditional(sigma, np.exp(np.linspace(np.log(0.11), np.log(256), 64)))
  y2 = (torch.randn((2, 4, 4, 8), generator=g) * 2).cuda().requires_grad_()
  _, lik2 = gc(y2, training=True)
  torch.log(lik2).sum().backward()
  assert float(y2.grad.abs().sum()) > 0 and float(sigma.grad.abs().sum()) > 0


@pytest.mark.parametrize("k,su,corr,act", [(9, 4, False, None), (9, 4, True, "relu"),
                                           (5, 4, False, None), (7, 3, False, None),
                                           (9, 2, False, None)])
def test_sub_pixel_path_matches_oracle(tfc, k, su, corr, act):
  """Thin-output transposed layers whose kh * kw * cout exceeds the scatter kernel's 128 columns
  (the 9x9 stride-4 last synthesis layer of bls2017) run as a stride-1 correlation with
  su * su * cout output channels + depth-to-space (kernel path 6): exact integer-valued operands
  (every product and sum is exact in bfloat16 x bfloat16 -> float32), so the result must EQUAL
  the numpy oracle, for float32 / bfloat16 / uint8 outputs, non-square images and batch > 1."""
  import ctypes
  from compression_b200 import _lib
  from compression_b200.python.layers.signal_conv import ConvDesc
  rng = np.random.default_rng(77)
  cin, cout = 128, 3
  kern = (rng.integers(-4, 5, size=(k, k, cin, cout)) / 8.0).astype(np.float32)
  b = (rng.integers(-8, 9, size=cout) / 4.0).astype(np.float32)
  layer = tfc.SignalConv2D(cout, (k, k), corr=corr, strides_up=su, padding="same_zeros",
                           use_bias=True, activation=torch.relu if act == "relu" else None,
                           kernel_parameterizer=None, kernel_initializer=_static(kern),
                           bias_initializer=_static(b))
  layer.build(cin, "cuda")
  for n, h, w in ((2, 16, 24), (3, 9, 5)):
    x = torch.from_numpy(rng.integers(-2, 3, size=(n, h, w, cin)).astype(np.float32)).cuda()
    xb = x.to(torch.bfloat16)
    want = tfc_np.signal_conv2d(x.cpu().numpy(), kern, b, corr, 1, su, "same_zeros")
    want = torch.relu(want).numpy() if act == "relu" else want.numpy()
    got = layer(xb, out_dtype=torch.float32)
    assert got.shape == (n, h * su, w * su, cout)
    np.testing.assert_array_equal(got.cpu().numpy(), want)
    got16 = layer(xb, out_dtype=torch.bfloat16)
    assert torch.equal(got16, torch.from_numpy(want).to(torch.bfloat16).cuda())
  # the path really is 6 for the bls2017 shape and the uint8 image comes out of it quantised
  d = ConvDesc(n=2, h=16, w=24, cin=cin, ho=16 * 4, wo=24 * 4, cout=3, kh=9, kw=9, sd=1, su=4,
               pad_top=4, pad_left=4, act=0, in_dtype=1, out_dtype=0, impl=0, flags=0)
  assert _lib.lib().tfc_signal_conv2d_path(ctypes.byref(d)) == 6
  if k == 9 and su == 4 and act is None:
    small = tfc.SignalConv2D(cout, (k, k), corr=corr, strides_up=su, padding="same_zeros",
                             use_bias=True, kernel_parameterizer=None,
                             kernel_initializer=_static(kern / 64), bias_initializer=_static(b / 8 + .5))
    small.build(cin, "cuda")
    y = small(xb, out_dtype=torch.float32)
    y8 = small(xb, out_dtype=torch.uint8)
    assert y8.dtype == torch.uint8
    assert torch.equal(y8, torch.clamp(torch.round(y * 255), 0, 255).to(torch.uint8))


def _separable_expected(x, kern, bias, corr, sd, su, padding, extra_pad_end=True):
  """signal_conv_test.py:47-78: channel c with its filter q lands in output channel c * M + q."""
  c, m = kern.shape[2], kern.shape[3]
  cols = []
  for ci in range(c):
    y = tfc_np.signal_conv2d(x[..., ci:ci + 1], kern[:, :, ci:ci + 1, :], None, corr, sd, su,
                             padding, extra_pad_end)
    cols.append(y)
  out = torch.stack(cols, dim=3).reshape(cols[0].shape[:3] + (c * m,))
  return out + (torch.as_tensor(bias, dtype=out.dtype) if bias is not None else 0)


@pytest.mark.parametrize("corr,sd,su,mult,padding,k", [
    (True, 1, 1, 2, "same_zeros", 3), (False, 2, 1, 3, "same_zeros", 5), (True, 2, 1, 1, "valid", 4),
    (False, 1, 2, 1, "same_zeros", 5), (True, 1, 3, 1, "valid", 3), (False, 1, 1, 2, "same_reflect", 3)])
def test_channel_separable_matches_oracle(tfc, corr, sd, su, mult, padding, k):
  """signal_conv.py:466-485 / :590-596: the depthwise layer against the dense oracle
  applied channel by channel, as the reference's own test builds its expectation."""
  rng = np.random.default_rng(5)
  cin = 6
  kern = rng.normal(size=(k, k, cin, mult)).astype(np.float32)
  b = rng.normal(size=cin * mult).astype(np.float32)
  layer = tfc.SignalConv2D(mult, (k, k), corr=corr, strides_down=sd, strides_up=su, padding=padding,
                           channel_separable=True, use_bias=True, kernel_parameterizer=None,
                           kernel_initializer=_static(kern), bias_initializer=_static(b))
  x = rng.normal(size=(2, 11, 14, cin)).astype(np.float32)
  got = layer(torch.from_numpy(x).cuda())
  want = _separable_expected(x, kern, b, corr, sd, su, padding)
  assert tuple(got.shape) == tuple(want.shape) == tuple(layer.compute_output_shape(x.shape))
  np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=1e-4, atol=1e-4)


def test_channel_separable_upsampled_needs_one_filter(tfc):
  layer = tfc.SignalConv2D(2, (3, 3), strides_up=2, channel_separable=True, padding="same_zeros")
  with pytest.raises(NotImplementedError):
    layer.build(4, "cuda")


@pytest.mark.parametrize("corr,sd,su,padding,k,sep", [
    (True, 1, 1, "same_zeros", 5, False), (False, 2, 1, "same_zeros", 5, False),
    (False, 1, 2, "same_zeros", 4, False), (False, 1, 3, "valid", 3, False),
    (True, 3, 1, "valid", 5, False), (False, 2, 1, "same_zeros", 3, True),
    (True, 1, 2, "same_zeros", 5, True)])
def test_signal_conv1d_matches_oracle(tfc, corr, sd, su, padding, k, sep):
  """signal_conv.py:848-863: 1-D layers, dense and channel separable, against the 2-D
  oracle over an image of height 1 with a kernel of height 1 (what the reference itself does
  where TF has no 1-D op)."""
  rng = np.random.default_rng(8)
  cin, filters = 5, (1 if sep else 7)
  kern = rng.normal(size=(1, k, cin, filters)).astype(np.float32)
  cout = cin * filters if sep else filters
  b = rng.normal(size=cout).astype(np.float32)
  layer = tfc.SignalConv1D(filters, k, corr=corr, strides_down=sd, strides_up=su, padding=padding,
                           channel_separable=sep, use_bias=True, kernel_parameterizer=None,
                           kernel_initializer=_static(kern), bias_initializer=_static(b))
  x = rng.normal(size=(3, 23, cin)).astype(np.float32)
  got = layer(torch.from_numpy(x).cuda())
  if sep:
    want = _separable_expected(x[:, None], kern, b, corr, sd, su, padding)[:, 0]
  else:
    want = tfc_np.signal_conv2d(x[:, None], kern, b, corr, sd, su, padding)[:, 0]
  assert tuple(got.shape) == tuple(want.shape) == tuple(layer.compute_output_shape(x.shape))
  np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=2e-2 if not sep else 1e-4,
                             atol=2e-2 if not sep else 1e-4)
  cf = tfc.SignalConv1D(filters, k, corr=corr, strides_down=sd, strides_up=su, padding=padding,
                        channel_separable=sep, use_bias=True, kernel_parameterizer=None,
                        kernel_initializer=_static(kern), bias_initializer=_static(b),
                        data_format="channels_first")
  got_cf = cf(torch.from_numpy(x).cuda().permute(0, 2, 1).contiguous())
  assert torch.equal(got_cf.permute(0, 2, 1), got)


def test_entropy_bottleneck_update_without_readback():
  """update() builds the tables for the max_length it saw last and checks the true one
  asynchronously (no host round trip in a training loop); a reader of the tables settles the
  check.  When the quantiles move the longest row — wider AND narrower — the tables read
  afterwards must be those of a layer that updates synchronously from the same variables."""
  from compression_b200.python.layers import entropy_models
  g = torch.Generator().manual_seed(21)
  eb = entropy_models.EntropyBottleneck()
  eb.build(16, "cuda", g)
  eb.update()                                    # first update: the one synchronous read
  for scale, shift in ((3.0, 0.0), (0.4, 0.0), (1.0, 2.5), (1.0, 0.0)):
    q = eb.quantiles.clone()
    q[:, 0, 0] = -10.0 * scale + shift
    q[:, 0, 2] = 10.0 * scale + shift
    q[3, 0, 2] += 7.0                            # one row longer than the others
    eb.quantiles = q
    eb.update()                                  # sized by the previous max_length
    eb.update()                                  # twice in a row: an unread check is dropped
    fresh = entropy_models.EntropyBottleneck()
    fresh.build(16, "cuda", torch.Generator().manual_seed(21))
    for name in ("matrices", "biases", "factors"):
      setattr(fresh, name, [t.clone() for t in getattr(eb, name)])
    fresh.quantiles = q.clone()
    fresh.update()                               # first update of a new layer: synchronous
    assert torch.equal(eb._cdf_length, fresh._cdf_length)
    assert eb._quantized_cdf.shape == fresh._quantized_cdf.shape
    assert torch.equal(eb._quantized_cdf, fresh._quantized_cdf)
    assert int(eb._cdf_length.max()) == eb._quantized_cdf.shape[1]


def test_thin_conv_tmem_epilogue_matches_smem_epilogue():
  """conv_thin_kernel's default GDN epilogue for bfloat16 outputs (x^2 handed to GEMM2 through
  tensor memory, bias and beta folded into the GEMMs as three bfloat16 pieces each, sixteen
  epilogue warps) against the older one (debug bit 32: x^2 through shared memory, bias and beta
  added in float32 by the epilogue): the same values up to the float32 summation order, i.e. the
  bfloat16 outputs are equal except where x sat on a rounding boundary, and then one or two ulps apart.
  float32 and uint8 images; float32 outputs take the older epilogue either way."""
  from compression_b200 import _lib
  from compression_b200.examples import bmshj2018
  model = bmshj2018.Bmshj2018Model(192, "cuda", seed=5)
  layer = model.analysis_transform.layers[0]
  with torch.no_grad():
    layer.bias_var.copy_(torch.linspace(-0.37, 0.41, 192))  # a bias the fold has to carry
  layer.invalidate()
  g = torch.Generator().manual_seed(9)
  xf = torch.rand((3, 72, 104, 3), generator=g).cuda()
  xu = torch.randint(0, 256, (3, 72, 104, 3), generator=g, dtype=torch.uint8).cuda()
  for x in (xf, xu):
    for out_dtype in (torch.bfloat16, torch.float32):
      got = layer(x, out_dtype=out_dtype)
      _lib.check(_lib.lib().tfc_configure(99, 32))
      try:
        want = layer(x, out_dtype=out_dtype)
        torch.cuda.synchronize()
      finally:
        _lib.check(_lib.lib().tfc_configure(99, 0))
      if out_dtype == torch.float32:
        assert torch.equal(got, want)
        continue
      g32, w32 = got.float(), want.float()
      assert (g32 == w32).float().mean().item() > 0.995
      assert ((g32 - w32).abs() <= w32.abs() * 2.0 ** -6 + 2e-6).all()


@pytest.mark.parametrize("model_name,shape", [
    ("bmshj2018", (1, 40, 40)),      # fewer tiles than CTAs: most CTAs run zero tiles
    ("bmshj2018", (40, 72, 104)),    # 800 tiles on 148 CTAs: five or six per CTA (odd and even counts)
    ("bmshj2018", (23, 100, 60)),    # ragged tiles on both borders, three or four per CTA
    ("bmshj2018", (9, 250, 330)),    # many tiles per CTA
    ("bls2017", (16, 128, 192)),     # 9x9 stride 4, 128 filters: the beta rows share W's last K step
    ("bls2017", (5, 100, 76)),
])
def test_thin_conv_tmem_epilogue_tile_counts(model_name, shape):
  """The pipelined tensor-memory epilogue of the thin-input GDN layer over CTAs with zero, one, odd
  and even numbers of tiles and ragged borders, both first layers (5x5 / 192 filters, 9x9 / 128
  filters), against the older epilogue (debug bit 32): equal except where x sat on a bfloat16
  rounding boundary (a few outputs in ten million), and then at most two ulps apart."""
  from compression_b200 import _lib
  from compression_b200.examples import bls2017, bmshj2018
  if model_name == "bmshj2018":
    layer = bmshj2018.Bmshj2018Model(192, "cuda", seed=3).analysis_transform.layers[0]
  else:
    layer = bls2017.Bls2017Model(128, "cuda", seed=3).analysis_transform.layers[0]
  with torch.no_grad():
    layer.bias_var.copy_(torch.linspace(-0.21, 0.33, layer.bias_var.numel()))
  layer.invalidate()
  g = torch.Generator().manual_seed(17)
  x = torch.rand(shape + (3,), generator=g).cuda()
  got = layer(x, out_dtype=torch.bfloat16)
  again = layer(x, out_dtype=torch.bfloat16)
  _lib.check(_lib.lib().tfc_configure(99, 32))
  try:
    want = layer(x, out_dtype=torch.bfloat16)
    torch.cuda.synchronize()
  finally:
    _lib.check(_lib.lib().tfc_configure(99, 0))
  assert torch.equal(got, again)
  g32, w32 = got.float(), want.float()
  assert torch.isfinite(g32).all()
  assert (g32 == w32).float().mean().item() > 0.995
  # x one bfloat16 ulp apart gives y up to two ulps apart after its own rounding; where bias and
  # the convolution sum cancel, x itself is only defined to the float32 rounding of its terms
  assert ((g32 - w32).abs() <= w32.abs() * 2.0 ** -6 + 2e-6).all()


@pytest.mark.parametrize("corr,sd,su,act", [(True, 2, 1, "gdn"), (False, 1, 2, "igdn"),
                                            (True, 1, 1, "relu"), (True, 2, 1, None)])
def test_cta_pair_conv_equals_single_cta(tfc, corr, sd, su, act):
  """TFC_CFG_CONV_PAIR: the implicit GEMM on CTA pairs (tcgen05.mma.cta_group::2, M = 256 over
  two adjacent tiles, each CTA holding half of the weight tile) sums the same products in the
  same order as the one-CTA kernel: identical outputs, bit for bit, including ragged tiles and an
  odd tile count per image (pairs then straddle images)."""
  rng = np.random.default_rng(41)
  k, cin, cout = (3 if act == "relu" else 5), 192, 192
  kern = (rng.integers(-8, 9, size=(k, k, cin, cout)) / 16.0).astype(np.float32)
  b = (rng.integers(-8, 9, size=cout) / 4.0).astype(np.float32)
  activation = {"gdn": tfc.GDN(), "igdn": tfc.GDN(inverse=True), "relu": torch.relu, None: None}[act]
  layer = tfc.SignalConv2D(cout, (k, k), corr=corr, strides_down=sd, strides_up=su,
                           padding="same_zeros", use_bias=True, activation=activation,
                           kernel_parameterizer=None, kernel_initializer=_static(kern),
                           bias_initializer=_static(b), impl=2)
  layer.build(cin, "cuda")
  for n, h, w in ((4, 40, 56), (3, 24, 40), (2, 16, 16)):
    x = torch.from_numpy(rng.integers(-3, 4, size=(n, h, w, cin)).astype(np.float32)).cuda()
    x = x.to(torch.bfloat16)
    outs = {}
    for mode in (1, 2):
      tfc.configure(conv_pair=mode)
      try:
        outs[mode] = layer(x, out_dtype=torch.bfloat16)
      finally:
        tfc.configure(conv_pair=0)
    assert torch.equal(outs[1], outs[2]), (n, h, w)
  want = tfc_np.signal_conv2d(x.float().cpu().numpy(), kern, b, corr, sd, su, "same_zeros")
  if act in ("gdn", "igdn"):
    beta, gamma = [t.cpu().numpy() for t in activation.effective()]
    want = tfc_np.gdn(want, beta, gamma, act == "igdn")
  elif act == "relu":
    want = torch.relu(want)
  want = want.numpy()
  got = outs[2].float().cpu().numpy()
  assert np.abs(got - want).max() / (np.abs(want).max() + 1e-6) <= 2e-2
