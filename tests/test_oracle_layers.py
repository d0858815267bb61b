"""Pins the float restatement in oracle/tfc_np.py to the properties and golden tables the
reference's own tests hold.  CPU only."""
import numpy as np
import pytest
import scipy.signal
import scipy.stats

from oracle import tfc_np
from tests.golden import tfc_goldens as G


def upsample(x, s, extra_pad_end):
  """signal_conv_test.py:32-45 (channels_last here)."""
  n, h, w, c = x.shape
  hu, wu = (h * s, w * s) if extra_pad_end else ((h - 1) * s + 1, (w - 1) * s + 1)
  up = np.zeros((n, hu, wu, c), x.dtype)
  up[:, ::s, ::s] = x
  return up


def scipy_valid(corr, x, k, sd, su, extra_pad_end):
  """signal_conv_test.py:47-79: per-channel scipy correlate/convolve of the zero-stuffed
  input, mode 'valid', then [::sd]."""
  f = scipy.signal.correlate if corr else scipy.signal.convolve
  if su != 1:
    x = upsample(x, su, extra_pad_end)
  n, _, _, cin = x.shape
  cout = k.shape[-1]
  ref = f(x[0, :, :, 0], k[:, :, 0, 0], mode="valid")[::sd, ::sd]
  out = np.zeros((n,) + ref.shape + (cout,), np.float64)
  for b in range(n):
    for o in range(cout):
      for c in range(cin):
        out[b, :, :, o] += f(x[b, :, :, c], k[:, :, c, o], mode="valid")[::sd, ::sd]
  return out


VALID_CASES = [
    # corr, kernel, sd, su, extra_pad_end
    (True, (3, 3), 1, 1, True), (False, (3, 3), 1, 1, True), (True, (5, 5), 2, 1, True),
    (False, (5, 5), 1, 2, True), (True, (5, 5), 1, 2, True), (False, (5, 5), 1, 2, False),
    (False, (4, 4), 1, 2, True), (False, (9, 9), 1, 4, True), (True, (9, 9), 4, 1, True),
    (False, (3, 5), 2, 3, True), (True, (2, 4), 2, 1, True), (False, (2, 2), 1, 1, True),
]


@pytest.mark.parametrize("corr,ks,sd,su,epe", VALID_CASES)
def test_conv_valid_matches_scipy(corr, ks, sd, su, epe):
  rng = np.random.default_rng(0)
  x = rng.integers(0, 32, size=(2, 13, 11, 3)).astype(np.float64)
  k = rng.integers(0, 16, size=ks + (3, 4)).astype(np.float64)
  got = tfc_np.signal_conv2d(x, k, None, corr, sd, su, "valid", epe).numpy()
  want = scipy_valid(corr, x, k, sd, su, epe)
  assert got.shape == want.shape
  np.testing.assert_allclose(got, want, rtol=0, atol=1e-3)


@pytest.mark.parametrize("corr", [True, False])
@pytest.mark.parametrize("ks,sd,su", [((5, 5), 2, 1), ((5, 5), 1, 2), ((3, 3), 1, 1),
                                      ((9, 9), 4, 1), ((9, 9), 1, 4), ((5, 5), 2, 2)])
def test_conv_same_identity_kernel(corr, ks, sd, su):
  """signal_conv_test.py:131-183: identity kernel + same_zeros reproduces the up/downsampled
  input."""
  x = np.arange(2 * 12 * 8, dtype=np.float64).reshape(2, 12, 8, 1)
  k = np.zeros(ks + (1, 1))
  k[ks[0] // 2, ks[1] // 2] = 1
  got = tfc_np.signal_conv2d(x, k, None, corr, sd, su, "same_zeros", True).numpy()
  want = upsample(x, su, True)[:, ::sd, ::sd]
  np.testing.assert_allclose(got, want, rtol=0, atol=1e-3)


def test_gdn_closed_forms():
  """gdn_test.py:49-75: default init beta=1, gamma=0.1*I."""
  x = np.random.default_rng(1).uniform(size=(1, 2, 3, 4))
  beta, gamma = np.ones(4), 0.1 * np.eye(4)
  np.testing.assert_allclose(tfc_np.gdn(x, beta, gamma).numpy(), x / np.sqrt(1 + .1 * x ** 2),
                             atol=1e-6)
  np.testing.assert_allclose(tfc_np.gdn(x, beta, gamma, True).numpy(),
                             x * np.sqrt(1 + .1 * x ** 2), atol=1e-6)


@pytest.mark.parametrize("kind", sorted(G.CONDITIONALS))
def test_conditional_tables_match_golden_cdfs(kind, oracle):
  """G5: the CDF arrays of entropy_models_test.py:519-526, :554-567, :595-608 are reproduced by
  SymmetricConditional.build restated + the compiled PerShard."""
  g = G.CONDITIONALS[kind]
  cdf, length, offset = tfc_np.conditional_tables(
      G.COND_SCALE_TABLE, lambda p: oracle.pmf_to_cdf(p[None], 16)[0], kind)
  np.testing.assert_array_equal(length, g["cdf_length"])
  np.testing.assert_array_equal(offset, g["offset"])
  np.testing.assert_array_equal(cdf, g["cdf"])


@pytest.mark.parametrize("kind,dist", [("gaussian", scipy.stats.norm),
                                       ("logistic", scipy.stats.logistic),
                                       ("laplacian", scipy.stats.laplace)])
def test_conditional_likelihood_vs_scipy(kind, dist):
  """entropy_models_test.py:448-463."""
  rng = np.random.default_rng(2)
  y = rng.normal(scale=5, size=(3, 50)).astype(np.float32)
  scale = np.exp(rng.normal(size=(3, 50))).astype(np.float32)
  values, lik = tfc_np.conditional_call(y, scale, [1e-3], kind, likelihood_bound=0)
  x = np.abs(values.astype(np.float64))
  want = dist.cdf(x + .5, scale=scale.astype(np.float64)) - dist.cdf(x - .5, scale=scale.astype(np.float64))
  np.testing.assert_allclose(lik, want, rtol=1e-5, atol=1e-12)


def test_conditional_indexes_definition():
  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))
  s = np.array([-1, 0, 0.1, 0.11, 0.12, 1.0, 255.9, 256, 300, np.float32(table[10])], np.float32)
  idx = tfc_np.conditional_indexes(s, table)
  t32 = table.astype(np.float32)
  for v, i in zip(np.maximum(s, t32[0]), idx):
    assert i == min(int(np.sum(t32[:-1] < v)), 63)


def test_eb_density_normalised_and_quantised():
  """entropy_models_test.py:42-72 (median offsets) and :231-242 (pmf sums to ~1)."""
  rng = np.random.default_rng(3)
  c = 4
  filters = (1, 3, 3, 3, 1)
  scale = 10 ** (1 / 4)
  mats = [np.full((c, filters[i + 1], filters[i]), np.log(np.expm1(1 / scale / filters[i + 1])))
          for i in range(4)]
  biases = [rng.uniform(-.5, .5, size=(c, filters[i + 1], 1)) for i in range(4)]
  factors = [np.zeros((c, filters[i + 1], 1)) for i in range(3)]
  m, b, f = tfc_np.eb_params64(mats, biases, factors)
  medians = rng.uniform(-.4, .4, size=c).astype(np.float32)
  x = rng.uniform(-20, 20, size=(5, 7, c)).astype(np.float32)
  values, lik = tfc_np.eb_call(x, medians, m, b, f)
  np.testing.assert_allclose(values - medians, np.round(values - medians), atol=1e-5)
  assert np.all(np.abs(values - x) <= .5 + 1e-5)
  grid = np.arange(-200, 201, dtype=np.float32)[:, None] * np.ones((1, c), np.float32)
  _, pm = tfc_np.eb_call(grid, np.zeros(c, np.float32), m, b, f, likelihood_bound=0)
  np.testing.assert_allclose(pm.sum(0), 1, atol=1e-4)
  sym = tfc_np.eb_symbols(x, medians)
  np.testing.assert_allclose(sym + medians, values, atol=1e-5)


def test_cpu_bls2017_codec_round_trip():
  """BASELINE configs[0] on the host (oracle/cpu_codec.py CpuBls2017, restating
  /root/reference/examples/bls2017.py:205-305): the decoder returns the coded symbols and an
  image of the input's size; different images give different strings."""
  from oracle import cpu_codec, oracle_lib
  coder = oracle_lib.load("best")
  codec = cpu_codec.CpuBls2017(cpu_codec.random_weights_bls2017(128, 0), coder, threads=2)
  rng = np.random.default_rng(0)
  x = rng.random((2, 72, 104, 3)).astype(np.float32)          # not multiples of 16
  c = codec.compress(x)
  assert tuple(c["y_shape"]) == (5, 7) and c["y_sym"].shape == (2, 5, 7, 128)
  x_hat, y_sym = codec.decompress(c)
  assert x_hat.shape == x.shape and np.isfinite(x_hat).all()
  np.testing.assert_array_equal(y_sym, c["y_sym"])
  n = c["string"][1]
  assert bytes(c["string"][0][0, :n[0]]) != bytes(c["string"][0][1, :n[1]])
