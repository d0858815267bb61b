"""Pins the CPU checkers to the reference's own golden vectors and the
port to the compiled reference on random streams.  CPU only."""
import os

import numpy as np
import pytest

from tests.golden import tfc_goldens as G


def test_eb_golden_decode_and_encode(oracle_any):
  # entropy_models_test.py:179-221
  for b in range(2):
    sym = oracle_any.ui_decode(G.EB_BITSTRINGS[b], G.EB_INDEX, G.EB_CDF, G.EB_CDF_LENGTH,
                               G.EB_OFFSET)
    np.testing.assert_array_equal(sym, G.EB_EXPECTED[b].astype(np.int32))
    enc = oracle_any.ui_encode(G.EB_EXPECTED[b].astype(np.int32), G.EB_INDEX, G.EB_CDF,
                               G.EB_CDF_LENGTH, G.EB_OFFSET)
    assert enc == G.EB_BITSTRINGS[b]


@pytest.mark.parametrize("name", sorted(G.CONDITIONALS))
def test_conditional_golden_decode_and_encode(oracle_any, name):
  # entropy_models_test.py:407-425 with :509-533, :544-574, :585-615
  g = G.CONDITIONALS[name]
  for b in range(2):
    sym = oracle_any.ui_decode(g["bitstrings"][b], G.COND_INDEXES[b], g["cdf"],
                               g["cdf_length"], g["offset"])
    np.testing.assert_array_equal(sym, G.COND_EXPECTED[b].astype(np.int32))
    enc = oracle_any.ui_encode(G.COND_EXPECTED[b].astype(np.int32), G.COND_INDEXES[b],
                               g["cdf"], g["cdf_length"], g["offset"])
    assert enc == g["bitstrings"][b]


def _random_tables(rng, rows, max_len, precision):
  total = 1 << precision
  cols = max_len
  cdf = np.zeros((rows, cols), np.int32)
  size = np.zeros(rows, np.int32)
  for r in range(rows):
    n = int(rng.integers(3, min(cols, total + 1) + 1))  # cdf_size in [3, cols]
    cuts = np.sort(rng.choice(np.arange(1, total), size=n - 2, replace=False)) if n > 2 else []
    cdf[r, :n] = np.concatenate([[0], cuts, [total]])
    size[r] = n
  offset = rng.integers(-5, 6, size=rows).astype(np.int32)
  return cdf, size, offset


@pytest.mark.skipif(not os.path.isdir("/root/reference") and
                    not os.path.exists(os.path.join(os.path.dirname(__file__), "..", "oracle",
                                                    "_ref", "libtfc_ref.so")),
                    reason="needs the compiled reference")
def test_port_matches_reference_random():
  from oracle import oracle_lib
  ref = oracle_lib.load("reference_dbg")
  port = oracle_lib.load("port")
  rng = np.random.default_rng(0)
  for trial in range(300):
    precision = int(rng.integers(1, 17))
    ow = int(rng.integers(1, 17))
    rows = int(rng.integers(1, 6))
    max_len = int(rng.integers(3, min(40, (1 << precision) + 1) + 1))
    cdf, size, offset = _random_tables(rng, rows, max_len, precision)
    n = int(rng.integers(0, 400))
    index = rng.integers(0, rows, size=n).astype(np.int32)
    # the reference's digit-count loop shifts a uint32 by widths*overflow_width, which is
    # undefined (and in practice never terminates) once that reaches 32: keep the zig-zagged
    # excess below 2**(ow*floor(31/ow))  (unbounded_index_range_coding_kernels.cc:233-236)
    limit = 1 << (ow * (31 // ow))
    spread = int(min(rng.choice([2, 10, 1000, 100000]), limit // 4 - 8))
    data = (rng.integers(-spread, spread + 1, size=n) + offset[index]).astype(np.int32)
    a = ref.ui_encode(data, index, cdf, size, offset, precision, ow)
    b = port.ui_encode(data, index, cdf, size, offset, precision, ow)
    assert a == b, (trial, precision, ow)
    np.testing.assert_array_equal(port.ui_decode(a, index, cdf, size, offset, precision, ow),
                                  data)
    np.testing.assert_array_equal(ref.ui_decode(a, index, cdf, size, offset, precision, ow),
                                  data)


def test_port_matches_reference_bounded_and_pmf():
  from oracle import oracle_lib
  if "reference" not in oracle_lib.available():
    pytest.skip("no compiled reference")
  ref = oracle_lib.load("reference_dbg")
  port = oracle_lib.load("port")
  rng = np.random.default_rng(1)
  for trial in range(100):
    precision = int(rng.integers(2, 17))
    m = int(rng.integers(2, min(30, 1 << precision) + 1))
    shape = tuple(int(x) for x in rng.integers(1, 6, size=int(rng.integers(1, 4))))
    lead = tuple(s if rng.random() < 0.5 else 1 for s in shape)
    pmf = rng.random(lead + (m,)).astype(np.float32) + 1e-3
    pmf /= pmf.sum(-1, keepdims=True)
    cdf_r = ref.pmf_to_cdf(pmf, precision)
    cdf_p = port.pmf_to_cdf(pmf, precision)
    np.testing.assert_array_equal(cdf_r, cdf_p)
    data = rng.integers(0, m, size=shape).astype(np.int16)
    a = ref.rc_encode(data, cdf_r, precision)
    assert a == port.rc_encode(data, cdf_r, precision)
    np.testing.assert_array_equal(port.rc_decode(a, shape, cdf_r, precision), data)
    np.testing.assert_array_equal(ref.rc_decode(a, shape, cdf_r, precision), data)


def test_pmf_port_matches_reference_ties_and_long_rows():
  """std::sort tie order matters: symmetric (tie-heavy) and long rows."""
  from oracle import oracle_lib
  if "reference" not in oracle_lib.available():
    pytest.skip("no compiled reference")
  ref = oracle_lib.load("reference")
  port = oracle_lib.load("port")
  rng = np.random.default_rng(2)
  for n in (2, 3, 16, 17, 22, 64, 257, 1481):
    k = np.arange(n, dtype=np.float64) - (n - 1) / 2
    for sigma in (0.3, 2.0, n / 6.0 + 0.1):
      p = np.exp(-0.5 * (k / sigma) ** 2)
      p = (p / p.sum()).astype(np.float32)
      np.testing.assert_array_equal(ref.pmf_to_cdf(p[None], 16), port.pmf_to_cdf(p[None], 16))
    p = rng.random((8, n)).astype(np.float32)
    p /= p.sum(-1, keepdims=True)
    np.testing.assert_array_equal(ref.pmf_to_cdf(p, 16), port.pmf_to_cdf(p, 16))
    p *= 1.3  # over-sum
    np.testing.assert_array_equal(ref.pmf_to_cdf(p, 12), port.pmf_to_cdf(p, 12))
