"""GPU parity tests of the entropy-coding kernels, through the C ABI, against the CPU checker
(oracle/_ref when built, else the port) and the reference's golden vectors."""
import numpy as np
import pytest
import torch

from tests.golden import tfc_goldens as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tfc():
  import compression_b200 as tfc_mod
  return tfc_mod


def test_eb_golden(tfc):
  # entropy_models_test.py:179-221 — decode the stored strings, and re-encode to the same bytes
  for b in range(2):
    sym = tfc.unbounded_index_range_decode(G.EB_BITSTRINGS[b], G.EB_INDEX, G.EB_CDF,
                                           G.EB_CDF_LENGTH, G.EB_OFFSET, 16, 4)
    np.testing.assert_array_equal(sym.cpu().numpy(), G.EB_EXPECTED[b].astype(np.int32))
    enc = tfc.unbounded_index_range_encode(G.EB_EXPECTED[b].astype(np.int32), G.EB_INDEX,
                                           G.EB_CDF, G.EB_CDF_LENGTH, G.EB_OFFSET, 16, 4)
    assert enc == G.EB_BITSTRINGS[b]


@pytest.mark.parametrize("name", sorted(G.CONDITIONALS))
def test_conditional_golden(tfc, name):
  g = G.CONDITIONALS[name]
  for b in range(2):
    sym = tfc.unbounded_index_range_decode(g["bitstrings"][b], G.COND_INDEXES[b], g["cdf"],
                                           g["cdf_length"], g["offset"], 16, 4)
    np.testing.assert_array_equal(sym.cpu().numpy(), G.COND_EXPECTED[b].astype(np.int32))
    enc = tfc.unbounded_index_range_encode(G.COND_EXPECTED[b].astype(np.int32),
                                           G.COND_INDEXES[b], g["cdf"], g["cdf_length"],
                                           g["offset"], 16, 4)
    assert enc == g["bitstrings"][b]


def _random_tables(rng, rows, max_len, precision):
  total = 1 << precision
  cdf = np.zeros((rows, max_len), np.int32)
  size = np.zeros(rows, np.int32)
  for r in range(rows):
    n = int(rng.integers(3, min(max_len, total + 1) + 1))
    cuts = np.sort(rng.choice(np.arange(1, total), size=n - 2, replace=False))
    cdf[r, :n] = np.concatenate([[0], cuts, [total]])
    size[r] = n
  offset = rng.integers(-5, 6, size=rows).astype(np.int32)
  return cdf, size, offset


@pytest.fixture(params=[0, 2], ids=["warps_per_stream", "lane_per_stream"])
def coder_lanes(request, tfc):
  """Both unbounded coder designs: 2-3 warps per stream (rc_v4.cuh) and one stream per lane
  (rc_lanes.cuh, automatic from 2048 streams on, forced here)."""
  tfc.configure(coder_lanes=request.param)
  yield request.param
  tfc.configure(coder_lanes=0)


def test_unbounded_random_sweep(tfc, oracle, coder_lanes):
  """>=10^4 streams over precision 1..16, overflow_width 1..16, escapes of both signs, ragged
  cdf_size, empty and tiny streams."""
  from compression_b200.python.ops import range_coding_ops as rc
  rng = np.random.default_rng(0)
  total_streams = 0
  for trial in range(120):
    precision = int(rng.integers(1, 17))
    ow = int(rng.integers(1, 17))
    rows = int(rng.integers(1, 9))
    max_len = int(rng.integers(3, min(80, (1 << precision) + 1) + 1))
    cdf, size, offset = _random_tables(rng, rows, max_len, precision)
    S = int(rng.choice([1, 3, 100, 150]))
    n = int(rng.choice([0, 1, 31, 32, 33, 127, 128, 129, 300]))
    shared = bool(rng.integers(0, 2))
    index = rng.integers(0, rows, size=(n,) if shared else (S, n)).astype(np.int32)
    limit = 1 << (ow * (31 // ow))
    spread = int(min(rng.choice([2, 10, 1000, 100000]), limit // 4 - 8))
    data = (rng.integers(-spread, spread + 1, size=(S, n)) + offset[index]).astype(np.int32)
    out, nbytes = rc.unbounded_index_range_encode_batch(data, index, cdf, size, offset,
                                                        precision, ow, debug_level=1)
    got = rc.strings_to_host(out, nbytes)
    for s in range(S):
      want = oracle.ui_encode(data[s], index if shared else index[s], cdf, size, offset,
                              precision, ow)
      assert got[s] == want, (trial, s, precision, ow)
    dec = rc.unbounded_index_range_decode_batch(out, nbytes, index, cdf, size, offset,
                                                precision, ow, debug_level=1,
                                                shared_index=shared)
    np.testing.assert_array_equal(dec.cpu().numpy(), data)
    total_streams += S
  assert total_streams >= 5000


@pytest.mark.parametrize("pack", [2, 3, 4, 8])
def test_packed_coder_ctas_match_reference(tfc, oracle, pack):
  """Packed coder CTAs (tfc.configure coder_streams_per_cta >= 2, what the pipeline runs): same
  bytes as the reference for odd and even stream counts (a last CTA that is not full), empty
  streams, streams whose put counts differ widely inside one CTA (escape bursts in some streams
  only), several rounds per stream, every overflow width; the decoder returns the symbols."""
  from compression_b200.python.ops import range_coding_ops as rc
  rng = np.random.default_rng(100 + pack)
  tfc.configure(coder_lanes=1, coder_streams_per_cta=pack)
  try:
    for trial in range(24):
      precision = int(rng.integers(8, 17))
      ow = int(rng.choice([1, 2, 4, 7, 16]))
      rows = int(rng.integers(1, 7))
      cdf, size, offset = _random_tables(rng, rows, int(rng.integers(3, 40)), precision)
      S = int(rng.choice([1, 2, 3, 5, 7, 8, 17, 64]))
      n = int(rng.choice([0, 1, 33, 500, 3000, 9000]))
      index = rng.integers(0, rows, size=(S, n)).astype(np.int32)
      limit = 1 << (ow * (31 // ow))
      data = rng.integers(-2, 3, size=(S, n))
      burst = rng.integers(0, S, size=max(1, S // 2))          # these streams escape all the time
      data[burst] = rng.integers(-min(50000, limit // 4 - 8), min(50000, limit // 4 - 8) + 1,
                                 size=(len(burst), n))
      data = (data + offset[index]).astype(np.int32)
      out, nbytes = rc.unbounded_index_range_encode_batch(data, index, cdf, size, offset,
                                                          precision, ow, debug_level=0)
      got = rc.strings_to_host(out, nbytes)
      for s_ in range(0, S, max(1, S // 6)):
        want = oracle.ui_encode(data[s_], index[s_], cdf, size, offset, precision, ow)
        assert got[s_] == want, (trial, s_, S, n, precision, ow)
      dec = rc.unbounded_index_range_decode_batch(out, nbytes, index, cdf, size, offset,
                                                  precision, ow, debug_level=0)
      np.testing.assert_array_equal(dec.cpu().numpy(), data)
  finally:
    tfc.configure(coder_lanes=0, coder_streams_per_cta=0)


def test_unbounded_many_streams_capacity_retry(tfc, oracle):
  """2100 ragged streams (the automatic lane-per-stream path), output slab deliberately too
  small: the kernel reports the size it needs per stream and the retry produces the
  reference's bytes."""
  from compression_b200.python.ops import range_coding_ops as rc
  rng = np.random.default_rng(11)
  cdf, size, offset = _random_tables(rng, 6, 40, 16)
  S, n = 2100, 77
  index = rng.integers(0, 6, size=(S, n)).astype(np.int32)
  data = (rng.integers(-60, 61, size=(S, n)) + offset[index]).astype(np.int32)
  out, nbytes = rc.unbounded_index_range_encode_batch(data, index, cdf, size, offset, 16, 4,
                                                      debug_level=0, cap=32)
  got = rc.strings_to_host(out, nbytes)
  assert max(len(g) for g in got) > 32
  for s in list(range(0, S, 50)) + [S - 1]:
    assert got[s] == oracle.ui_encode(data[s], index[s], cdf, size, offset, 16, 4), s
  dec = rc.unbounded_index_range_decode_batch(out, nbytes, index, cdf, size, offset, 16, 4, 0)
  np.testing.assert_array_equal(dec.cpu().numpy(), data)


def test_unbounded_sweep_size_round_trip(tfc, oracle):
  """BASELINE configs[3] scale: 16 384 streams x 4096 symbols over the bmshj2018 Gaussian table
  (rows of 5..1481 entries, ~0.2 % escapes): decode(encode(x)) == x for every stream, and an
  evenly spaced sample of streams is byte-identical to the CPU checker."""
  from compression_b200.python.layers import entropy_models
  from compression_b200.python.ops import range_coding_ops as rc
  dev = torch.device("cuda", 0)
  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))
  gc = entropy_models.GaussianConditional(None, table)
  gc.build_tables(dev)
  cdf, size, off = gc._quantized_cdf, gc._cdf_length, gc._offset
  g = torch.Generator(device=dev).manual_seed(5)
  S, n = 16384, 4096
  idx = torch.randint(0, 64, (S, n), generator=g, dtype=torch.int32, device=dev)
  sig = torch.as_tensor(table, dtype=torch.float32, device=dev)[idx.long()]
  sym = torch.floor(torch.randn((S, n), generator=g, device=dev) * sig + .5).to(torch.int32)
  out, nbytes = rc.unbounded_index_range_encode_batch(sym, idx, cdf, size, off, 16, 4, 0)
  dec = rc.unbounded_index_range_decode_batch(out, nbytes, idx, cdf, size, off, 16, 4, 0)
  assert torch.equal(dec, sym)
  pick = torch.linspace(0, S - 1, 48).long().to(dev)
  got = rc.strings_to_host(out[pick], nbytes[pick])
  c, s_, o = cdf.cpu().numpy(), size.cpu().numpy(), off.cpu().numpy()
  sym_h, idx_h = sym[pick].cpu().numpy(), idx[pick].cpu().numpy()
  for k in range(len(got)):
    assert got[k] == oracle.ui_encode(sym_h[k], idx_h[k], c, s_, o, 16, 4), int(pick[k])


def test_unbounded_long_rows_and_carry(tfc, oracle, coder_lanes):
  """Rows longer than the 32-wide search window (Gaussian-like, up to 1481 entries) and long
  runs of near-certain symbols that keep the encoder in its carry-pending state
  (range_coder.cc:169-210)."""
  from compression_b200.python.ops import range_coding_ops as rc
  rng = np.random.default_rng(3)
  lens = [5, 40, 300, 1481]
  cdf = np.zeros((len(lens), max(lens)), np.int32)
  for r, n in enumerate(lens):
    k = np.arange(n - 2) - (n - 3) / 2.0
    p = np.exp(-0.5 * (k / (n / 8.0 + 0.2)) ** 2)
    p = np.concatenate([p / p.sum() * (1 - 2.0 ** -8), [2.0 ** -8]]).astype(np.float32)
    cdf[r, :n] = oracle.pmf_to_cdf(p[None], 16)[0]
  size = np.array(lens, np.int32)
  offset = -((size - 3) // 2)
  S, n = 37, 2000
  index = rng.integers(0, len(lens), size=(S, n)).astype(np.int32)
  sigma = (size[index] / 8.0 + 0.2)
  data = np.rint(rng.standard_normal((S, n)) * sigma * rng.choice([0.05, 1, 3], size=(S, 1)))
  data = data.astype(np.int32)
  out, nbytes = rc.unbounded_index_range_encode_batch(data, index, cdf, size, offset, 16, 4, 1)
  got = rc.strings_to_host(out, nbytes)
  for s in range(S):
    assert got[s] == oracle.ui_encode(data[s], index[s], cdf, size, offset, 16, 4), s
  dec = rc.unbounded_index_range_decode_batch(out, nbytes, index, cdf, size, offset, 16, 4, 1)
  np.testing.assert_array_equal(dec.cpu().numpy(), data)
  # carry path: symbol 1 of a [0, 1, 65535, 65536] table has probability ~1
  cdf2 = np.array([[0, 1, 65535, 65536], [0, 32768, 65535, 65536]], np.int32)
  size2 = np.array([4, 4], np.int32)
  off2 = np.zeros(2, np.int32)
  data2 = np.ones((8, 5000), np.int32)
  data2[:, ::977] = 0
  data2[3, 100:120] = 7  # escapes in the middle of a run
  idx2 = np.zeros((8, 5000), np.int32)
  idx2[:, ::13] = 1
  out, nbytes = rc.unbounded_index_range_encode_batch(data2, idx2, cdf2, size2, off2, 16, 4, 1)
  got = rc.strings_to_host(out, nbytes)
  for s in range(8):
    assert got[s] == oracle.ui_encode(data2[s], idx2[s], cdf2, size2, off2, 16, 4), s
  dec = rc.unbounded_index_range_decode_batch(out, nbytes, idx2, cdf2, size2, off2, 16, 4, 1)
  np.testing.assert_array_equal(dec.cpu().numpy(), data2)


def test_bounded_broadcast_patterns(tfc, oracle, coder_lanes):
  """range_coding_kernels_test.cc:245-321: no broadcast, 1-axis and 2-axis cdf broadcast."""
  rng = np.random.default_rng(4)
  for trial in range(60):
    precision = int(rng.integers(2, 17))
    m = int(rng.integers(2, min(50, 1 << precision) + 1))
    rank = int(rng.integers(1, 5))
    shape = tuple(int(x) for x in rng.integers(1, 7, size=rank))
    lead = tuple(s if rng.random() < 0.5 else 1 for s in shape)
    pmf = rng.random(lead + (m,)).astype(np.float32) + 1e-3
    pmf /= pmf.sum(-1, keepdims=True)
    cdf = oracle.pmf_to_cdf(pmf, precision)
    data = rng.integers(0, m, size=shape).astype(np.int16)
    enc = tfc.range_encode(data, cdf, precision)
    assert enc == oracle.rc_encode(data, cdf, precision), (trial, shape, lead)
    dec = tfc.range_decode(enc, shape, cdf, precision)
    np.testing.assert_array_equal(dec.cpu().numpy(), data)


def test_bounded_batch_per_stream_tables(tfc, oracle, coder_lanes):
  """Batched bounded coder: 70 ragged streams, one shared table and one table per stream."""
  from compression_b200.python.ops import range_coding_ops as rc
  rng = np.random.default_rng(12)
  S, n, m, precision = 70, 333, 40, 13
  pmf = rng.random((S, 1, m)).astype(np.float32) + 1e-3
  pmf /= pmf.sum(-1, keepdims=True)
  cdfs = oracle.pmf_to_cdf(pmf, precision)                 # [S, 1, m + 1]
  data = rng.integers(0, m, size=(S, n)).astype(np.int16)
  for shared in (True, False):
    cdf = cdfs[0] if shared else cdfs
    out, nbytes = rc.range_encode_batch(data, cdf, precision, debug_level=0, shared_cdf=shared)
    got = rc.strings_to_host(out, nbytes)
    for s in range(S):
      assert got[s] == oracle.rc_encode(data[s], cdfs[0 if shared else s], precision), (shared, s)
    dec = rc.range_decode_batch(out, nbytes, (n,), cdf, precision, debug_level=0,
                                shared_cdf=shared)
    np.testing.assert_array_equal(dec.cpu().numpy(), data)


def test_bounded_readme_example(tfc, coder_lanes):
  """python/ops/range_coding_ops_test.py:33-46 / docs/range_coding.md:52-71."""
  rng = np.random.default_rng(5)
  data = rng.integers(0, 10, size=(4, 6, 6)).astype(np.int16)
  hist = np.bincount(data.ravel(), minlength=10).astype(np.float32)
  from compression_b200.python.ops import range_coding_ops as rc
  cdf = rc.pmf_to_quantized_cdf(hist / hist.sum(), 14)
  cdf = cdf.reshape(1, 1, 1, -1)
  enc = tfc.range_encode(data, cdf, 14)
  dec = tfc.range_decode(enc, data.shape, cdf, 14)
  np.testing.assert_array_equal(dec.cpu().numpy(), data)
  assert len(enc) < data.size  # < 8 bits per symbol


def test_debug_level_errors(tfc):
  """Error wording of unbounded_index_range_coding_kernels.cc:58-104 and
  range_coding_kernels.cc:153-176,250-255."""
  cdf = np.array([[0, 10, 65536, 0], [0, 5, 100, 65536]], np.int32)
  size = np.array([3, 4], np.int32)
  off = np.zeros(2, np.int32)
  data = np.zeros(4, np.int32)
  with pytest.raises(tfc.InvalidArgumentError, match=r"'index' has a value not in \[0, 2\)"):
    tfc.unbounded_index_range_encode(data, np.array([0, 1, 2, 0], np.int32), cdf, size, off,
                                     16, 4)
  idx = np.zeros(4, np.int32)
  with pytest.raises(tfc.InvalidArgumentError, match=r"'cdf_size' has a value not in \[3, 4\]"):
    tfc.unbounded_index_range_encode(data, idx, cdf, np.array([2, 4], np.int32), off, 16, 4)
  bad = cdf.copy(); bad[0, 2] = 65535
  with pytest.raises(tfc.InvalidArgumentError, match="Each cdf should start from 0 and end at"):
    tfc.unbounded_index_range_encode(data, idx, bad, size, off, 16, 4)
  bad = cdf.copy(); bad[1, 2] = 5
  with pytest.raises(tfc.InvalidArgumentError, match="CDF is not monotonic"):
    tfc.unbounded_index_range_decode(b"\0\0", idx, bad, size, off, 16, 4)
  with pytest.raises(tfc.InvalidArgumentError, match="should match the number of rows"):
    tfc.unbounded_index_range_encode(data, idx, cdf, size[:1], off, 16, 4)
  # debug_level=0 skips the value checks
  tfc.unbounded_index_range_encode(data, idx, cdf, size, off, 16, 4, debug_level=0)
  # bounded op
  d16 = np.array([0, 1, 2, 3], np.int16)
  c = np.array([[0, 100, 1000, 65536]], np.int32)
  with pytest.raises(tfc.InvalidArgumentError, match=r"value not in \[0, 3\)"):
    tfc.range_encode(d16, c, 16)
  with pytest.raises(tfc.InvalidArgumentError, match="CDF should start from 0 and end at"):
    tfc.range_encode(d16[:3], np.array([[0, 100, 1000, 65535]], np.int32), 16)
  with pytest.raises(tfc.InvalidArgumentError, match="CDF size should be > 2"):
    tfc.range_encode(np.zeros(3, np.int16), np.array([[0, 65536]], np.int32), 16)
  with pytest.raises(tfc.InvalidArgumentError, match="Cannot broadcast shape"):
    tfc.range_encode(np.zeros((2, 3), np.int16), np.zeros((2, 4, 5), np.int32), 16)


def test_pmf_to_quantized_cdf_matches_checker(tfc, oracle):
  """Bit-exact against the compiled PerShard, including tie-heavy symmetric rows and rows long
  enough for every std::sort code path; plus the reference's own property test
  (range_coding_helper_kernels_test.cc:70-142)."""
  rng = np.random.default_rng(6)
  for n in (2, 3, 16, 17, 22, 64, 257, 1481):
    rows = []
    k = np.arange(n, dtype=np.float64) - (n - 1) / 2
    for sigma in (0.3, 2.0, n / 6.0 + 0.1):
      p = np.exp(-0.5 * (k / sigma) ** 2)
      rows.append(p / p.sum())
    r = rng.random((13, n)) + 1e-4
    rows.extend(r / r.sum(-1, keepdims=True))
    pmf = np.array(rows, np.float32)
    for precision, scale in ((16, 1.0), (12, 1.3), (10, 0.8)):
      if n + 1 > (1 << precision):
        continue
      got = tfc.pmf_to_quantized_cdf(pmf * scale, precision).cpu().numpy()
      want = oracle.pmf_to_cdf(pmf * np.float32(scale), precision)
      np.testing.assert_array_equal(got, want)
      assert (got[:, 0] == 0).all() and (got[:, -1] == 1 << precision).all()
      assert (np.diff(got, axis=-1) > 0).all()


def test_pmf_long_rows_warp_sort(tfc, oracle):
  """Rows long enough for the warp-cooperative replay of std::sort (ranges above 64 items are
  partitioned by the whole warp, csrc/pmf_to_cdf.cu warp_sort) and for the ring re-insertion:
  few distinct masses (almost every comparison is a tie), uniform rows, sorted and reversed
  masses, and totals far above / below 2^precision so that hundreds of adjustments are replayed."""
  rng = np.random.default_rng(61)
  for n in (65, 129, 700, 1480, 3000):
    rows = [np.full(n, 1.0 / n)]
    for distinct in (2, 5, 40):
      r = rng.integers(1, distinct + 1, n).astype(np.float64)
      rows.append(r / r.sum())
    r = np.sort(rng.random(n) + 1e-3)
    rows.append(r / r.sum())
    rows.append(r[::-1] / r.sum())
    k = np.arange(n, dtype=np.float64) - (n - 1) / 2
    for sigma in (1.5, n / 10.0):
      p = np.exp(-0.5 * (k / sigma) ** 2) + 1e-9
      rows.append(p / p.sum())
    pmf = np.array(rows, np.float32)
    for precision, scale in ((16, 1.0), (16, 1.02), (16, 0.97), (13, 1.0)):
      if 2 * n > (1 << precision):
        continue
      got = tfc.pmf_to_quantized_cdf(pmf * scale, precision).cpu().numpy()
      want = oracle.pmf_to_cdf(pmf * np.float32(scale), precision)
      np.testing.assert_array_equal(got, want)


def test_pmf_gaussian_model_table(tfc, oracle):
  """The 64-level bmshj2018 scale table (bmshj2018.py:46-48, entropy_models.py:827-858): the
  rows on which a stable sort is known to differ from std::sort (SURVEY probe 4)."""
  import scipy.stats
  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))
  mult = -scipy.stats.norm.ppf(2.0 ** -8 / 2)
  center = np.ceil(table * mult).astype(int)
  for r in (0, 20, 40, 63):
    n = 2 * center[r] + 1
    s = np.abs(np.arange(n) - center[r]).astype(np.float32)
    sc = np.float32(table[r])
    up = 0.5 * scipy.special.erfc(-(2 ** -0.5) * ((np.float32(.5) - s) / sc).astype(np.float32))
    lo = 0.5 * scipy.special.erfc(-(2 ** -0.5) * ((np.float32(-.5) - s) / sc).astype(np.float32))
    pmf = np.concatenate([(up - lo), [2 * lo[0]]]).astype(np.float32)
    got = tfc.pmf_to_quantized_cdf(pmf, 16).cpu().numpy()
    np.testing.assert_array_equal(got, oracle.pmf_to_cdf(pmf[None], 16)[0])


def test_decoder_clamps_oversize_nbytes(tfc, oracle, coder_lanes):
  """an encode that overflowed its slab reports the size it NEEDED (> cap).  A decoder
  handed that (slab, nbytes) pair must stay inside the slab row and flag the stream corrupt —
  here the slab is the very end of an allocation guarded by a canary row."""
  from compression_b200.python.ops import range_coding_ops as rc
  rng = np.random.default_rng(17)
  cdf, size, offset = _random_tables(rng, 4, 30, 16)
  S, n, cap = 5, 400, 64
  index = rng.integers(0, 4, size=(S, n)).astype(np.int32)
  data = (rng.integers(-3, 20, size=(S, n)) + offset[index]).astype(np.int32)
  with rc.deferred_checks() as checks:
    checks.tag = "enc"
    out, nbytes = rc.unbounded_index_range_encode_batch(data, index, cdf, size, offset, 16, 4,
                                                        debug_level=0, cap=cap)
    assert int(nbytes.max()) > cap                       # the need, not what was written
    with pytest.raises(tfc.CudaError if hasattr(tfc, "CudaError") else Exception):
      checks.verify("enc")
    checks.tag = "dec"
    dec = rc.unbounded_index_range_decode_batch(out, nbytes, index, cdf, size, offset, 16, 4, 0)
    with pytest.raises(rc.InvalidArgumentError, match="corrupt"):
      checks.verify("dec")
  assert tuple(dec.shape) == (S, n)
  # outside deferred mode the same call raises by itself
  with pytest.raises(rc.InvalidArgumentError, match="corrupt"):
    rc.unbounded_index_range_decode_batch(out, nbytes, index, cdf, size, offset, 16, 4, 0)
  # bounded decoder: status is looked at too
  cdf1 = np.array([[0, 20000, 50000, 65536]], np.int32)
  d16 = rng.integers(0, 3, size=(3, 200)).astype(np.int16)
  o2, n2 = rc.range_encode_batch(d16, cdf1, 16, debug_level=0)
  back = rc.range_decode_batch(o2, n2, (200,), cdf1, 16, debug_level=0)
  np.testing.assert_array_equal(back.cpu().numpy(), d16)
  with pytest.raises(rc.InvalidArgumentError, match="corrupt"):
    rc.range_decode_batch(o2[:, :16].contiguous(), n2, (200,), cdf1, 16, debug_level=0)


@pytest.mark.parametrize("smem_kb,table_kb", [(54, 24), (54, 8), (32, 8)])
def test_compact_coder_ctas_code_the_same_bytes(tfc, oracle, smem_kb, table_kb):
  """TFC_CFG_CODER_SMEM_KB / TFC_CFG_CODER_TABLE_KB: small coder CTAs (shorter rounds, long
  table rows read through L1 instead of staged) produce the reference's bytes and decode them."""
  from compression_b200.python.layers import entropy_models
  from compression_b200.python.ops import range_coding_ops as rc
  dev = torch.device("cuda", 0)
  table = np.exp(np.linspace(np.log(0.11), np.log(256), 64))
  gc = entropy_models.GaussianConditional(None, table)
  gc.build_tables(dev)
  cdf, size, off = gc._quantized_cdf, gc._cdf_length, gc._offset
  rng = np.random.default_rng(smem_kb + table_kb)
  S, n = 6, 20011
  idx = rng.integers(0, 64, size=(S, n)).astype(np.int32)
  idx[:3] = np.minimum(rng.geometric(0.6, size=(3, n)) - 1, 12)
  sym = np.floor(rng.standard_normal((S, n)) * table[idx] * 1.3 + .5).astype(np.int32)
  tfc.configure(coder_lanes=1, coder_smem_kb=smem_kb, coder_table_kb=table_kb)
  try:
    out, nbytes = rc.unbounded_index_range_encode_batch(sym, idx, cdf, size, off, 16, 4, 0)
    dec = rc.unbounded_index_range_decode_batch(out, nbytes, idx, cdf, size, off, 16, 4, 0)
  finally:
    tfc.configure(coder_lanes=0, coder_smem_kb=0, coder_table_kb=0)
  got = rc.strings_to_host(out, nbytes)
  c, s_, o = cdf.cpu().numpy(), size.cpu().numpy(), off.cpu().numpy()
  for s in range(S):
    assert got[s] == oracle.ui_encode(sym[s], idx[s], c, s_, o, 16, 4), s
  np.testing.assert_array_equal(dec.cpu().numpy(), sym)
