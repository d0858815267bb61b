"""tfc_pmf_to_quantized_cdf_static: the allocation-free variant of PmfToQuantizedCdf (scratch in a
file-scope device array; no counterpart in the reference).
Called through the C ABI and compared, entry for entry, with the existing
tfc_pmf_to_quantized_cdf on two calls in a row (the second finds the first call's contents in the
shared scratch).  Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _pmfs(rows, n, seed, kind):
  g = torch.Generator(device="cuda").manual_seed(seed)
  if kind == "uniform":
    p = torch.rand((rows, n), generator=g, device="cuda") + 1e-6
  else:                                   # symmetric: exact ties, the std::sort replay matters
    sig = torch.rand((rows, 1), generator=g, device="cuda") * (n / 6.0) + 0.2
    k = torch.arange(n, device="cuda", dtype=torch.float32) - (n - 1) / 2
    p = torch.exp(-0.5 * (k / sig) ** 2)
  return (p / p.sum(-1, keepdim=True)).contiguous()


def _static_call(L, _lib, pmf, precision, out):
  rows, n = pmf.shape
  return L.tfc_pmf_to_quantized_cdf_static(
      _lib.ptr(pmf), ctypes.c_int64(rows), ctypes.c_int32(n), int(precision), _lib.ptr(out),
      _lib.stream_ptr(pmf.device))


def test_static_scratch_variant_matches_existing_entry_point():
  from compression_b200 import _lib
  from compression_b200.python.ops import range_coding_ops as rc
  L = _lib.lib()
  L.tfc_pmf_to_quantized_cdf_static_capacity.restype = ctypes.c_size_t
  cap = int(L.tfc_pmf_to_quantized_cdf_static_capacity())
  assert cap == 8 << 20
  # two calls in a row; the second is smaller in both dimensions, so the scratch beyond and
  # between its items still holds the first call's sort items
  cases = [(3000, 40, 16, "gaussian"), (300, 700, 16, "uniform"), (77, 23, 12, "gaussian")]
  inputs = [_pmfs(r, n, 100 + i, kind) for i, (r, n, _, kind) in enumerate(cases)]
  outs = [torch.full((r, n + 1), -7, dtype=torch.int32, device="cuda") for r, n, _, _ in cases]
  want = [rc.pmf_to_quantized_cdf(p, prec) for p, (_, _, prec, _) in zip(inputs, cases)]
  torch.cuda.synchronize()
  for rep in range(2):
    for p, o, w, (r, n, prec, _) in zip(inputs, outs, want, cases):
      o.fill_(-7)
      free0 = torch.cuda.mem_get_info()[0]
      _lib.check(_static_call(L, _lib, p, prec, o))
      torch.cuda.synchronize()
      assert torch.equal(o, w), (rep, r, n)                  # integer op: exact equality
      assert bool((o[:, 0] == 0).all()) and bool((o[:, -1] == (1 << prec)).all())
      if rep == 1:                                           # set-up done: no device allocation
        assert torch.cuda.mem_get_info()[0] >= free0
  # beyond the maximum: a clear error, nothing launched, output untouched
  rows, n = 8192, 65                                         # 8192 * 65 items > 524 288
  big = _pmfs(rows, n, 7, "uniform")
  out = torch.full((rows, n + 1), -7, dtype=torch.int32, device="cuda")
  with pytest.raises(_lib.InvalidArgumentError, match="static scratch array holds"):
    _lib.check(_static_call(L, _lib, big, 16, out))
  torch.cuda.synchronize()
  assert bool((out == -7).all())
  # exactly at the maximum works
  rows, n = 8192, 64
  edge = _pmfs(rows, n, 8, "gaussian")
  out = torch.empty((rows, n + 1), dtype=torch.int32, device="cuda")
  _lib.check(_static_call(L, _lib, edge, 16, out))
  assert torch.equal(out, rc.pmf_to_quantized_cdf(edge, 16))
  with pytest.raises(_lib.InvalidArgumentError, match="precision"):
    _lib.check(_static_call(L, _lib, edge, 17, out))
