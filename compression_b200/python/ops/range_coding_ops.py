"""Range coding ops — same names and argument meaning as the reference's
`tensorflow_compression/python/ops/range_coding_ops.py:26-31` (which exports the ops
registered in `cc/ops/range_coding_ops.cc`), over torch CUDA tensors and the C ABI.

Single-stream forms mirror the TF ops one to one and return `bytes` / tensors:
  range_encode, range_decode, unbounded_index_range_encode, unbounded_index_range_decode,
  pmf_to_quantized_cdf.
`*_batch` forms code S independent streams in one launch (what `tf.map_fn` over the batch does
in `python/layers/entropy_models.py:242-244,282-283`) and keep the strings on the device.
"""
import ctypes

import numpy as np
import torch

from compression_b200 import _lib
from compression_b200._lib import InvalidArgumentError  # re-export  pylint:disable=unused-import

__all__ = [
    "range_encode", "range_decode", "unbounded_index_range_encode",
    "unbounded_index_range_decode", "pmf_to_quantized_cdf",
    "unbounded_index_range_encode_batch", "unbounded_index_range_decode_batch",
    "range_encode_batch", "range_decode_batch", "strings_to_host", "strings_to_device",
    "deferred_checks",
]

_i64 = ctypes.c_int64


class deferred_checks(object):
  """Context manager: inside it the batched coder calls do not synchronise with the host to
  look at their per-stream status words (capacity overflow on encode, corrupt string on
  decode).  Each call's status words are copied to a small pinned host buffer on the stream that
  produced them (one 4*S-byte copy and an event, no reduction kernel) and filed under the
  current `tag`; the owner of the block verifies a tag when its results are about to be used —
  `verify(tag)` waits for that call's event only — and everything still pending is verified on
  exit.  This is what lets several batches be in flight on different CUDA streams while no
  result is delivered, and no decode queued, before the status of what it depends on has been
  looked at.  An encode that overflowed its slab cannot be retried transparently in this mode:
  verify raises and the caller re-runs the batch with a larger `cap`.  Per thread."""
  _tls = __import__("threading").local()

  def __enter__(self):
    self._prev = getattr(self._tls, "pending", None)
    self._tls.pending = self
    self._by_tag = {}        # tag -> [(what, pinned host words, count, event)]
    self._order = []         # tags in the order they first appeared
    self._pool = []          # pinned int32 buffers ready for reuse
    self.tag = None
    return self

  def __exit__(self, exc_type, exc, tb):
    self._tls.pending = self._prev
    if exc_type is None:
      self.flush()
    return False

  def add(self, what, status):
    n = int(status.numel())
    host = None
    for i, buf in enumerate(self._pool):
      if buf.numel() >= n:
        host = self._pool.pop(i)
        break
    if host is None:
      host = torch.empty((max(n, 256),), dtype=torch.int32, pin_memory=True)
    host[:n].copy_(status.reshape(-1), non_blocking=True)
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(status.device))
    if self.tag not in self._by_tag:
      self._by_tag[self.tag] = []
      self._order.append(self.tag)
    self._by_tag[self.tag].append((what, host, n, ev))

  def verify(self, tag):
    """Waits for the calls filed under `tag` and raises if any of their streams failed."""
    entries = self._by_tag.pop(tag, None)
    if entries is None:
      return
    self._order.remove(tag)
    bad = None
    for what, host, n, ev in entries:
      ev.synchronize()
      if bad is None and n and bool(host[:n].numpy().any()):
        bad = what
      self._pool.append(host)
    if bad == "encode":
      raise _lib.CudaError("range encode output slab too small (deferred check): re-run "
                           "the batch with a larger `cap`")
    if bad is not None:
      raise InvalidArgumentError(
          "corrupt range-coded string (decoder ran past the end of a CDF)")

  def drain(self, keep):
    """Verifies the oldest tags until at most `keep` are pending (bounds the backlog of a long
    pipeline; the calls verified are `keep` tags old and have long finished)."""
    while len(self._order) > max(0, int(keep)):
      self.verify(self._order[0])

  def flush(self):
    self.drain(0)

  @classmethod
  def collector(cls):
    return getattr(cls._tls, "pending", None)


def _shape_arr(shape):
  return (ctypes.c_int64 * max(len(shape), 1))(*[int(s) for s in shape])


def _tables(cdf, cdf_size, offset, device):
  cdf = _lib.to_device(cdf, torch.int32, device)
  if cdf.dim() != 2:
    raise InvalidArgumentError(
        "'cdf' should be 2-D and cdf.dim_size(1) >= 3: %s" % (list(cdf.shape),))
  cdf_size = _lib.to_device(cdf_size, torch.int32, cdf.device)
  offset = _lib.to_device(offset, torch.int32, cdf.device)
  if cdf_size.dim() != 1 or cdf_size.shape[0] != cdf.shape[0]:
    raise InvalidArgumentError(
        "'cdf_size' should be 1-D and its length should match the number of rows in 'cdf': %s"
        % (list(cdf_size.shape),))
  if offset.dim() != 1 or offset.shape[0] != cdf.shape[0]:
    raise InvalidArgumentError(
        "'offset' should be 1-D and its length should match the number of rows in 'cdf': "
        "offset.shape=%s, cdf.shape=%s" % (list(offset.shape), list(cdf.shape)))
  return cdf, cdf_size, offset


def default_capacity(n):
  """Output slab size per stream used when the caller gives none: 2 bytes per symbol covers
  any stream without escapes at precision 16; the kernel reports the exact need otherwise."""
  return int(2 * n + 64 + (-(2 * n + 64)) % 16)


# ---------------------------------------------------------------------------------------
# batched forms (device in, device out)
# ---------------------------------------------------------------------------------------
def unbounded_index_range_encode_batch(data, index, cdf, cdf_size, offset, precision,
                                       overflow_width, debug_level=1, cap=None):
  """data [S, ...] int32, index same shape or data.shape[1:] (shared by all streams).

  Returns (out [S, cap] uint8, nbytes [S] int32), both on the device.  If a stream does not
  fit in `cap` the call is retried once with the exact size the kernel reported.
  """
  data = _lib.to_device(data, torch.int32)
  dev = data.device
  index = _lib.to_device(index, torch.int32, dev)
  cdf, cdf_size, offset = _tables(cdf, cdf_size, offset, dev)
  S = data.shape[0]
  n = data[0].numel() if S else 0
  if index.shape == data.shape:
    stride = n
  elif tuple(index.shape) == tuple(data.shape[1:]):
    stride = 0
  else:
    raise InvalidArgumentError(
        "`data` and `index` should have the same shape: data.shape=%s, index.shape=%s"
        % (list(data.shape), list(index.shape)))
  cap = int(cap) if cap else default_capacity(n)
  L = _lib.lib()
  for attempt in range(2):
    out = torch.empty((S, cap), dtype=torch.uint8, device=dev)
    nbytes = torch.empty((S,), dtype=torch.int32, device=dev)
    status = torch.empty((S,), dtype=torch.int32, device=dev)
    _lib.check(L.tfc_unbounded_index_range_encode(
        _lib.ptr(data), _lib.ptr(index), _i64(stride), _i64(S), _i64(n), _lib.ptr(cdf),
        ctypes.c_int32(cdf.shape[0]), ctypes.c_int32(cdf.shape[1]), _lib.ptr(cdf_size),
        _lib.ptr(offset), int(precision), int(overflow_width), int(debug_level), _lib.ptr(out),
        _i64(cap), _lib.ptr(nbytes), _lib.ptr(status), _lib.stream_ptr(dev)))
    pending = deferred_checks.collector()
    if pending is not None:
      pending.add("encode", status)
      break
    if attempt == 0 and S and bool((status != 0).any()):
      cap = int(nbytes.max().item())
      cap += (-cap) % 16
      continue
    break
  return out, nbytes


def unbounded_index_range_decode_batch(enc, nbytes, index, cdf, cdf_size, offset, precision,
                                       overflow_width, debug_level=1, shared_index=False,
                                       check=True):
  """enc [S, cap] uint8 + nbytes [S]; index [S, ...] or [...] with shared_index=True.
  Returns int32 [S, ...]."""
  enc = _lib.to_device(enc, torch.uint8)
  dev = enc.device
  nbytes = _lib.to_device(nbytes, torch.int32, dev)
  index = _lib.to_device(index, torch.int32, dev)
  cdf, cdf_size, offset = _tables(cdf, cdf_size, offset, dev)
  S, cap = enc.shape
  if shared_index:
    n, stride, oshape = index.numel(), 0, (S,) + tuple(index.shape)
  else:
    if index.shape[0] != S:
      raise InvalidArgumentError("index batch %d != number of strings %d" % (index.shape[0], S))
    n = index[0].numel() if S else 0
    stride, oshape = n, tuple(index.shape)
  out = torch.empty(oshape, dtype=torch.int32, device=dev)
  status = torch.empty((S,), dtype=torch.int32, device=dev)
  _lib.check(_lib.lib().tfc_unbounded_index_range_decode(
      _lib.ptr(enc), _i64(cap), _lib.ptr(nbytes), _lib.ptr(index), _i64(stride), _i64(S),
      _i64(n), _lib.ptr(cdf), ctypes.c_int32(cdf.shape[0]), ctypes.c_int32(cdf.shape[1]),
      _lib.ptr(cdf_size), _lib.ptr(offset), int(precision), int(overflow_width),
      int(debug_level), _lib.ptr(out), _lib.ptr(status), _lib.stream_ptr(dev)))
  pending = deferred_checks.collector()
  if pending is not None:
    pending.add("decode", status)
  elif check and S and n and bool((status != 0).any()):
    raise InvalidArgumentError("corrupt range-coded string (decoder ran past the end of a CDF)")
  return out


def range_encode_batch(data, cdf, precision, debug_level=1, cap=None, shared_cdf=True):
  """data [S, *data_shape] int16; cdf [*cdf_shape] shared, or [S, *cdf_shape] per stream."""
  data = _lib.to_device(data, torch.int16)
  dev = data.device
  cdf = _lib.to_device(cdf, torch.int32, dev)
  S = data.shape[0]
  dshape = tuple(data.shape[1:])
  cshape = tuple(cdf.shape) if shared_cdf else tuple(cdf.shape[1:])
  cstride = 0 if shared_cdf else int(np.prod(cshape, dtype=np.int64))
  n = int(np.prod(dshape, dtype=np.int64))
  cap = int(cap) if cap else default_capacity(n)
  L = _lib.lib()
  for attempt in range(2):
    out = torch.empty((S, cap), dtype=torch.uint8, device=dev)
    nbytes = torch.empty((S,), dtype=torch.int32, device=dev)
    status = torch.empty((S,), dtype=torch.int32, device=dev)
    _lib.check(L.tfc_range_encode(
        _lib.ptr(data), _i64(S), _shape_arr(dshape), len(dshape), _lib.ptr(cdf),
        _shape_arr(cshape), len(cshape), _i64(cstride), int(precision), int(debug_level),
        _lib.ptr(out), _i64(cap), _lib.ptr(nbytes), _lib.ptr(status), _lib.stream_ptr(dev)))
    if attempt == 0 and S and bool((status != 0).any()):
      cap = int(nbytes.max().item())
      cap += (-cap) % 16
      continue
    break
  return out, nbytes


def range_decode_batch(enc, nbytes, shape, cdf, precision, debug_level=1, shared_cdf=True):
  enc = _lib.to_device(enc, torch.uint8)
  dev = enc.device
  nbytes = _lib.to_device(nbytes, torch.int32, dev)
  cdf = _lib.to_device(cdf, torch.int32, dev)
  S, cap = enc.shape
  dshape = tuple(int(s) for s in shape)
  cshape = tuple(cdf.shape) if shared_cdf else tuple(cdf.shape[1:])
  cstride = 0 if shared_cdf else int(np.prod(cshape, dtype=np.int64))
  out = torch.empty((S,) + dshape, dtype=torch.int16, device=dev)
  status = torch.empty((S,), dtype=torch.int32, device=dev)
  _lib.check(_lib.lib().tfc_range_decode(
      _lib.ptr(enc), _i64(cap), _lib.ptr(nbytes), _i64(S), _shape_arr(dshape), len(dshape),
      _lib.ptr(cdf), _shape_arr(cshape), len(cshape), _i64(cstride), int(precision),
      int(debug_level), _lib.ptr(out), _lib.ptr(status), _lib.stream_ptr(dev)))
  pending = deferred_checks.collector()
  if pending is not None:
    pending.add("decode", status)
  elif S and out.numel() and bool((status != 0).any()):
    raise InvalidArgumentError("corrupt range-coded string (decoder ran past the end of a CDF)")
  return out


def strings_to_host(out, nbytes):
  """Device slab + lengths -> list of Python bytes (one D2H copy of the used prefix)."""
  nb = nbytes.cpu().tolist()
  width = max(nb) if nb else 0
  host = out[:, :width].cpu().numpy() if width else np.zeros((len(nb), 0), np.uint8)
  return [host[i, :nb[i]].tobytes() for i in range(len(nb))]


def strings_to_device(strings, device=None):
  """List of bytes -> (slab [S, cap] uint8, nbytes [S] int32) on the device."""
  S = len(strings)
  cap = max([len(s) for s in strings] + [1])
  cap += (-cap) % 16
  host = np.zeros((S, cap), np.uint8)
  for i, s in enumerate(strings):
    host[i, :len(s)] = np.frombuffer(s, np.uint8)
  nb = np.array([len(s) for s in strings], np.int32)
  return (_lib.to_device(host, torch.uint8, device), _lib.to_device(nb, torch.int32, device))


# ---------------------------------------------------------------------------------------
# single-stream forms == the TF ops
# ---------------------------------------------------------------------------------------
def unbounded_index_range_encode(data, index, cdf, cdf_size, offset, precision, overflow_width,
                                 debug_level=1):
  """cc/ops/range_coding_ops.cc:125-135.  Returns the encoded string as `bytes`."""
  data = _lib.to_device(data, torch.int32)
  index = _lib.to_device(index, torch.int32, data.device)
  if data.shape != index.shape:
    raise InvalidArgumentError(
        "`data` and `index` should have the same shape: data.shape=%s, index.shape=%s"
        % (list(data.shape), list(index.shape)))
  out, nbytes = unbounded_index_range_encode_batch(
      data.reshape(1, -1), index.reshape(1, -1), cdf, cdf_size, offset, precision,
      overflow_width, debug_level)
  return strings_to_host(out, nbytes)[0]


def unbounded_index_range_decode(encoded, index, cdf, cdf_size, offset, precision,
                                 overflow_width, debug_level=1):
  """cc/ops/range_coding_ops.cc:203-216.  Returns int32 tensor of `index.shape`."""
  index = _lib.to_device(index, torch.int32)
  enc, nbytes = strings_to_device([bytes(encoded)], index.device)
  out = unbounded_index_range_decode_batch(
      enc, nbytes, index.reshape(1, -1), cdf, cdf_size, offset, precision, overflow_width,
      debug_level)
  return out.reshape(index.shape)


def range_encode(data, cdf, precision, debug_level=1):
  """cc/ops/range_coding_ops.cc:29-35.  data int16, cdf int32 with one more axis."""
  data = _lib.to_device(data, torch.int16)
  out, nbytes = range_encode_batch(data.unsqueeze(0), cdf, precision, debug_level)
  return strings_to_host(out, nbytes)[0]


def range_decode(encoded, shape, cdf, precision, debug_level=1):
  """cc/ops/range_coding_ops.cc:91-103."""
  cdf = _lib.to_device(cdf, torch.int32)
  if isinstance(shape, torch.Tensor):
    shape = shape.cpu().tolist()
  enc, nbytes = strings_to_device([bytes(encoded)], cdf.device)
  return range_decode_batch(enc, nbytes, list(shape), cdf, precision, debug_level)[0]


def pmf_to_quantized_cdf(pmf, precision):
  """cc/ops/range_coding_ops.cc:244-257: float32 [..., n] -> int32 [..., n+1]."""
  pmf = _lib.to_device(pmf, torch.float32)
  if pmf.dim() < 1:
    raise InvalidArgumentError("`pmf` should be at least 1-D.")
  n = pmf.shape[-1]
  if n <= 1:
    raise InvalidArgumentError("`pmf` size should be at least 2 in the last axis.")
  rows = pmf.numel() // n
  cdf = torch.empty(tuple(pmf.shape[:-1]) + (n + 1,), dtype=torch.int32, device=pmf.device)
  L = _lib.lib()
  ws_bytes = int(L.tfc_pmf_to_quantized_cdf_workspace(_i64(rows), ctypes.c_int32(n)))
  ws = torch.empty((max(ws_bytes, 16),), dtype=torch.uint8, device=pmf.device)
  _lib.check(L.tfc_pmf_to_quantized_cdf(
      _lib.ptr(pmf), _i64(rows), ctypes.c_int32(n), int(precision), _lib.ptr(cdf),
      _lib.ptr(ws), ctypes.c_size_t(ws_bytes), _lib.stream_ptr(pmf.device)))
  return cdf
