"""TEST INFRASTRUCTURE ONLY — ctypes loader for the CPU checkers (see oracle/oracle_api.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (compression_b200/) never does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libtfc_ref.so")
REF_DBG_SO = os.path.join(_HERE, "_ref", "libtfc_ref_dbg.so")
PORT_SO = os.path.join(_HERE, "libtfc_port.so")

_i32p = ctypes.POINTER(ctypes.c_int32)
_i16p = ctypes.POINTER(ctypes.c_int16)
_i64p = ctypes.POINTER(ctypes.c_int64)
_u8p = ctypes.POINTER(ctypes.c_uint8)
_f32p = ctypes.POINTER(ctypes.c_float)


def build(reference=True, port=True):
  """Compile the checkers (port always; reference only where /root/reference exists)."""
  if port:
    subprocess.check_call(["make", "-s", "-C", _HERE, "port"])
  if reference and os.path.isdir("/root/reference"):
    subprocess.check_call(["make", "-s", "-C", _HERE, "reference"])


def _ptr(a, t):
  return a.ctypes.data_as(t) if a is not None else None


class Oracle(object):
  """One loaded checker library."""

  def __init__(self, path):
    self.path = path
    self.lib = ctypes.CDLL(path, mode=ctypes.RTLD_LOCAL)
    L = self.lib
    L.orc_kind.restype = ctypes.c_char_p
    L.orc_ui_encode.restype = ctypes.c_int64
    L.orc_ui_encode.argtypes = [_i32p, _i32p, ctypes.c_int64, _i32p, ctypes.c_int32,
                                ctypes.c_int32, _i32p, _i32p, ctypes.c_int, ctypes.c_int,
                                _u8p, ctypes.c_int64]
    L.orc_ui_decode.restype = ctypes.c_int
    L.orc_ui_decode.argtypes = [_u8p, ctypes.c_int64, _i32p, ctypes.c_int64, _i32p,
                                ctypes.c_int32, ctypes.c_int32, _i32p, _i32p, ctypes.c_int,
                                ctypes.c_int, _i32p]
    L.orc_rc_encode.restype = ctypes.c_int64
    L.orc_rc_encode.argtypes = [_i16p, _i64p, ctypes.c_int64, _i32p, ctypes.c_int32,
                                ctypes.c_int, _u8p, ctypes.c_int64]
    L.orc_rc_decode.restype = ctypes.c_int
    L.orc_rc_decode.argtypes = [_u8p, ctypes.c_int64, _i64p, ctypes.c_int64, _i32p,
                                ctypes.c_int32, ctypes.c_int, _i16p]
    L.orc_pmf_to_cdf.restype = None
    L.orc_pmf_to_cdf.argtypes = [_f32p, ctypes.c_int64, ctypes.c_int32, ctypes.c_int, _i32p,
                                 ctypes.c_int]
    L.orc_ui_encode_batch.restype = None
    L.orc_ui_encode_batch.argtypes = [ctypes.c_int64, _i32p, _i32p, ctypes.c_int64,
                                      ctypes.c_int64, _i32p, ctypes.c_int32, ctypes.c_int32,
                                      _i32p, _i32p, ctypes.c_int, ctypes.c_int, _u8p,
                                      ctypes.c_int64, _i64p, ctypes.c_int]
    L.orc_ui_decode_batch.restype = None
    L.orc_ui_decode_batch.argtypes = [ctypes.c_int64, _u8p, ctypes.c_int64, _i64p, _i32p,
                                      ctypes.c_int64, ctypes.c_int64, _i32p, ctypes.c_int32,
                                      ctypes.c_int32, _i32p, _i32p, ctypes.c_int,
                                      ctypes.c_int, _i32p, ctypes.c_int]

  @property
  def kind(self):
    return self.lib.orc_kind().decode()

  # -- single-stream forms -------------------------------------------------------------
  def ui_encode(self, data, index, cdf, cdf_size, offset, precision=16, overflow_width=4):
    data = np.ascontiguousarray(data, np.int32).ravel()
    index = np.ascontiguousarray(np.broadcast_to(index, data.shape) if np.ndim(index) == 0
                                 else index, np.int32).ravel()
    assert index.size == data.size
    cdf = np.ascontiguousarray(cdf, np.int32)
    cdf_size = np.ascontiguousarray(cdf_size, np.int32)
    offset = np.ascontiguousarray(offset, np.int32)
    cap = 8 * data.size + 64
    out = np.empty(cap, np.uint8)
    n = self.lib.orc_ui_encode(_ptr(data, _i32p), _ptr(index, _i32p), data.size,
                               _ptr(cdf, _i32p), cdf.shape[0], cdf.shape[1],
                               _ptr(cdf_size, _i32p), _ptr(offset, _i32p), precision,
                               overflow_width, _ptr(out, _u8p), cap)
    assert n >= 0
    return out[:n].tobytes()

  def ui_decode(self, encoded, index, cdf, cdf_size, offset, precision=16, overflow_width=4):
    index = np.ascontiguousarray(index, np.int32)
    cdf = np.ascontiguousarray(cdf, np.int32)
    cdf_size = np.ascontiguousarray(cdf_size, np.int32)
    offset = np.ascontiguousarray(offset, np.int32)
    enc = np.frombuffer(bytes(encoded) + b"\0", np.uint8).copy()
    out = np.empty(index.shape, np.int32)
    self.lib.orc_ui_decode(_ptr(enc, _u8p), len(encoded), _ptr(index, _i32p), index.size,
                           _ptr(cdf, _i32p), cdf.shape[0], cdf.shape[1],
                           _ptr(cdf_size, _i32p), _ptr(offset, _i32p), precision,
                           overflow_width, _ptr(out, _i32p))
    return out

  @staticmethod
  def _rows(data_shape, cdf):
    """Row of cdf.reshape(-1, chip) used by each element of data (numpy broadcasting)."""
    lead = cdf.shape[:-1]
    nrows = int(np.prod(lead)) if lead else 1
    ids = np.arange(nrows, dtype=np.int64).reshape(lead)
    return np.ascontiguousarray(np.broadcast_to(ids, data_shape)).ravel()

  def rc_encode(self, data, cdf, precision):
    data = np.ascontiguousarray(data, np.int16)
    cdf = np.ascontiguousarray(cdf, np.int32)
    rows = self._rows(data.shape, cdf)
    cap = 4 * data.size + 64
    out = np.empty(cap, np.uint8)
    n = self.lib.orc_rc_encode(_ptr(data, _i16p), _ptr(rows, _i64p), data.size,
                               _ptr(cdf, _i32p), cdf.shape[-1], precision, _ptr(out, _u8p), cap)
    assert n >= 0
    return out[:n].tobytes()

  def rc_decode(self, encoded, shape, cdf, precision):
    cdf = np.ascontiguousarray(cdf, np.int32)
    shape = tuple(int(s) for s in shape)
    rows = self._rows(shape, cdf)
    enc = np.frombuffer(bytes(encoded) + b"\0", np.uint8).copy()
    out = np.empty(shape, np.int16)
    self.lib.orc_rc_decode(_ptr(enc, _u8p), len(encoded), _ptr(rows, _i64p), out.size,
                           _ptr(cdf, _i32p), cdf.shape[-1], precision, _ptr(out, _i16p))
    return out

  def pmf_to_cdf(self, pmf, precision=16, threads=1):
    pmf = np.ascontiguousarray(pmf, np.float32)
    n = pmf.shape[-1]
    rows = pmf.size // n
    out = np.empty(pmf.shape[:-1] + (n + 1,), np.int32)
    self.lib.orc_pmf_to_cdf(_ptr(pmf, _f32p), rows, n, precision, _ptr(out, _i32p), threads)
    return out

  # -- batched forms (CPU baseline) ----------------------------------------------------
  def ui_encode_batch(self, data, index, cdf, cdf_size, offset, precision=16,
                      overflow_width=4, threads=1, cap=None):
    data = np.ascontiguousarray(data, np.int32)
    S, n = data.shape[0], data[0].size
    index = np.ascontiguousarray(index, np.int32)
    stride = n if index.size == data.size else 0
    assert index.size in (n, data.size)
    cdf = np.ascontiguousarray(cdf, np.int32)
    cdf_size = np.ascontiguousarray(cdf_size, np.int32)
    offset = np.ascontiguousarray(offset, np.int32)
    cap = cap or (4 * n + 64)
    out = np.empty((S, cap), np.uint8)
    nbytes = np.empty(S, np.int64)
    self.lib.orc_ui_encode_batch(S, _ptr(data, _i32p), _ptr(index, _i32p), stride, n,
                                 _ptr(cdf, _i32p), cdf.shape[0], cdf.shape[1],
                                 _ptr(cdf_size, _i32p), _ptr(offset, _i32p), precision,
                                 overflow_width, _ptr(out, _u8p), cap, _ptr(nbytes, _i64p),
                                 threads)
    return out, nbytes

  def ui_decode_batch(self, enc, nbytes, index, n, cdf, cdf_size, offset, precision=16,
                      overflow_width=4, threads=1):
    enc = np.ascontiguousarray(enc, np.uint8)
    S, cap = enc.shape
    nbytes = np.ascontiguousarray(nbytes, np.int64)
    index = np.ascontiguousarray(index, np.int32)
    stride = n if index.size == S * n and S > 1 else (0 if index.size == n else n)
    cdf = np.ascontiguousarray(cdf, np.int32)
    cdf_size = np.ascontiguousarray(cdf_size, np.int32)
    offset = np.ascontiguousarray(offset, np.int32)
    out = np.empty((S, n), np.int32)
    self.lib.orc_ui_decode_batch(S, _ptr(enc, _u8p), cap, _ptr(nbytes, _i64p),
                                 _ptr(index, _i32p), stride, n, _ptr(cdf, _i32p),
                                 cdf.shape[0], cdf.shape[1], _ptr(cdf_size, _i32p),
                                 _ptr(offset, _i32p), precision, overflow_width,
                                 _ptr(out, _i32p), threads)
    return out


_cache = {}


def load(kind="best"):
  """kind: 'reference' (oracle/_ref), 'reference_dbg', 'port', or 'best' (reference if built)."""
  if kind == "best":
    kind = "reference" if os.path.exists(REF_SO) else "port"
  if kind not in _cache:
    path = {"reference": REF_SO, "reference_dbg": REF_DBG_SO, "port": PORT_SO}[kind]
    if not os.path.exists(path):
      build(reference=kind.startswith("reference"), port=(kind == "port"))
    _cache[kind] = Oracle(path)
  return _cache[kind]


def available():
  kinds = []
  if os.path.exists(REF_SO) or os.path.isdir("/root/reference"):
    kinds.append("reference")
  kinds.append("port")
  return kinds
