"""`.tfci` container — mirror of `tfc.PackedTensors`
(`/root/reference/tensorflow_compression/python/util/packed_tensors.py:30-105`).

The reference stores the compressed strings and the shape vectors in a `tf.train.Example`
protocol buffer: feature `chr(1)`, `chr(2)`, ... for the tensors in call order (strings as
`bytes_list`, integer vectors as `int64_list`) and an optional `"MD"` bytes feature with a model
identifier.  TensorFlow is not needed to emit or read those bytes: this module writes the
protobuf wire format directly (tensorflow/core/example/{example,feature}.proto):

  Example  { Features features = 1; }
  Features { map<string, Feature> feature = 1; }      map entry: key = 1, value = 2
  Feature  { oneof kind { BytesList bytes_list = 1; FloatList float_list = 2;
                          Int64List int64_list = 3; } }
  BytesList { repeated bytes value = 1; }
  Int64List { repeated int64 value = 1 [packed = true]; }

Map entries are written in sorted key order (protobuf's deterministic serialisation); any order
is accepted when parsing, as are unpacked int64 lists.
"""
import numpy as np

__all__ = ["PackedTensors"]

_LEN, _VARINT = 2, 0


def _varint(v):
  v &= (1 << 64) - 1               # negative int64 -> 10-byte two's complement varint
  out = bytearray()
  while True:
    b = v & 0x7F
    v >>= 7
    if v:
      out.append(b | 0x80)
    else:
      out.append(b)
      return bytes(out)


def _field(number, payload):
  return _varint((number << 3) | _LEN) + _varint(len(payload)) + payload


def _read_varint(buf, at):
  shift = v = 0
  while True:
    if at >= len(buf):
      raise ValueError("truncated varint")
    b = buf[at]
    at += 1
    v |= (b & 0x7F) << shift
    shift += 7
    if not b & 0x80:
      return v, at
    if shift > 70:
      raise ValueError("varint too long")


def _fields(buf):
  """Yields (field number, wire type, value) of one message; LEN fields as bytes."""
  at = 0
  while at < len(buf):
    key, at = _read_varint(buf, at)
    number, wire = key >> 3, key & 7
    if wire == _VARINT:
      v, at = _read_varint(buf, at)
    elif wire == _LEN:
      n, at = _read_varint(buf, at)
      if at + n > len(buf):
        raise ValueError("truncated field")
      v, at = bytes(buf[at:at + n]), at + n
    elif wire == 1:
      v, at = bytes(buf[at:at + 8]), at + 8
    elif wire == 5:
      v, at = bytes(buf[at:at + 4]), at + 4
    else:
      raise ValueError("unsupported wire type %d" % wire)
    yield number, wire, v


def _signed(v):
  return v - (1 << 64) if v >> 63 else v


def _dtype_of(spec):
  """dtype of an array-like spec; classes (np.int32, bytes, ...) and dtypes stand for themselves."""
  return spec if isinstance(spec, type) else getattr(spec, "dtype", spec)


def _is_string_spec(spec):
  if isinstance(spec, str):
    return spec in ("string", "bytes")
  if spec in (bytes, str, object, np.object_, np.bytes_):
    return True
  dt = _dtype_of(spec)
  try:
    return np.dtype(dt).kind in ("O", "S")
  except TypeError:
    return False


def _is_integer_spec(spec):
  if isinstance(spec, str):
    return spec.startswith(("int", "uint"))
  dt = _dtype_of(spec)
  if hasattr(dt, "is_floating_point"):                     # torch.dtype
    import torch
    return dt in (torch.int8, torch.int16, torch.int32, torch.int64, torch.uint8)
  try:
    return np.dtype(dt).kind in ("i", "u")
  except TypeError:
    return False


def _np_dtype(spec):
  if isinstance(spec, str):
    return np.dtype(spec)
  dt = _dtype_of(spec)
  if hasattr(dt, "is_floating_point"):
    import torch
    return {torch.int8: np.int8, torch.int16: np.int16, torch.int32: np.int32,
            torch.int64: np.int64, torch.uint8: np.uint8}[dt]
  return np.dtype(dt)


class PackedTensors(object):
  """Packed representation of compressed tensors (rank-1 integer or string vectors) with an
  optional model identifier.  `tensors` arguments only supply dtypes, as in the reference: a
  torch / numpy array or dtype, or the names "string" / "int32" / ...
  """

  def __init__(self, string=None):
    self._features = {}            # key -> ("bytes", [bytes...]) | ("int64", [int...])
    if string:
      self.string = string

  # ---- packed_tensors.py:45-58 ---------------------------------------------------------------
  @property
  def model(self):
    return self._features["MD"][1][0].decode("ascii")

  @model.setter
  def model(self, value):
    self._features["MD"] = ("bytes", [value.encode("ascii")])

  @model.deleter
  def model(self):
    self._features.pop("MD", None)

  # ---- packed_tensors.py:60-67 ---------------------------------------------------------------
  @property
  def string(self):
    entries = b""
    for key in sorted(self._features):
      kind, values = self._features[key]
      if kind == "bytes":
        feature = _field(1, b"".join(_field(1, v) for v in values))
      else:
        packed = b"".join(_varint(int(v)) for v in values)
        feature = _field(3, _field(1, packed) if values else b"")
      entries += _field(1, _field(1, key.encode("utf-8")) + _field(2, feature))
    return _field(1, entries)

  @string.setter
  def string(self, value):
    feats = {}
    for number, wire, features in _fields(value):
      if number != 1 or wire != _LEN:
        continue
      for n2, w2, entry in _fields(features):
        if n2 != 1 or w2 != _LEN:
          continue
        key, feature = "", b""
        for n3, w3, v in _fields(entry):
          if n3 == 1 and w3 == _LEN:
            key = v.decode("utf-8")
          elif n3 == 2 and w3 == _LEN:
            feature = v
        kind, values = "bytes", []
        for n4, w4, lst in _fields(feature):
          if w4 != _LEN:
            continue
          if n4 == 1:
            kind = "bytes"
            values = [v for n5, w5, v in _fields(lst) if n5 == 1 and w5 == _LEN]
          elif n4 == 3:
            kind, values = "int64", []
            for n5, w5, v in _fields(lst):
              if n5 != 1:
                continue
              if w5 == _LEN:                                 # packed
                at = 0
                while at < len(v):
                  x, at = _read_varint(v, at)
                  values.append(_signed(x))
              elif w5 == _VARINT:
                values.append(_signed(v))
        feats[key] = (kind, values)
    self._features = feats

  # ---- packed_tensors.py:69-91 ---------------------------------------------------------------
  def pack(self, tensors, arrays):
    if len(tensors) != len(arrays):
      raise ValueError("`tensors` and `arrays` must have same length.")
    i = 1
    for tensor, array in zip(tensors, arrays):
      array = np.asarray(array, dtype=object) if _is_string_spec(tensor) else np.asarray(array)
      if array.ndim != 1:
        raise RuntimeError("Unexpected tensor rank: {}.".format(array.ndim))
      if _is_integer_spec(tensor):
        self._features[chr(i)] = ("int64", [int(v) for v in array])
      elif _is_string_spec(tensor):
        self._features[chr(i)] = ("bytes", [bytes(v) for v in array])
      else:
        raise RuntimeError("Unexpected tensor dtype: '{}'.".format(_dtype_of(tensor)))
      i += 1
    while chr(i) in self._features:          # delete any remaining, previously set arrays
      del self._features[chr(i)]
      i += 1

  # ---- packed_tensors.py:93-105 --------------------------------------------------------------
  def unpack(self, tensors):
    arrays = []
    for i, tensor in enumerate(tensors):
      kind, values = self._features.get(chr(i + 1), ("bytes", []))
      if _is_integer_spec(tensor):
        arrays.append(np.array(values if kind == "int64" else [], dtype=_np_dtype(tensor)))
      elif _is_string_spec(tensor):
        arrays.append(np.array(values if kind == "bytes" else [], dtype=object))
      else:
        raise RuntimeError("Unexpected tensor dtype: '{}'.".format(_dtype_of(tensor)))
    return arrays
