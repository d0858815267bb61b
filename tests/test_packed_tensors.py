"""`.tfci` container: the reference's own tests
(`python/util/packed_tensors_test.py:28-53`) re-expressed, plus byte equality of our hand-written
protobuf wire format with the protobuf runtime serialising the same `tf.train.Example` schema
(tensorflow/core/example/{example,feature}.proto, rebuilt here from a descriptor — no TF)."""
import numpy as np
import pytest

from compression_b200.python.util import packed_tensors


def test_pack_unpack():
  """packed_tensors_test.py:28-44."""
  string = np.array(["xyz".encode("ascii")], dtype=object)
  shape = np.array([1, 3], dtype=np.int32)
  tensors = ["string", np.int32]
  packed = packed_tensors.PackedTensors()
  packed.pack(tensors, [string, shape])
  packed = packed_tensors.PackedTensors(packed.string)
  string_u, shape_u = packed.unpack(tensors)
  assert list(string_u) == list(string)
  assert shape_u.dtype == np.int32 and list(shape_u) == [1, 3]


def test_model():
  """packed_tensors_test.py:46-52."""
  packed = packed_tensors.PackedTensors()
  packed.model = "xyz"
  packed = packed_tensors.PackedTensors(packed.string)
  assert packed.model == "xyz"
  del packed.model
  assert packed.string == packed_tensors.PackedTensors().string


def test_pack_errors_and_repack_drops_stale_entries():
  packed = packed_tensors.PackedTensors()
  with pytest.raises(ValueError, match="must have same length"):
    packed.pack(["string"], [])
  with pytest.raises(RuntimeError, match="Unexpected tensor rank: 2"):
    packed.pack([np.int32], [np.zeros((2, 2), np.int32)])
  with pytest.raises(RuntimeError, match="Unexpected tensor dtype"):
    packed.pack([np.float32], [np.zeros(2, np.float32)])
  packed.pack(["string", np.int32, np.int32], [[b"a"], [1], [2]])
  packed.pack(["string"], [[b"b"]])
  again = packed_tensors.PackedTensors(packed.string)
  assert list(again.unpack(["string", np.int32])[0]) == [b"b"]
  assert again.unpack(["string", np.int32])[1].size == 0


def _example_class():
  """tf.train.Example rebuilt from its schema with the protobuf runtime."""
  from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
  F = descriptor_pb2.FieldDescriptorProto
  fd = descriptor_pb2.FileDescriptorProto(name="tfc_example_test.proto", package="tensorflow",
                                          syntax="proto3")

  def msg(name):
    m = fd.message_type.add()
    m.name = name
    return m

  bl = msg("BytesList")
  bl.field.add(name="value", number=1, type=F.TYPE_BYTES, label=F.LABEL_REPEATED)
  fl = msg("FloatList")
  fl.field.add(name="value", number=1, type=F.TYPE_FLOAT, label=F.LABEL_REPEATED)
  il = msg("Int64List")
  il.field.add(name="value", number=1, type=F.TYPE_INT64, label=F.LABEL_REPEATED)
  ft = msg("Feature")
  ft.oneof_decl.add(name="kind")
  for name, num, t in (("bytes_list", 1, "BytesList"), ("float_list", 2, "FloatList"),
                       ("int64_list", 3, "Int64List")):
    ft.field.add(name=name, number=num, type=F.TYPE_MESSAGE, label=F.LABEL_OPTIONAL,
                 type_name=".tensorflow." + t, oneof_index=0)
  fs = msg("Features")
  entry = fs.nested_type.add()
  entry.name = "FeatureEntry"
  entry.options.map_entry = True
  entry.field.add(name="key", number=1, type=F.TYPE_STRING, label=F.LABEL_OPTIONAL)
  entry.field.add(name="value", number=2, type=F.TYPE_MESSAGE, label=F.LABEL_OPTIONAL,
                  type_name=".tensorflow.Feature")
  fs.field.add(name="feature", number=1, type=F.TYPE_MESSAGE, label=F.LABEL_REPEATED,
               type_name=".tensorflow.Features.FeatureEntry")
  ex = msg("Example")
  ex.field.add(name="features", number=1, type=F.TYPE_MESSAGE, label=F.LABEL_OPTIONAL,
               type_name=".tensorflow.Features")
  pool = descriptor_pool.DescriptorPool()
  pool.Add(fd)
  desc = pool.FindMessageTypeByName("tensorflow.Example")
  if hasattr(message_factory, "GetMessageClass"):
    return message_factory.GetMessageClass(desc)
  return message_factory.MessageFactory(pool).GetPrototype(desc)


def test_wire_format_matches_protobuf_runtime():
  pytest.importorskip("google.protobuf")
  Example = _example_class()
  rng = np.random.default_rng(0)
  strings = [bytes(rng.integers(0, 256, size=n, dtype=np.uint8)) for n in (0, 1, 127, 128, 70000)]
  cases = [
      (["string", "string", np.int32, np.int32, np.int32],
       [[strings[4]], [strings[2]], [512, 768], [32, 48], [8, 12]], "bmshj2018-factorized-mse-1"),
      (["string", np.int32, np.int32], [[strings[3]], [256, 256], [16, 16]], None),
      (["string", np.int64], [strings, [0, -1, 2 ** 40, -2 ** 62, 127, 128]], "m"),
      ([np.int32], [np.zeros((0,), np.int32)], None),
  ]
  for tensors, arrays, model in cases:
    ours = packed_tensors.PackedTensors()
    if model is not None:
      ours.model = model
    ours.pack(tensors, arrays)
    ref = Example()
    if model is not None:
      ref.features.feature["MD"].bytes_list.value[:] = [model.encode("ascii")]
    for i, (t, a) in enumerate(zip(tensors, arrays)):
      feat = ref.features.feature[chr(i + 1)]
      if t == "string":
        feat.bytes_list.value[:] = list(a)
      else:
        feat.int64_list.value[:] = [int(v) for v in a]
        feat.int64_list.SetInParent()
    want = ref.SerializeToString(deterministic=True)
    assert ours.string == want
    # and the runtime's bytes parse back through our reader, in any map order
    back = packed_tensors.PackedTensors(ref.SerializeToString())
    for got, a, t in zip(back.unpack(tensors), arrays, tensors):
      assert [x for x in got] == [bytes(x) if t == "string" else int(x) for x in a]
    if model is not None:
      assert back.model == model


def test_example_models_pack_unpack_tfci():
  from compression_b200.examples import bls2017, bmshj2018
  c = bmshj2018.Compressed([b"abc", b"defg"], [b"s0", b"s1"], (512, 768), (32, 48), (8, 12))
  one = bmshj2018.Bmshj2018Model.unpack_tfci(bmshj2018.Bmshj2018Model.pack_tfci(c, image=1))
  assert one == bmshj2018.Compressed([b"defg"], [b"s1"], (512, 768), (32, 48), (8, 12))
  c = bls2017.Compressed([b"xyz"], (256, 256), (16, 16))
  assert bls2017.Bls2017Model.unpack_tfci(bls2017.Bls2017Model.pack_tfci(c, model_id="bls")) == c
