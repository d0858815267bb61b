"""CPU-only checks of the C-ABI boundary: the library loads, exports every symbol the header
declares, and its host-side logic (attribute checks, MergeAxes, error wording) matches the
reference (range_coding_kernels_util.cc:36-92; range_coding_kernels.cc:137-151,180-189;
unbounded_index_range_coding_kernels.cc:145-157)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def L():
  from compression_b200 import build, _lib
  build.build()
  return _lib.lib()


def test_exports_every_declared_symbol(L):
  hdr = open(os.path.join(ROOT, "include", "tfc_b200.h")).read()
  hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
  names = sorted(set(re.findall(r"\b(tfc_[a-z0-9_]+)\s*\(", hdr)))
  assert len(names) >= 10
  missing = [n for n in names if not hasattr(L, n)]
  assert not missing, missing


def _merge(L, dshape, cshape):
  n = len(dshape)
  d = (ctypes.c_int64 * max(n, 1))(*dshape)
  c = (ctypes.c_int64 * (n + 1))(*cshape)
  md = (ctypes.c_int64 * (n + 1))()
  mc = (ctypes.c_int64 * (n + 2))()
  mr = ctypes.c_int(0)
  rc = L.tfc_merge_axes(d, n, c, n + 1, md, mc, ctypes.byref(mr))
  if rc:
    return rc, L.tfc_last_error().decode()
  return list(md[:mr.value]), list(mc[:mr.value + 1])


def test_merge_axes(L):
  # no broadcasting: everything collapses into one axis
  assert _merge(L, [2, 3, 4], [2, 3, 4, 9]) == ([24], [24, 9])
  # full broadcasting
  assert _merge(L, [2, 3, 4], [1, 1, 1, 9]) == ([24], [1, 9])
  # alternating
  assert _merge(L, [2, 3, 4], [2, 1, 4, 9]) == ([2, 3, 4], [2, 1, 4, 9])
  assert _merge(L, [2, 3, 4, 5], [2, 3, 1, 1, 7]) == ([6, 20], [6, 1, 7])
  # size-1 data axes merge with anything
  assert _merge(L, [2, 1, 4], [2, 1, 4, 9]) == ([8], [8, 9])
  rc, msg = _merge(L, [2, 3], [2, 4, 9])
  assert rc == 1 and "Cannot broadcast shape" in msg


def test_attr_errors_need_no_gpu(L):
  i64 = ctypes.c_int64
  z = ctypes.c_void_p(0)
  rc = L.tfc_unbounded_index_range_encode(z, z, i64(0), i64(1), i64(1), z, 1, 3, z, z, 17, 4,
                                          0, z, i64(16), z, z, z)
  assert rc == 1 and "`precision` must be in [1, 16]" in L.tfc_last_error().decode()
  rc = L.tfc_unbounded_index_range_encode(z, z, i64(0), i64(1), i64(1), z, 1, 3, z, z, 16, 0,
                                          0, z, i64(16), z, z, z)
  assert rc == 1 and "`overflow_width` must be in [1, 16]" in L.tfc_last_error().decode()
  rc = L.tfc_unbounded_index_range_decode(z, i64(16), z, z, i64(0), i64(1), i64(1), z, 1, 3, z,
                                          z, 16, 4, 2, z, z, z)
  assert rc == 1 and "`debug_level` must be 0 or 1" in L.tfc_last_error().decode()
  rc = L.tfc_unbounded_index_range_encode(z, z, i64(0), i64(1), i64(1), z, 1, 2, z, z, 16, 4,
                                          0, z, i64(16), z, z, z)
  assert rc == 1 and "'cdf' should be 2-D and cdf.dim_size(1) >= 3" in L.tfc_last_error().decode()
  d = (ctypes.c_int64 * 2)(4, 5)
  c = (ctypes.c_int64 * 2)(4, 9)
  rc = L.tfc_range_encode(z, i64(1), d, 2, z, c, 2, i64(0), 16, 0, z, i64(16), z, z, z)
  assert rc == 1 and "`cdf` should have one more axis than `data`" in L.tfc_last_error().decode()
  c3 = (ctypes.c_int64 * 3)(4, 5, 1)
  rc = L.tfc_range_encode(z, i64(1), d, 2, z, c3, 3, i64(0), 16, 0, z, i64(16), z, z, z)
  assert rc == 1 and "The last dimension of `cdf` should be > 1" in L.tfc_last_error().decode()
  d8 = (ctypes.c_int64 * 8)(2, 2, 2, 2, 2, 2, 2, 2)
  c9 = (ctypes.c_int64 * 9)(2, 1, 2, 1, 2, 1, 2, 1, 5)
  rc = L.tfc_range_decode(z, i64(16), z, i64(1), d8, 8, z, c9, 9, i64(0), 16, 0, z, z, z)
  assert rc == 1 and "Irregular broadcast pattern" in L.tfc_last_error().decode()
  rc = L.tfc_pmf_to_quantized_cdf(z, i64(1), 1, 16, z, z, ctypes.c_size_t(0), z)
  assert rc == 1 and "`pmf` size should be at least 2" in L.tfc_last_error().decode()


def test_stdsort_replay_matches_libstdcxx(tmp_path):
  """The device sort replays libstdc++ std::sort exactly, ties included."""
  import subprocess
  exe = str(tmp_path / "stdsort_check")
  subprocess.check_call(["g++", "-std=c++17", "-O1", "-I",
                         os.path.join(ROOT, "compression_b200", "csrc"),
                         os.path.join(ROOT, "tests", "cpp", "stdsort_check.cc"), "-o", exe])
  out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
  assert out.returncode == 0, out.stdout[-2000:]
  assert "bad=0" in out.stdout


def test_tf_op_shell_compiles_and_links(L, tmp_path):
  """The GPU kernel registrations a maintainer adds to the reference's op library
  (integration/tf_ops) compile against the stub of the TF op API and
  every tfc_* call in them resolves against the built library with the header's signatures."""
  import subprocess
  so = str(tmp_path / "b200_tf_kernels.so")
  cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
  subprocess.check_call([
      "g++", "-std=c++14", "-Wall", "-Werror", "-fPIC", "-shared",
      "-I", os.path.join(ROOT, "integration", "tf_stub"), "-I", os.path.join(ROOT, "include"),
      "-I", os.path.join(cuda, "include"),
      os.path.join(ROOT, "integration", "tf_ops", "b200_range_coding_kernels.cc"),
      "-o", so, "-Wl,--no-undefined", "-L", os.path.join(ROOT, "compression_b200"),
      "-ltfc_b200", "-L", os.path.join(cuda, "lib64"), "-lcudart"])
  src = open(os.path.join(ROOT, "integration", "tf_ops", "b200_range_coding_kernels.cc")).read()
  for op in ("UnboundedIndexRangeEncode", "UnboundedIndexRangeDecode", "RangeEncode", "RangeDecode",
             "PmfToQuantizedCdf"):
    assert 'Name("%s").Device(DEVICE_GPU)' % op in src


def test_no_cpu_fallback():
  import torch
  if torch.cuda.is_available():
    pytest.skip("GPU present")
  import compression_b200 as tfc
  with pytest.raises(tfc.CudaError):
    tfc.pmf_to_quantized_cdf([[0.5, 0.5]], 16)
