"""ctypes binding of libtfc_b200.so (the C ABI declared in include/tfc_b200.h).

There is no CPU fallback: if the shared library is missing this module raises at import of the
first symbol, and every compute entry point needs a CUDA device.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# TFC_B200_SO: another build of the same library (A/B timing of kernel variants)
SO_PATH = os.environ.get("TFC_B200_SO") or os.path.join(_HERE, "libtfc_b200.so")

OK, INVALID_ARGUMENT, CUDA_ERROR, CAPACITY, DECODE, UNIMPLEMENTED = range(6)
STREAM_CAPACITY, STREAM_CORRUPT = 1, 2


class InvalidArgumentError(ValueError):
  """Counterpart of tf.errors.InvalidArgumentError raised by the reference's op kernels."""


class CudaError(RuntimeError):
  pass


_lib = None

c_i64 = ctypes.c_int64
c_i32 = ctypes.c_int32
c_int = ctypes.c_int
c_vp = ctypes.c_void_p
c_sz = ctypes.c_size_t
c_f32 = ctypes.c_float


def lib():
  global _lib
  if _lib is None:
    if not os.path.exists(SO_PATH):
      raise ImportError(
          "libtfc_b200.so not found at %s — build it with `python compression_b200/build.py` "
          "(there is no CPU fallback)" % SO_PATH)
    L = ctypes.CDLL(SO_PATH)
    L.tfc_last_error.restype = ctypes.c_char_p
    L.tfc_version.restype = ctypes.c_char_p
    L.tfc_launch_count.restype = c_i64
    L.tfc_pmf_to_quantized_cdf_workspace.restype = c_sz
    L.tfc_pmf_to_quantized_cdf_workspace.argtypes = [c_i64, c_i32]
    _lib = L
  return _lib


def check(rc):
  if rc == OK:
    return
  msg = lib().tfc_last_error().decode(errors="replace")
  if rc == INVALID_ARGUMENT:
    raise InvalidArgumentError(msg)
  if rc == UNIMPLEMENTED:
    raise NotImplementedError(msg)
  raise CudaError("tfc_b200 status %d: %s" % (rc, msg))


def ptr(t):
  """Device (or host) address of a torch tensor as a ctypes void pointer; None -> NULL."""
  if t is None:
    return c_vp(0)
  assert t.is_contiguous(), "tfc_b200 wants contiguous tensors"
  return c_vp(t.data_ptr())


def stream_ptr(device=None):
  return c_vp(torch.cuda.current_stream(device).cuda_stream)


def launch_count():
  return int(lib().tfc_launch_count())


def require_cuda(t, name):
  if not t.is_cuda:
    raise CudaError("`%s` must live on a CUDA device (no CPU fallback)" % name)


def to_device(x, dtype, device=None):
  """Accepts torch tensors / numpy arrays / lists; returns a contiguous CUDA tensor."""
  if not torch.cuda.is_available():
    raise CudaError("tfc_b200 needs a CUDA device (no CPU fallback)")
  if isinstance(x, torch.Tensor):
    t = x
  else:
    import numpy as np
    t = torch.from_numpy(np.ascontiguousarray(x))
  if device is None:
    device = t.device if t.is_cuda else torch.device("cuda", torch.cuda.current_device())
  return t.to(device=device, dtype=dtype).contiguous()


CFG_SM_BUDGET, CFG_CODER_STREAMS_PER_CTA, CFG_TILES_PER_CTA, CFG_CONV_CLUSTER = 1, 2, 3, 4
CFG_CODER_LANES = 5
CFG_CONV_STAGES, CFG_CODER_TABLE_KB, CFG_CODER_SMEM_KB, CFG_CONV_PAIR = 6, 7, 8, 9


def configure(sm_budget=None, coder_streams_per_cta=None, tiles_per_cta=None, conv_cluster=None,
              coder_lanes=None, conv_stages=None, coder_table_kb=None, coder_smem_kb=None,
              conv_pair=None):
  """Process-wide scheduling knobs of the library (tfc_configure): `sm_budget` = SMs the
  persistent transform kernels may occupy (0: all), `coder_streams_per_cta` = range-coder
  streams packed into one CTA (0: automatic), `tiles_per_cta` = chunked instead of persistent
  transform kernels, `conv_cluster` = thread-block cluster size of the implicit-GEMM kernel
  (0: automatic), `coder_lanes` = one range-coder stream per lane (0: automatic, for thousands
  of streams; 1: never; 2: always).  No effect on results."""
  L = lib()
  if sm_budget is not None:
    check(L.tfc_configure(CFG_SM_BUDGET, int(sm_budget)))
  if coder_streams_per_cta is not None:
    check(L.tfc_configure(CFG_CODER_STREAMS_PER_CTA, int(coder_streams_per_cta)))
  if tiles_per_cta is not None:
    check(L.tfc_configure(CFG_TILES_PER_CTA, int(tiles_per_cta)))
  if conv_cluster is not None:
    check(L.tfc_configure(CFG_CONV_CLUSTER, int(conv_cluster)))
  if coder_lanes is not None:
    check(L.tfc_configure(CFG_CODER_LANES, int(coder_lanes)))
  if conv_stages is not None:      # operand-ring stages of the implicit-GEMM kernel (0: 4)
    check(L.tfc_configure(CFG_CONV_STAGES, int(conv_stages)))
  if conv_pair is not None:        # cta_group::2 implicit GEMM: 0 auto, 1 off, 2 on
    check(L.tfc_configure(CFG_CONV_PAIR, int(conv_pair)))
  if coder_smem_kb is not None:    # dynamic shared memory per warps-per-stream coder CTA (0: 226 KB)
    check(L.tfc_configure(CFG_CODER_SMEM_KB, int(coder_smem_kb)))
  if coder_table_kb is not None:   # shared-memory budget of the coder's staged tables (0: 72 KB)
    check(L.tfc_configure(CFG_CODER_TABLE_KB, int(coder_table_kb)))
